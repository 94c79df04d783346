"""Device times of the other BASELINE configurations (C4a DTU, C4b video, C5 stress) next to C2, on ONE B200.
bench.py quotes the metric on C2 only; these are the same kernels at the other shapes, with the same roofline
arithmetic (bytes_per_view of bench.py, forward-only form for the evaluation configs).  One JSON line per config.
    python scripts/config_times.py [c2 c4a c4b c5]
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from transplat_b200 import _lib  # noqa: E402
from transplat_b200.decoder import DecoderSplattingCUDA, Gaussians  # noqa: E402
from transplat_b200.synthetic import make_scene, upstream_grads  # noqa: E402

CONFIGS = {
    # name: (make_scene kwargs, backward?, steps, description)
    "c2": (dict(n_ctx=2, h=256, w=256, n_views=4, seed=1236), True, 40, "RE10K train step: 131072 Gaussians, 4 views 256x256, fwd+bwd"),
    "c4a": (dict(n_ctx=3, h=256, w=256, n_views=1, seed=1238, near=2.125, far=4.525, d0=2.2, d1=4.4), False, 40,
            "DTU evaluation: 196608 Gaussians (3 context views), 1 view 256x256, forward"),
    "c4b": (dict(n_ctx=2, h=256, w=256, n_views=67, seed=1239), False, 20,
            "RE10K video: 131072 Gaussians, 67 views 256x256 (mean of the video index), forward"),
    "c5": (dict(n_ctx=2, h=1024, w=1024, n_views=16, seed=1240), True, 3,
           "stress: 2097152 Gaussians, 1024x1024, fwd+bwd; 16 of the 64 views = one rank's share when the views are sharded over 4 GPUs"),
}


def bytes_per_view(G, D, P, T, backward):
    """fwd+bwd 1232 G + 260 D + 48 P + 8 T; forward-only 420 G + 216 D + 24 P + 8 T."""
    return (1232.0 * G + 260.0 * D + 48.0 * P + 8.0 * T) if backward else (420.0 * G + 216.0 * D + 24.0 * P + 8.0 * T)


def run(name):
    kw, backward, steps, desc = CONFIGS[name]
    dev = torch.device("cuda")
    sc = make_scene(**kw)
    V, (H, W) = sc.extrinsics.shape[0], sc.image_shape
    G = sc.means.shape[0]
    leaf = lambda t: t[None].to(dev).requires_grad_(backward)
    s = dict(means=leaf(sc.means), cov=leaf(sc.covariances), sh=leaf(sc.harmonics), op=leaf(sc.opacities))
    cam = tuple(t[None].to(dev) for t in (sc.extrinsics, sc.intrinsics, sc.near, sc.far))
    dec = DecoderSplattingCUDA().to(dev)
    target = dLd = None
    if backward:
        target, dLd = (t[None].to(dev) for t in upstream_grads(V, H, W, seed=99))
    one = torch.ones((), device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def step():
        g = Gaussians(s["means"], s["cov"], s["sh"], s["op"])
        if backward:
            out, loss = dec.forward_mse(g, *cam, (H, W), target, depth_mode="depth")
            return torch.autograd.grad([loss, out.depth], [s["means"], s["cov"], s["sh"], s["op"]], [one, dLd])
        with torch.no_grad():
            return dec.forward(g, *cam, (H, W), depth_mode="depth")

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    _lib.profile_enable(False)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for i in range(steps):
        flush.zero_()
        ev[i][0].record()
        step()
        ev[i][1].record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in ev)
    med = ms[steps // 2]
    # second pass with the library's per-stage events, and the pair count of the last call
    _lib.profile_enable(True)
    for _ in range(min(steps, 5)):
        step()
    torch.cuda.synchronize()
    st = _lib.profile_read()
    _lib.profile_enable(False)
    calls = max(int(st.get("calls", 1)), 1)
    from transplat_b200 import rasterize as rz
    rz._capacity.sweep(wait=True)
    D = rz._capacity.last_num_rendered
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = peaks.get("hbm_gbs", 6650.0)
    T = ((H + 15) // 16) * ((W + 15) // 16)
    line = {"config": name, "what": desc, "gaussians": G, "views": V, "hw": [H, W], "backward": backward,
            "ms_per_step": round(med, 4), "ms_per_view": round(med / V, 4), "views_per_s": round(V / med * 1e3, 1),
            "gaussians_per_s": round(G * V / med * 1e3, 0), "steps": steps,
            "stage_ms": {k: round(v, 4) for k, v in st.items() if k != "calls" and v > 0}}
    if D is not None:
        b = bytes_per_view(G, D / V, H * W, T, backward) * V
        line.update(pairs_per_view=round(D / V), algorithmic_bytes=int(b), achieved_gbs=round(b / (med * 1e-3) / 1e9, 1),
                    hbm_frac=round(b / (med * 1e-3) / 1e9 / peak, 4), hbm_peak_gbs=peak)
    print(json.dumps(line), flush=True)
    del s, cam, target, dLd, flush
    torch.cuda.empty_cache()


if __name__ == "__main__":
    for n in (sys.argv[1:] or list(CONFIGS)):
        run(n)
