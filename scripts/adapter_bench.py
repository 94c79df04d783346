"""Device time of the GaussianAdapter kernels at the C2 size (2 context views x 65536 pixels, SH degree 4) against the
reference's torch op chain (oracle/adapter_ref.adapter_restated moved to the GPU: the same ~40 ops the reference
launches) - CUDA events, L2 flushed between iterations.  Run on the B200 box; prints one JSON line."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from transplat_b200.adapter import GaussianAdapter, GaussianAdapterCfg
from oracle.adapter_ref import adapter_restated

dev = torch.device("cuda")
N, n, deg, hw = 2, 65536, 4, (256, 256)
g = torch.Generator().manual_seed(1)
ext = torch.eye(4).repeat(N, 1, 1); ext[1, 0, 3] = 0.3
intr = torch.tensor([[0.9, 0, 0.5], [0, 0.9, 0.5], [0, 0, 1.0]]).repeat(N, 1, 1)
wig = [torch.linalg.qr(torch.randn(N, 2 * l + 1, 2 * l + 1, generator=g))[0].contiguous().to(dev) for l in range(deg + 1)]
coords = torch.rand(1, N, n, 1, 1, 2, generator=g).to(dev).requires_grad_(True)
depths = (1 + 9 * torch.rand(1, N, n, 1, 1, generator=g)).to(dev).requires_grad_(True)
opac = torch.rand(1, N, n, 1, 1, generator=g).to(dev)
raw = torch.randn(1, N, n, 1, 1, 82, generator=g).to(dev).requires_grad_(True)
ad = GaussianAdapter(GaussianAdapterCfg(0.5, 15.0, deg)).to(dev)
ext_d, intr_d = ext.to(dev), intr.to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
gm, gc, gh = (torch.randn(1, N, n, 1, 1, *s, device=dev) for s in ((3,), (3, 3), (3, 25)))


def ours():
    G = ad.forward(ext_d.reshape(1, N, 1, 1, 1, 4, 4), intr_d.reshape(1, N, 1, 1, 1, 3, 3), coords, depths, opac, raw, hw,
                   wigner_matrices=[w[None] for w in wig])
    return (G.means, G.covariances, G.harmonics), (gm, gc, gh)


torch.set_default_device("cpu")
def ref():
    import oracle.adapter_ref as R
    m, c, h, *_ = adapter_restated_gpu()
    return (m, c, h), (gm.reshape(N, n, 3), gc.reshape(N, n, 3, 3), gh.reshape(N, n, 3, 25))


def adapter_restated_gpu():
    import oracle.adapter_ref as R
    old = R.sh_mask
    R.sh_mask = lambda d: old(d).to(dev)
    pix = torch.tensor
    try:
        with torch.device(dev):
            return adapter_restated(ext_d, intr_d, coords.reshape(N, n, 2), depths.reshape(N, n), opac.reshape(N, n),
                                    raw.reshape(N, n, 82), wig, hw, 0.5, 15.0, deg)
    finally:
        R.sh_mask = old


def timed(fn, iters=30):
    fw, bw = [], []
    for i in range(iters + 5):
        flush.zero_()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        outs, gs = fn()
        e[1].record()
        torch.autograd.grad(outs, [coords, depths, raw], gs)
        e[2].record()
        torch.cuda.synchronize()
        if i >= 5:
            fw.append(e[0].elapsed_time(e[1])); bw.append(e[1].elapsed_time(e[2]))
    fw.sort(); bw.sort()
    return fw[len(fw) // 2], bw[len(bw) // 2]


G_ = N * n
of, ob = timed(ours)
rf, rb = timed(ref)
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
bytes_f, bytes_b = G_ * (340 + 376), G_ * (1012 + 340)
print(json.dumps({"workload": "GaussianAdapter, 2 context views x 65536 px, SH degree 4 (C2)", "gaussians": G_,
                  "ours_ms": {"forward": round(of, 4), "backward": round(ob, 4)},
                  "torch_op_chain_ms": {"forward": round(rf, 4), "backward": round(rb, 4)},
                  "roofline": {"bound": "hbm", "peak": peak, "unit": "GB/s",
                               "forward": {"algorithmic_bytes": bytes_f, "achieved": round(bytes_f / of / 1e6, 1), "frac": round(bytes_f / of / 1e6 / peak, 4)},
                               "backward": {"algorithmic_bytes": bytes_b, "achieved": round(bytes_b / ob / 1e6, 1), "frac": round(bytes_b / ob / 1e6 / peak, 4)}},
                  "how": "CUDA events around the Python call (includes launch + allocation), 256 MiB L2 flush between iterations, median of 30"}))
