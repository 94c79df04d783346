"""Kernel timeline of one C2 step (CUPTI through torch.profiler): start offsets, durations and the
idle gaps between consecutive kernels.  Run on the B200 box."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

from transplat_b200.decoder import DecoderSplattingCUDA, Gaussians  # noqa: E402
from transplat_b200.synthetic import make_scene, upstream_grads  # noqa: E402

dev = torch.device("cuda")
sc = make_scene(n_ctx=2, h=256, w=256, n_views=4, seed=1236)
target, dLd = (t[None].to(dev) for t in upstream_grads(4, 256, 256, seed=99))
leaf = lambda t: t[None].to(dev).requires_grad_(True)
s = dict(means=leaf(sc.means), cov=leaf(sc.covariances), sh=leaf(sc.harmonics), op=leaf(sc.opacities),
         ext=sc.extrinsics[None].to(dev), intr=sc.intrinsics[None].to(dev), near=sc.near[None].to(dev), far=sc.far[None].to(dev))
dec = DecoderSplattingCUDA().to(dev)
numel = float(4 * 3 * 256 * 256)


def step():
    g = Gaussians(s["means"], s["cov"], s["sh"], s["op"])
    out = dec.forward(g, s["ext"], s["intr"], s["near"], s["far"], (256, 256), depth_mode="depth")
    dLc = (out.color - target) * (2.0 / numel)
    return torch.autograd.grad([out.color, out.depth], [s["means"], s["cov"], s["sh"], s["op"]], [dLc, dLd])


for _ in range(20):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(6):
        step()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
# take the 4th step: find kernel names pattern
names = [e.name for e in evs]
starts = [i for i, n in enumerate(names) if "pack_cameras" in n]
i0, i1 = starts[3], starts[4]
t0 = evs[i0].time_range.start
prev_end = None
tot_k = tot_gap = 0.0
for e in evs[i0:i1]:
    st, en = e.time_range.start - t0, e.time_range.end - t0
    gap = (st - prev_end) if prev_end is not None else 0.0
    print(f"{st:9.1f} us  dur {en - st:8.1f}  gap {gap:6.1f}  {e.name[:70]}")
    tot_k += en - st
    tot_gap += max(gap, 0.0)
    prev_end = max(en, prev_end or 0)
print(f"step span {evs[i1].time_range.start - t0:.1f} us, kernels {tot_k:.1f} us, gaps {tot_gap:.1f} us")
