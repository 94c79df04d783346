"""Where the HOST time of one C2 step goes (run on the B200 box): cProfile over N steps of
DecoderSplattingCUDA.forward + autograd.grad, plus wall/device ms per step."""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

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


for _ in range(10):
    step()
torch.cuda.synchronize()
N = int(sys.argv[1]) if len(sys.argv) > 1 else 200
t0 = time.perf_counter()
for _ in range(N):
    step()
t_issue = time.perf_counter() - t0
torch.cuda.synchronize()
t_all = time.perf_counter() - t0
print(f"steps={N} host-issue {1e3 * t_issue / N:.3f} ms/step, wall incl. drain {1e3 * t_all / N:.3f} ms/step")
# host-only cost: sync after every step so the queue never back-pressures
t_host = 0.0
for _ in range(N):
    t1 = time.perf_counter()
    step()
    t_host += time.perf_counter() - t1
    torch.cuda.synchronize()
print(f"host time per step with an empty queue: {1e3 * t_host / N:.3f} ms")
pr = cProfile.Profile()
pr.enable()
for _ in range(N):
    step()
    torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
