"""Timeline of the host-buffer pipeline (bench.py's e2e leg): per step, when its H2D, kernels and D2H start and end."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from transplat_b200.decoder import DecoderSplattingCUDA, Gaussians
from transplat_b200.streaming import HostPipeline
from transplat_b200.synthetic import make_scene, upstream_grads

HW, V = 256, 4
dev = torch.device("cuda")
scenes = [make_scene(n_ctx=2, h=HW, w=HW, n_views=V, seed=1236 + i) for i in range(4)]
target_h, dLd_h = upstream_grads(V, HW, HW, seed=99)
pin = lambda t: t.contiguous().pin_memory()
host = [dict(means=pin(s.means[None]), cov=pin(s.covariances[None]), sh=pin(s.harmonics[None]), op=pin(s.opacities[None]),
             ext=pin(s.extrinsics[None]), intr=pin(s.intrinsics[None]), near=pin(s.near[None]), far=pin(s.far[None]),
             target=pin(target_h[None]), dLd=pin(dLd_h[None])) for s in scenes]
G = scenes[0].means.shape[0]
outs = [dict(color=torch.empty((1, V, 3, HW, HW)).pin_memory(), depth=torch.empty((1, V, HW, HW)).pin_memory(),
             g_means=torch.empty((1, G, 3)).pin_memory(), g_cov=torch.empty((1, G, 3, 3)).pin_memory(),
             g_sh=torch.empty((1, G, 3, 25)).pin_memory(), g_op=torch.empty((1, G)).pin_memory()) for _ in range(4)]
dec = DecoderSplattingCUDA().to(dev)
numel = float(V * 3 * HW * HW)


def step_fn(d):
    leaves = [d[k].detach().requires_grad_(True) for k in ("means", "cov", "sh", "op")]
    out = dec.forward(Gaussians(*leaves), d["ext"], d["intr"], d["near"], d["far"], (HW, HW), depth_mode="depth")
    dLc = (out.color - d["target"]) * (2.0 / numel)
    g = torch.autograd.grad([out.color, out.depth], leaves, [dLc, d["dLd"]])
    return dict(color=out.color, depth=out.depth, g_means=g[0], g_cov=g[1], g_sh=g[2], g_op=g[3])


slots = int(sys.argv[1]) if len(sys.argv) > 1 else 3
pipe = HostPipeline(dev, slots=slots)
for i in range(8):
    pipe.submit(step_fn, host[i % 4], outs[i % 4])
pipe.drain()
pipe = HostPipeline(dev, slots=slots, trace=True)
pipe._dev_in = [None] * slots
torch.cuda.synchronize()
t0 = time.perf_counter()
hs = []
for i in range(16):
    a = time.perf_counter()
    pipe.submit(step_fn, host[i % 4], outs[i % 4])
    hs.append((a - t0, time.perf_counter() - t0))
pipe.drain()
torch.cuda.synchronize()
base = pipe.trace[0][0]
print("step  host_submit[start,end]   h2d[start,end]   compute[start,end]   d2h[start,end]  (ms)")
for i, tr in enumerate(pipe.trace):
    t = [base.elapsed_time(e) for e in tr]
    print(f"{i:3d}  [{hs[i][0]*1e3:6.2f},{hs[i][1]*1e3:6.2f}]  [{t[0]:6.2f},{t[1]:6.2f}]  [{t[2]:6.2f},{t[3]:6.2f}]  [{t[4]:6.2f},{t[5]:6.2f}]")
print("period ms/step over the last 8:", (base.elapsed_time(pipe.trace[-1][5]) - base.elapsed_time(pipe.trace[-9][5])) / 8)
