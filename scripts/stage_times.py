"""Per-stage device times of the C2 step (1 scene x 4 views, colour+depth fwd+bwd) for the library selected by
TSPLAT_LIB_PATH (kernel experiments).  Prints one line.  Run on the B200 box."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from transplat_b200 import _lib
from transplat_b200.decoder import DecoderSplattingCUDA, Gaussians
from transplat_b200.synthetic import make_scene, upstream_grads

G_CTX, HW, V = int(os.environ.get("N_CTX", 2)), int(os.environ.get("HW", 256)), int(os.environ.get("VIEWS", 4))
dev = torch.device("cuda")
scs = [make_scene(n_ctx=G_CTX, h=HW, w=HW, n_views=V, seed=1236 + i) for i in range(2)]
target, dLd = (t[None].to(dev) for t in upstream_grads(V, HW, HW, seed=99))
leaf = lambda t: t[None].to(dev).requires_grad_(True)
S = [dict(means=leaf(sc.means), cov=leaf(sc.covariances), sh=leaf(sc.harmonics), op=leaf(sc.opacities),
          ext=sc.extrinsics[None].to(dev), intr=sc.intrinsics[None].to(dev), near=sc.near[None].to(dev),
          far=sc.far[None].to(dev)) for sc in scs]
dec = DecoderSplattingCUDA().to(dev)
numel = float(V * 3 * HW * HW)
one = torch.ones((), device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def step(s):
    g = Gaussians(s["means"], s["cov"], s["sh"], s["op"])
    if os.environ.get("FUSED", "1") == "1":
        out, loss = dec.forward_mse(g, s["ext"], s["intr"], s["near"], s["far"], (HW, HW), target, depth_mode="depth")
        return torch.autograd.grad([loss, out.depth], [s["means"], s["cov"], s["sh"], s["op"]], [one, dLd])
    out = dec.forward(g, s["ext"], s["intr"], s["near"], s["far"], (HW, HW), depth_mode="depth")
    dLc = (out.color - target) * (2.0 / numel)
    return torch.autograd.grad([out.color, out.depth], [s["means"], s["cov"], s["sh"], s["op"]], [dLc, dLd])


for i in range(10):
    step(S[i % 2])
torch.cuda.synchronize()
PROF = os.environ.get('PROFILE', '1') == '1'
_lib.profile_enable(PROF)
n = int(os.environ.get("STEPS", 40))
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
for i in range(n):
    flush.zero_()
    ev[i][0].record()
    step(S[i % 2])
    ev[i][1].record()
torch.cuda.synchronize()
st = _lib.profile_read() if PROF else {}
ms = sorted(a.elapsed_time(b) for a, b in ev)
print(os.environ.get("TAG", "?"), "step_ms median %.4f mean %.4f |" % (ms[n // 2], sum(ms) / n),
      " ".join(f"{k}={v:.4f}" for k, v in st.items() if k != "calls" and v > 0))
