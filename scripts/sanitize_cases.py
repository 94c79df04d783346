"""The small parity cases through libtsplat.so, forward + backward, both binning paths and the fused-MSE entry points -
the workload for compute-sanitizer (one tool per gpurun call):
    compute-sanitizer --tool memcheck|racecheck|synccheck python scripts/sanitize_cases.py
"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from cases import case
from transplat_b200.rasterize import backward_raw, forward_raw

names = sys.argv[1:] or ["tiny_l1", "ragged_l2", "dense_big", "coplanar"]
for name in names:
    for sort in ("bucket", "onesweep"):
        for fused in (False, True):
            means, cov, sh, op, cams, cfg = case(name)
            cfg.sort, cfg.check = sort, "sync"
            args = [t.cuda() for t in (means, cov, sh, op, cams)]
            N = cams.shape[0]
            g = torch.Generator().manual_seed(3)
            target = torch.rand(N, 3, cfg.H, cfg.W, generator=g).cuda() if fused else None
            color, depth, radii, state = forward_raw(cfg, *args, mse_target=target, mse_psnr=fused)
            dLc = torch.randn(N, 3, cfg.H, cfg.W, generator=g).cuda()
            dLd = torch.randn(N, cfg.H, cfg.W, generator=g).cuda() if cfg.depth_mode else None
            grads = backward_raw(cfg, state, dLc, dLd, grad_loss=torch.ones(1, device="cuda") if fused else None)
            torch.cuda.synchronize()
            assert all(torch.isfinite(x).all() for x in grads[:4]) and torch.isfinite(color).all()
            print(name, sort, "fused" if fused else "plain", "ok", float(color.mean()), flush=True)
print("done")
