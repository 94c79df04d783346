"""Distribution of the render jobs' list lengths at C2 and the balance of the static deal (run on the B200 box)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from cases import case
from transplat_b200.rasterize import forward_raw, inspect_workspace
means, cov, sh, op, cams, cfg = case(sys.argv[1] if len(sys.argv) > 1 else "c2")
cfg.check = "sync"; cfg.want_radii = False
color, depth, _, state = forward_raw(cfg, *(t.cuda() for t in (means, cov, sh, op, cams)))
ins = inspect_workspace(state)
jobs = ins["block_jobs"].cpu().numpy().astype(np.int64)
cnt, last = jobs[:, 1], jobs[:, 3]
print("pairs D", ins["num_rendered"], "block entries E", ins["num_block_entries"], "E/D %.3f" % (ins["num_block_entries"] / ins["num_rendered"]))
print("jobs", len(cnt), "mean", cnt.mean(), "median", np.median(cnt), "p90", np.percentile(cnt, 90), "p99", np.percentile(cnt, 99), "max", cnt.max())
print("bwd (cut at deepest contributor): sum", last.clip(0, None).sum(), "mean", last.mean(), "max", last.max())
for name, w in (("fwd count", cnt), ("bwd last", np.minimum(cnt, last))):
    for W in (4736, 5920):
        per = np.zeros(W)
        for i in range(0, len(w), W):
            seg = w[i:i + W]
            per[:len(seg)] += seg
        print(f"  {name}: static deal over {W} warps: max/mean per warp = {per.max() / per.mean():.2f}, max {per.max():.0f}, mean {per.mean():.0f}")
    # per-SM sums (32 or 40 warps per SM: consecutive warps of 4-5 CTAs; CTA c on SM c % 148 roughly)
    srt = np.sort(w)[::-1]
    print(f"  {name}: heaviest jobs", srt[:10])
np.save(os.path.join(ROOT, "gpurun_out", "block_jobs_c2.npy"), jobs)
