"""Stage-by-stage CUDA-vs-oracle report (run on the B200 box): prints one JSON line per case and
never stops at the first mismatch.  `python scripts/gpu_check.py [case ...]`"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

from cases import case  # noqa: E402
from parity import run_both  # noqa: E402

names = sys.argv[1:] or ["tiny_l1", "deg2_l1", "small_l2_depth", "ragged_l2", "mid_l2"]
for name in names:
    t0 = time.time()
    try:
        means, cov, sh, op, cams, cfg = case(name)
        fwd, bwd = run_both(means, cov, sh, op, cams, cfg)
        print(json.dumps({"case": name, "G": means.shape[1], "fwd": fwd, "bwd": bwd, "sec": round(time.time() - t0, 2)}), flush=True)
    except Exception as e:  # keep going: one report per case
        import traceback
        traceback.print_exc()
        print(json.dumps({"case": name, "error": repr(e)}), flush=True)
        if "CUDA" in repr(e) or "cuda" in repr(e):
            break
