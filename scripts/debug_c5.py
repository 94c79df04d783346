"""Localise a colour mismatch at the C5 shape: prints the full parity report and the worst pixels."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from cases import case
from parity import oracle_batch
from transplat_b200.rasterize import forward_raw, inspect_workspace

means, cov, sh, op, cams, cfg = case("c5_stress")
cfg.check = "sync"; cfg.want_radii = False
cfg.sort = sys.argv[1] if len(sys.argv) > 1 else "bucket"
dev = "cuda"
color, depth, _, state = forward_raw(cfg, *(t.to(dev) for t in (means, cov, sh, op, cams)))
ins = inspect_workspace(state)
states, _ = oracle_batch(means, cov, sh, op, cams, cfg)
c = color.cpu().numpy()
for n, st in enumerate(states):
    d = np.abs(c[n] - st.color).max(0)
    nc = ins["n_contrib"][n].cpu().numpy().astype(np.uint32)
    ft = ins["final_T"][n].cpu().numpy()
    bad = np.argwhere(d > 1e-5)
    print("view", n, "max", d.max(), "bad pixels", len(bad), "n_contrib mismatches", int((nc != st.n_contrib).sum()),
          "final_T max diff", np.abs(ft - st.final_T).max())
    tiles = {}
    for y, x in bad[:2000]:
        tiles[(y // 16, x // 16)] = tiles.get((y // 16, x // 16), 0) + 1
    print(" tiles with bad pixels:", len(tiles), sorted(tiles.items(), key=lambda kv: -kv[1])[:10])
    r = ins["ranges"][n].cpu().numpy()
    for (ty, tx), cnt in sorted(tiles.items(), key=lambda kv: -kv[1])[:5]:
        t = ty * 64 + tx
        print("  tile", (ty, tx), "list length", int(r[t, 1] - r[t, 0]), "bad", cnt)
    for y, x in bad[:8]:
        print("  pixel", (y, x), "gpu", c[n][:, y, x], "oracle", st.color[:, y, x], "n_contrib gpu/oracle", nc[y, x], st.n_contrib[y, x],
              "T", ft[y, x], st.final_T[y, x])
