"""Per-warp timeline of render_forward at C2 from a -DTSPLAT_TIMELINE build (development only)."""
import ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from cases import case
from transplat_b200 import _lib
from transplat_b200.rasterize import forward_raw
means, cov, sh, op, cams, cfg = case("c2")
cfg.check = "sync"; cfg.want_radii = False
args = [t.cuda() for t in (means, cov, sh, op, cams)]
for _ in range(3):
    forward_raw(cfg, *args)
torch.cuda.synchronize()
n = 8192 * 4
buf = (C.c_ulonglong * n)()
assert _lib.lib().tsplat_debug_timeline(buf, n) == 0
t = np.frombuffer(buf, dtype=np.uint64).reshape(-1, 4).astype(np.int64)
t = t[t[:, 1] > 0]
smid = t[:, 2] >> 32
t[:, 2] &= 0xffffffff
t0 = t[:, 0].min()
end = (t[:, 1] - t0) / 1e3
dur = (t[:, 1] - t[:, 0]) / 1e3
print("warps", len(t), "kernel span us", end.max())
print("end time percentiles us: p10 %.1f p50 %.1f p90 %.1f p99 %.1f max %.1f" % tuple(np.percentile(end, [10, 50, 90, 99, 100])))
print("jobs per warp: mean %.2f min %d max %d" % (t[:, 2].mean(), t[:, 2].min(), t[:, 2].max()))
print("refs per warp: mean %.0f p90 %.0f max %d ; total %d" % (t[:, 3].mean(), np.percentile(t[:, 3], 90), t[:, 3].max(), t[:, 3].sum()))
late = np.argsort(end)[-10:]
for i in late:
    print("  late warp: end %.1f us, jobs %d, refs %d, ns per ref %.0f" % (end[i], t[i, 2], t[i, 3], dur[i] * 1e3 / max(t[i, 3], 1)))
early = np.argsort(end)[:5]
for i in early:
    print("  early warp: end %.1f us, jobs %d, refs %d, ns per ref %.0f" % (end[i], t[i, 2], t[i, 3], dur[i] * 1e3 / max(t[i, 3], 1)))
# correlation between refs processed and end time
print("corr(refs, end) = %.3f" % np.corrcoef(t[:, 3], end)[0, 1])

# per SM: how many warps, their references, when the SM's last warp ends; does CTA c sit on SM c % #SM ?
sms = np.unique(smid)
rows = []
for sm in sms:
    m = smid == sm
    rows.append((end[m].max(), int(sm), int(m.sum()), int(t[m, 3].sum()), int(t[m, 2].sum())))
rows.sort()
print("SMs", len(sms), "smid range", int(sms.min()), int(sms.max()))
print("SM end time us: min %.1f p50 %.1f max %.1f ; refs per SM: min %d mean %.0f max %d ; corr(refs per SM, SM end) = %.3f" % (
    rows[0][0], rows[len(rows) // 2][0], rows[-1][0], min(r[3] for r in rows), np.mean([r[3] for r in rows]),
    max(r[3] for r in rows), np.corrcoef([r[3] for r in rows], [r[0] for r in rows])[0, 1]))
for r in rows[-5:]:
    print("  late SM %d: end %.1f us, warps %d, refs %d, jobs %d" % (r[1], r[0], r[2], r[3], r[4]))
for r in rows[:3]:
    print("  early SM %d: end %.1f us, warps %d, refs %d, jobs %d" % (r[1], r[0], r[2], r[3], r[4]))
cta = np.arange(len(smid)) // 8
first = {}
for c, sm in zip(cta, smid):
    first.setdefault(int(c), int(sm))
seq = [first[c] for c in sorted(first)][:300]
print("SM of CTA 0..15:", seq[:16], " CTA 148..155:", seq[148:156])
