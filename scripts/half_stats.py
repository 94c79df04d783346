"""CPU replay of the C2 block lists (oracle preprocess/binning, test infrastructure only; development helper, no GPU):
how many visits a render warp would make if the two 4x4 halves of its 8x4 block walked the batch's entries in
lock-step (left lanes take the next entry that touches the left half, right lanes the next one that touches the
right half) instead of every lane visiting every entry of the block's list."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import oracle
from transplat_b200.synthetic import make_scene
from transplat_b200.cameras import pack_cameras

V = 4
sc = make_scene(n_ctx=2, h=256, w=256, n_views=V, seed=1236)
cams = pack_cameras(sc.extrinsics, sc.intrinsics, sc.near, sc.far, torch.zeros(V, 3)).numpy()
S = oracle.Scene(sc.means.numpy(), sc.covariances.numpy(), sc.harmonics.numpy(), sc.opacities.numpy(), 4,
                 oracle.SH_G3M, oracle.COV_FULL9)
view = int(sys.argv[1]) if len(sys.argv) > 1 else 1
st = oracle.forward(S, cams[view], 256, 256, depth_mode="depth")
xy, co = st.xy, st.conic_opacity
tot = dict(list=0, lock=0, lock_all=0, both=0, quarter=0)


def cont(x, y, A, B, Cc, tau, ok, X0, Y0, bw, bh):
    a_lo = X0 - x; a_hi = a_lo + bw - 1; b_lo = Y0 - y; b_hi = b_lo + bh - 1
    ax = np.minimum(np.maximum(0, a_lo), a_hi); by = np.minimum(np.maximum(0, b_lo), b_hi)
    dy1 = np.minimum(np.maximum(-B * ax / Cc, b_lo), b_hi); dx2 = np.minimum(np.maximum(-B * by / A, a_lo), a_hi)
    q1 = 0.5 * (A * ax * ax + Cc * dy1 * dy1) + B * ax * dy1; q2 = 0.5 * (A * dx2 * dx2 + Cc * by * by) + B * dx2 * by
    return (np.minimum(q1, q2) <= (tau + 2e-4) * 1.003) & ok


for tile in range(256):
    lo, hi = st.ranges[tile]
    if hi <= lo:
        continue
    ids = st.vals[lo:hi]
    ty, tx = divmod(tile, 16)
    ncon = st.n_contrib[ty * 16:ty * 16 + 16, tx * 16:tx * 16 + 16].astype(np.int64)
    x, y = xy[ids, 0], xy[ids, 1]
    A, B, Cc, o = (co[ids, k] for k in range(4))
    det = A * Cc - B * B
    tau = np.log(np.maximum(255 * o, 1e-30))
    ok = (o >= 1 / 255) & (det > 0)
    pos = np.arange(len(ids))
    for blk in range(8):
        X0, Y0 = tx * 16 + (blk & 1) * 8, ty * 16 + (blk >> 1) * 4
        last = ncon[(blk >> 1) * 4:(blk >> 1) * 4 + 4, (blk & 1) * 8:(blk & 1) * 8 + 8].max()
        inlist = cont(x, y, A, B, Cc, tau, ok, X0, Y0, 8, 4) & (pos < last)
        L = cont(x, y, A, B, Cc, tau, ok, X0, Y0, 4, 4)[inlist]
        R = cont(x, y, A, B, Cc, tau, ok, X0 + 4, Y0, 4, 4)[inlist]
        q = [cont(x, y, A, B, Cc, tau, ok, X0 + 4 * (k & 1), Y0 + 2 * (k >> 1), 4, 2)[inlist] for k in range(4)]
        n = int(inlist.sum())
        tot["list"] += n
        tot["both"] += int((L & R).sum())
        tot["lock_all"] += max(int(L.sum()), int(R.sum()))
        for b0 in range(0, n, 32):
            tot["lock"] += max(int(L[b0:b0 + 32].sum()), int(R[b0:b0 + 32].sum()))
            tot["quarter"] += max(int(qq[b0:b0 + 32].sum()) for qq in q)
print(tot)
print("visits, halves in lock-step per batch of 32: %.3f of the list; over whole lists: %.3f; entries touching both halves: %.3f; "
      "quarters (4x2) per batch: %.3f" % (tot["lock"] / tot["list"], tot["lock_all"] / tot["list"], tot["both"] / tot["list"],
                                          tot["quarter"] / tot["list"]))
