"""CPU replay of the C2 tile lists (oracle preprocess/binning, test infrastructure only): how many (warp, entry)
visits the render-backward kernel makes under different culling granularities.  No GPU."""
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
print("D", st.num_rendered)
xy, co = st.xy, st.conic_opacity
tot = dict(pairs=0, cur=0, exact=0, valid_lanes=0, half_bbox=0, half_exact=0, quarter_bbox=0, quarter_exact=0,
           rounds=0, cur_fwdorder=0)
for tile in range(256):
    lo, hi = st.ranges[tile]
    if hi <= lo: continue
    ids = st.vals[lo:hi]
    ty, tx = divmod(tile, 16)
    py, px = np.mgrid[ty*16:ty*16+16, tx*16:tx*16+16]
    pxf, pyf = px.reshape(-1).astype(np.float32), py.reshape(-1).astype(np.float32)
    ncon = st.n_contrib[ty*16:ty*16+16, tx*16:tx*16+16].reshape(-1).astype(np.int64)
    E = len(ids)
    x, y = xy[ids, 0][:, None], xy[ids, 1][:, None]
    A, B, Cc, o = (co[ids, k][:, None] for k in range(4))
    dx, dy = x - pxf[None], y - pyf[None]
    power = -0.5 * (A*dx*dx + Cc*dy*dy) - B*dx*dy
    alpha = np.minimum(0.99, o*np.exp(power))
    pos = np.arange(E)[:, None]
    valid = (pos < ncon[None]) & (power <= 0) & (alpha >= 1/255)          # [E,256]
    # extents
    det = (A*Cc - B*B)[:, 0]
    tau = np.log(np.maximum(255*o[:, 0], 1e-30))
    ok = (o[:, 0] >= 1/255) & (det > 0)
    hx = np.where(ok, np.sqrt(np.maximum(2*tau*Cc[:, 0]/np.where(det > 0, det, 1), 0))*1.0002+0.01, -np.inf)
    hy = np.where(ok, np.sqrt(np.maximum(2*tau*A[:, 0]/np.where(det > 0, det, 1), 0))*1.0002+0.01, -np.inf)
    pxi, pyi = (px.reshape(-1) - tx*16), (py.reshape(-1) - ty*16)
    def blocks(bw, bh):
        bid = (pyi // bh) * (16 // bw) + (pxi // bw)
        nb = (16//bw)*(16//bh)
        X0 = tx*16 + (np.arange(nb) % (16//bw))*bw; Y0 = ty*16 + (np.arange(nb) // (16//bw))*bh
        last = np.array([ncon[bid == k].max() for k in range(nb)])
        bbox = ((x + hx[:, None] >= X0[None]) & (x - hx[:, None] <= X0[None]+bw-1) &
                (y + hy[:, None] >= Y0[None]) & (y - hy[:, None] <= Y0[None]+bh-1) & (pos < last[None]))
        ex = np.stack([valid[:, bid == k].any(1) for k in range(nb)], 1)
        if bw == 8:
            a_lo = X0[None] - x; a_hi = a_lo + bw - 1; b_lo = Y0[None] - y; b_hi = b_lo + bh - 1
            ax = np.minimum(np.maximum(0, a_lo), a_hi); by = np.minimum(np.maximum(0, b_lo), b_hi)
            dy1 = np.minimum(np.maximum(-B*ax/Cc, b_lo), b_hi); dx2 = np.minimum(np.maximum(-B*by/A, a_lo), a_hi)
            q1 = 0.5*(A*ax*ax + Cc*dy1*dy1) + B*ax*dy1; q2 = 0.5*(A*dx2*dx2 + Cc*by*by) + B*dx2*by
            cont = (np.minimum(q1, q2) <= (tau[:, None] + 2e-4)*1.003) & ok[:, None] & (pos < last[None])
            tot["cont"] = tot.get("cont", 0) + cont.sum()
            assert not (ex & ~cont).any()
            xl = np.ceil(np.maximum(x - hx[:, None], X0[None])); xh = np.floor(np.minimum(x + hx[:, None], X0[None] + bw - 1))
            yl = np.ceil(np.maximum(y - hy[:, None], Y0[None])); yh = np.floor(np.minimum(y + hy[:, None], Y0[None] + bh - 1))
            disc = (xl <= xh) & (yl <= yh) & (pos < last[None])
            assert not (ex & ~disc).any()
            tot["disc"] = tot.get("disc", 0) + disc.sum()
            tot["disc_cont"] = tot.get("disc_cont", 0) + (disc & cont).sum()
            # q at the clamped-rounded centre pixel and min over discrete bbox when it holds <= 2x2 pixels
            small = ((xh - xl) <= 1) & ((yh - yl) <= 1)
            qs = np.full(disc.shape, np.inf)
            for ix in (xl, xh):
                for iy in (yl, yh):
                    ddx, ddy = ix - x, iy - y
                    qs = np.minimum(qs, 0.5*(A*ddx*ddx + Cc*ddy*ddy) + B*ddx*ddy)
            sm = disc & cont & (~small | (qs <= (tau[:, None] + 2e-4)*1.003))
            assert not (ex & ~sm).any()
            tot["disc_cont_small"] = tot.get("disc_cont_small", 0) + sm.sum()
        return bbox, ex
    b84, e84 = blocks(8, 4)
    tot["pairs"] += E
    tot["cur"] += b84.sum(); tot["exact"] += e84.sum(); tot["valid_lanes"] += valid.sum()
    b44, e44 = blocks(4, 4)   # 16 blocks: warp w (8x4 at bx=(w&1)*8, by=(w>>1)*4) = 4x4 blocks (by, 2*(w&1)), (by, 2*(w&1)+1)
    b42, e42 = blocks(4, 2)
    R = (E + 31)//32
    tot["rounds"] += R*8
    padE = R*32
    def steps(mask, group):   # mask [E, nblk]; group: list of lists of block ids per warp
        m = np.zeros((padE, mask.shape[1]), bool); m[:E] = mask
        cnt = m.reshape(R, 32, -1).sum(1)      # [R, nblk]
        return sum(cnt[:, g].max(1).sum() for g in group)
    g44 = [[(w >> 1)*4 + 2*(w & 1), (w >> 1)*4 + 2*(w & 1) + 1] for w in range(8)]
    tot["half_bbox"] += steps(b44, g44); tot["half_exact"] += steps(e44, g44)
    g42 = [[(2*(w >> 1))*4 + 2*(w & 1), (2*(w >> 1))*4 + 2*(w & 1)+1, (2*(w >> 1)+1)*4 + 2*(w & 1), (2*(w >> 1)+1)*4 + 2*(w & 1)+1] for w in range(8)]
    tot["quarter_bbox"] += steps(b42, g42); tot["quarter_exact"] += steps(e42, g42)
print({k: int(v) for k, v in tot.items()})
print("cur visits per pair-warp %.3f, exact %.3f, valid lanes/visit %.1f" % (tot["cur"]/(8*tot["pairs"]), tot["exact"]/(8*tot["pairs"]), tot["valid_lanes"]/tot["cur"]))
for k in ("half_bbox", "half_exact", "quarter_bbox", "quarter_exact"):
    print(k, "steps vs cur: %.3f" % (tot[k]/tot["cur"]))
