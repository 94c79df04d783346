"""The C oracle's hand-written backward against float64 autograd of a dense
formulation of the same forward math (oracle/torch_dense.py), which takes its own discrete decisions
(visibility, radius, tile rectangle, depth order) in binary64: those are COMPARED with the oracle's
integer outputs here, not borrowed from them.  CPU only."""
import numpy as np
import pytest
import torch

import oracle
from oracle import torch_dense as td
from transplat_b200.cameras import pack_cameras
from transplat_b200.synthetic import make_scene

H, W = 40, 48  # W is a multiple of 16, H is not: partial tiles are covered


def _setup():
    sc = make_scene(n_ctx=1, h=24, w=20, n_views=2, render_hw=(H, W), seed=5, near=1.5, far=50.0,
                    d0=3.0, d1=8.0)
    G = sc.means.shape[0]
    bg = torch.tensor([[0.2, 0.5, 0.7]]).expand(2, 3)
    cams = pack_cameras(sc.extrinsics, sc.intrinsics, sc.near, sc.far, bg).numpy()
    cov6 = sc.covariances.reshape(G, 9)[:, [0, 1, 2, 4, 5, 8]].contiguous() * 9
    sh = sc.harmonics.permute(0, 2, 1).contiguous()
    sh[:, 0, :] -= 1.0  # force some colour clamping
    return sc, cams, cov6, sh


@pytest.mark.parametrize("depth_mode", [0, 1, 2, 3, 4])
def test_oracle_backward_matches_float64_autograd(depth_mode):
    sc, cams, cov6, sh = _setup()
    scene = oracle.Scene(sc.means.numpy(), cov6.numpy(), sh.numpy(), sc.opacities.numpy(), 4)
    cam = cams[1]
    st = oracle.forward(scene, cam, H, W, depth_mode=depth_mode)
    assert (st.clamped > 0).sum() > 10 and st.num_rendered > 1000
    g = torch.Generator().manual_seed(1)
    dLc = torch.randn(3, H, W, generator=g, dtype=torch.float64)
    dLd = torch.randn(H, W, generator=g, dtype=torch.float64)
    acc = oracle.backward(scene, cam, st, dLc.numpy(), dLd.numpy() if depth_mode else None,
                          depth_mode=depth_mode)
    m = sc.means.double().requires_grad_(True)
    c6 = cov6.double().requires_grad_(True)
    s = sh.double().requires_grad_(True)
    o = sc.opacities.double().requires_grad_(True)
    geo = {}
    col, dep = td.render_dense(m, c6, s, o, cam, H, W, 4, depth_mode=depth_mode, geometry_out=geo)
    _compare_geometry(geo, st)
    assert np.abs(col.detach().numpy() - st.color).max() < 1e-5  # tolerance: 1e-5 abs (north_star)
    loss = (col * dLc).sum()
    if depth_mode:
        # oracle depth = ((x+x)+x)/3 of the rendered channel (cuda_splatting.py:417)
        assert np.abs(dep.detach().numpy() - st.depth).max() < 1e-5
        loss = loss + (dep * dLd).sum()
    loss.backward()
    for a, b in ((acc.means, m.grad), (acc.cov, c6.grad), (acc.sh, s.grad), (acc.opacity, o.grad)):
        b = b.numpy()
        assert np.abs(a - b).max() <= 1e-4 * np.abs(b).max()  # tolerance: 1e-4 relative (north_star)


def _compare_geometry(geo, st):
    """The dense formulation's own binary64 radii / visibility / rectangles / depth order against the oracle's
    binary32 ones: identical, except where the real-valued quantity lies within 1e-4 of the decision boundary."""
    rr = geo["radius_real"]
    safe = np.abs(rr - np.round(rr)) > 1e-4
    assert safe.mean() > 0.99
    both = safe & (np.abs(geo["tz"] - 0.2) > 1e-6)
    assert np.array_equal(geo["radii"][both] > 0, st.radii[both] > 0)
    vis = both & (st.radii > 0)
    assert vis.sum() > 100
    assert np.array_equal(geo["radii"][vis], st.radii[vis])
    assert np.abs(geo["px"][vis] - st.xy[vis, 0]).max() < 1e-3 and np.abs(geo["py"][vis] - st.xy[vis, 1]).max() < 1e-3
    x0, y0, x1, y1 = geo["rect"][vis].T
    # rectangle corners move only where (p +- r) / 16 is within rounding distance of an integer
    edge = np.stack(((geo["px"][vis] - st.radii[vis]) / 16, (geo["px"][vis] + st.radii[vis] + 15) / 16,
                     (geo["py"][vis] - st.radii[vis]) / 16, (geo["py"][vis] + st.radii[vis] + 15) / 16), -1)
    firm = (np.abs(edge - np.round(edge)) > 1e-5).all(-1)
    assert firm.mean() > 0.99
    assert np.array_equal(((x1 - x0) * (y1 - y0))[firm], st.tiles_touched[vis][firm].astype(np.int64))
    # depth order: the oracle's sorted list of any tile is a subsequence of this order
    if np.array_equal(geo["radii"] > 0, st.radii > 0):
        rank = np.empty(st.G, np.int64)
        rank[geo["order"]] = np.arange(geo["order"].size)
        for lo, hi in st.ranges:
            r = rank[st.vals[lo:hi].astype(np.int64)]
            z = geo["tz"][st.vals[lo:hi].astype(np.int64)]
            bad = (r[1:] < r[:-1]) & (np.abs(z[1:] - z[:-1]) > 1e-6 * np.abs(z[1:]))
            assert not bad.any()


def test_colors_precomp_and_degree_sweep():
    sc, cams, cov6, sh = _setup()
    for deg in (0, 1, 2, 3):
        scene = oracle.Scene(sc.means.numpy(), cov6.numpy(), sh.numpy(), sc.opacities.numpy(), deg)
        st = oracle.forward(scene, cams[0], H, W)
        col, _ = td.render_dense(sc.means.double(), cov6.double(), sh.double(), sc.opacities.double(),
                                 cams[0], H, W, deg)
        assert np.abs(col.numpy() - st.color).max() < 1e-5
    pre = torch.rand(sc.means.shape[0], 3)
    scene = oracle.Scene(sc.means.numpy(), cov6.numpy(), pre.numpy(), sc.opacities.numpy(), 0)
    st = oracle.forward(scene, cams[0], H, W)
    col, _ = td.render_dense(sc.means.double(), cov6.double(), pre.double(), sc.opacities.double(),
                             cams[0], H, W, 0)
    assert np.abs(col.numpy() - st.color).max() < 1e-5
