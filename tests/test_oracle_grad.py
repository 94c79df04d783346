"""The C oracle's hand-written backward against float64 autograd of a dense
formulation of the same forward math (oracle/torch_dense.py).  CPU only."""
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
    col, dep = td.render_dense(m, c6, s, o, cam, H, W, 4, st.radii, st.depths, depth_mode=depth_mode)
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


def test_colors_precomp_and_degree_sweep():
    sc, cams, cov6, sh = _setup()
    for deg in (0, 1, 2, 3):
        scene = oracle.Scene(sc.means.numpy(), cov6.numpy(), sh.numpy(), sc.opacities.numpy(), deg)
        st = oracle.forward(scene, cams[0], H, W)
        col, _ = td.render_dense(sc.means.double(), cov6.double(), sh.double(), sc.opacities.double(),
                                 cams[0], H, W, deg, st.radii, st.depths)
        assert np.abs(col.numpy() - st.color).max() < 1e-5
    pre = torch.rand(sc.means.shape[0], 3)
    scene = oracle.Scene(sc.means.numpy(), cov6.numpy(), pre.numpy(), sc.opacities.numpy(), 0)
    st = oracle.forward(scene, cams[0], H, W)
    col, _ = td.render_dense(sc.means.double(), cov6.double(), pre.double(), sc.opacities.double(),
                             cams[0], H, W, 0, st.radii, st.depths)
    assert np.abs(col.numpy() - st.color).max() < 1e-5
