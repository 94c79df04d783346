"""tsplat_adapter_forward / backward (csrc/adapter.cu) through transplat_b200.adapter.GaussianAdapter against
(1) the fixture recorded from the UNMODIFIED reference GaussianAdapter (tests/golden/ref_adapter_small.npz) and
(2) the CPU restatement (oracle/adapter_ref.py, itself held to that fixture) at the C2 size and on ragged sizes.
Floating point: outputs and gradients within 1e-5 of each tensor's magnitude (north_star: 1e-5 abs / 1e-4 rel)."""
import numpy as np
import pytest
import torch

from oracle.adapter_ref import adapter_restated
from test_adapter_cpu import load, rel

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _run_gpu(ext, intr, coords, depths, opac, raw, wig, deg, hw, smin, smax, weights=None, lead=None):
    from transplat_b200.adapter import GaussianAdapter, GaussianAdapterCfg
    N, n = depths.shape
    lead = lead or (1, N, n, 1, 1)
    b, v = lead[:2]
    cu = lambda t: t.cuda()
    leaves = [cu(t).reshape(*lead, *t.shape[2:]).clone().requires_grad_(True) for t in (coords, depths, opac, raw)]
    ad = GaussianAdapter(GaussianAdapterCfg(smin, smax, deg)).cuda()
    G = ad.forward(cu(ext).reshape(b, v, 1, 1, 1, 4, 4), cu(intr).reshape(b, v, 1, 1, 1, 3, 3), *leaves, hw,
                   wigner_matrices=[cu(w).reshape(b, v, *w.shape[1:]) for w in wig])
    outs = {k: getattr(G, k).reshape(N, n, *getattr(G, k).shape[5:]) for k in
            ("means", "covariances", "harmonics", "opacities", "scales", "rotations")}
    grads = None
    if weights is not None:
        loss = sum((outs[k] * cu(weights[k])).sum() for k in weights)
        loss.backward()
        grads = [l.grad.reshape(N, n, *l.shape[5:]) for l in leaves]
    return outs, grads


@pytest.mark.parametrize("tag", ["d4", "d2"])
def test_adapter_matches_reference_fixture(tag):
    z, t, wig, deg, hw, smin, smax = load(tag)
    _, b, v, h, w, _, _ = z[f"{tag}_cfg"]
    weights = {k: t(f"w_{k}") for k in ("means", "covariances", "harmonics", "opacities")}
    outs, grads = _run_gpu(t("extrinsics"), t("intrinsics"), t("coordinates"), t("depths"), t("opacities"), t("raw"), wig,
                           deg, hw, smin, smax, weights, lead=(int(b), int(v), int(h * w), 1, 1))
    for k, o in outs.items():
        assert rel(o, z[f"{tag}_out_{k}"]) <= TOL, (tag, k)
    for g, k in zip(grads, ("coordinates", "depths", "opacities", "raw")):
        assert rel(g, z[f"{tag}_g_{k}"]) <= TOL, (tag, k)


@pytest.mark.parametrize("N,n,deg,hw", [(2, 65536, 4, (256, 256)), (3, 1000, 3, (20, 50)), (1, 77, 0, (7, 11)), (2, 130, 1, (10, 13))])
def test_adapter_matches_restatement(N, n, deg, hw):
    """C2 size (2 context views x 65536 pixels, degree 4) and ragged sizes (partial CTAs, rows that break the 16-B
    alignment of the TMA path, degrees 0-3)."""
    g = torch.Generator().manual_seed(7 + n)
    d_sh = (deg + 1) ** 2
    q, _ = torch.linalg.qr(torch.randn(N, 3, 3, generator=g))
    ext = torch.eye(4).repeat(N, 1, 1)
    ext[:, :3, :3] = q
    ext[:, :3, 3] = torch.randn(N, 3, generator=g)
    intr = torch.eye(3).repeat(N, 1, 1)
    intr[:, 0, 0], intr[:, 1, 1], intr[:, 0, 2], intr[:, 1, 2] = 0.9, 0.85, 0.5, 0.5
    wig = [torch.linalg.qr(torch.randn(N, 2 * l + 1, 2 * l + 1, generator=g))[0].contiguous() for l in range(deg + 1)]
    coords = torch.rand(N, n, 2, generator=g)
    depths = 1 + 9 * torch.rand(N, n, generator=g)
    opac = torch.rand(N, n, generator=g)
    raw = torch.randn(N, n, 7 + 3 * d_sh, generator=g)
    weights = dict(means=torch.randn(N, n, 3, generator=g), covariances=torch.randn(N, n, 3, 3, generator=g),
                   harmonics=torch.randn(N, n, 3, d_sh, generator=g), opacities=torch.randn(N, n, generator=g))
    outs, grads = _run_gpu(ext, intr, coords, depths, opac, raw, wig, deg, hw, 0.5, 15.0, weights)
    leaves = [t.clone().requires_grad_(True) for t in (coords, depths, opac, raw)]
    want = adapter_restated(ext, intr, *leaves, wig, hw, 0.5, 15.0, deg)
    names = ("means", "covariances", "harmonics", "opacities", "scales", "rotations")
    for o, k in zip(want, names):
        assert rel(outs[k], o.detach().numpy()) <= TOL, k
    sum((o * weights[k]).sum() for o, k in zip(want[:4], names[:4])).backward()
    for gg, leaf, k in zip(grads, leaves, ("coordinates", "depths", "opacities", "raw")):
        assert rel(gg, leaf.grad.numpy()) <= TOL, k


def test_adapter_feeds_the_rasterizer_unchanged():
    """Adapter outputs, flattened as the encoder does (encoder_trans.py:336-353), render through DecoderSplattingCUDA."""
    from transplat_b200.adapter import GaussianAdapter, GaussianAdapterCfg
    from transplat_b200.decoder import DecoderSplattingCUDA, Gaussians
    b, v, h, w, deg = 1, 2, 32, 32, 4
    g = torch.Generator().manual_seed(3)
    ext = torch.eye(4).repeat(b, v, 1, 1)
    ext[:, 1, 0, 3] = 0.3
    intr = torch.tensor([[0.9, 0, 0.5], [0, 0.9, 0.5], [0, 0, 1.0]]).repeat(b, v, 1, 1)
    ys, xs = torch.meshgrid((torch.arange(h) + 0.5) / h, (torch.arange(w) + 0.5) / w, indexing="ij")
    coords = torch.stack((xs, ys), -1).reshape(1, 1, h * w, 1, 1, 2).expand(b, v, -1, -1, -1, -1).contiguous()
    depths = (2 + 3 * torch.rand(b, v, h * w, 1, 1, generator=g)).cuda().requires_grad_(True)
    raw = torch.randn(b, v, h * w, 1, 1, 7 + 75, generator=g).cuda().requires_grad_(True)
    opac = torch.rand(b, v, h * w, 1, 1, generator=g).cuda()
    wig = [torch.eye(2 * l + 1).repeat(b, v, 1, 1).cuda() for l in range(deg + 1)]
    ad = GaussianAdapter(GaussianAdapterCfg(0.5, 15.0, deg)).cuda()
    G = ad.forward(ext.cuda()[:, :, None, None, None], intr.cuda()[:, :, None, None, None], coords.cuda(), depths, opac, raw,
                   (h, w), wigner_matrices=wig)
    flat = Gaussians(G.means.reshape(b, -1, 3), G.covariances.reshape(b, -1, 3, 3), G.harmonics.reshape(b, -1, 3, 25),
                     G.opacities.reshape(b, -1))
    dec = DecoderSplattingCUDA().cuda()
    out = dec.forward(flat, ext.cuda(), intr.cuda(), torch.ones(b, v).cuda(), torch.full((b, v), 100.0).cuda(), (h, w),
                      depth_mode="depth")
    assert float(out.color.detach().mean()) > 0.01 and float(out.depth.detach().max()) > 1.0
    (out.color.sum() + out.depth.sum()).backward()
    assert float(raw.grad.abs().max()) > 0 and float(depths.grad.abs().max()) > 0
    assert torch.isfinite(raw.grad).all() and torch.isfinite(depths.grad).all()
