"""Generates tests/golden/ref_adapter_small.npz IN THE AUTHORING CONTAINER (needs /root/reference): outputs and
autograd gradients of the UNMODIFIED reference GaussianAdapter.forward (gaussian_adapter.py:50-98) called with the
shapes of its only call site (encoder_trans.py:307-318), for sh_degree 4 and 2.  Wigner-D matrices are random
orthogonal stand-ins fed through the reference's own rotate_sh_with_precomputed (e3nn is absent; oracle/adapter_ref.py).
    python tests/golden/make_golden_adapter.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.adapter_ref import load_reference_adapter  # noqa: E402

ref = load_reference_adapter()
out = {}
for tag, deg, (b, v, h, w), (smin, smax) in (("d4", 4, (2, 2, 12, 20), (0.5, 15.0)), ("d2", 2, (1, 3, 8, 8), (0.1, 4.0))):
    g = torch.Generator().manual_seed(100 + deg)
    d_sh = (deg + 1) ** 2
    n = h * w
    rnd = lambda *s: torch.randn(*s, generator=g)
    # context cameras: random rotations (QR), translations; normalised intrinsics
    q, _ = torch.linalg.qr(rnd(b, v, 3, 3))
    q = q * torch.linalg.det(q)[..., None, None].sign()
    ext = torch.eye(4).repeat(b, v, 1, 1)
    ext[..., :3, :3] = q
    ext[..., :3, 3] = rnd(b, v, 3)
    intr = torch.eye(3).repeat(b, v, 1, 1)
    intr[..., 0, 0] = 0.8 + 0.3 * torch.rand(b, v, generator=g)
    intr[..., 1, 1] = 0.8 + 0.3 * torch.rand(b, v, generator=g)
    intr[..., 0, 2] = 0.5
    intr[..., 1, 2] = 0.5
    wig = []
    for l in range(deg + 1):
        m, _ = torch.linalg.qr(rnd(b, v, 2 * l + 1, 2 * l + 1))
        wig.append(m.contiguous())
    coords = torch.rand(b, v, n, 1, 1, 2, generator=g).requires_grad_(True)
    depths = (1.0 + 9.0 * torch.rand(b, v, n, 1, 1, generator=g)).requires_grad_(True)
    opac = torch.rand(b, v, n, 1, 1, generator=g).requires_grad_(True)
    raw = rnd(b, v, n, 1, 1, 7 + 3 * d_sh).requires_grad_(True)
    adapter = ref.GaussianAdapter(ref.GaussianAdapterCfg(smin, smax, deg))
    ref.WIGNER_HOLDER.WIGNER = wig
    G = adapter.forward(ext[:, :, None, None, None], intr[:, :, None, None, None], coords, depths, opac, raw, (h, w))
    ws = dict(means=rnd(*G.means.shape), covariances=rnd(*G.covariances.shape), harmonics=rnd(*G.harmonics.shape),
              opacities=rnd(*G.opacities.shape))
    loss = sum((getattr(G, k) * ws[k]).sum() for k in ws)
    loss.backward()
    flat = lambda t: t.detach().reshape(b * v, n, *t.shape[5:])
    out.update({f"{tag}_cfg": np.array([deg, b, v, h, w, smin, smax], np.float64),
                f"{tag}_extrinsics": ext.reshape(b * v, 4, 4), f"{tag}_intrinsics": intr.reshape(b * v, 3, 3),
                f"{tag}_coordinates": flat(coords), f"{tag}_depths": flat(depths), f"{tag}_opacities": flat(opac),
                f"{tag}_raw": flat(raw)})
    for l, m in enumerate(wig):
        out[f"{tag}_wigner{l}"] = m.reshape(b * v, 2 * l + 1, 2 * l + 1)
    for k in ("means", "covariances", "harmonics", "opacities", "scales", "rotations"):
        out[f"{tag}_out_{k}"] = flat(getattr(G, k))
    for k in ws:
        out[f"{tag}_w_{k}"] = flat(ws[k])
    out.update({f"{tag}_g_coordinates": flat(coords.grad), f"{tag}_g_depths": flat(depths.grad),
                f"{tag}_g_opacities": flat(opac.grad), f"{tag}_g_raw": flat(raw.grad)})
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_adapter_small.npz")
np.savez_compressed(path, **{k: (v.numpy() if torch.is_tensor(v) else v) for k, v in out.items()})
print("wrote", path, os.path.getsize(path), "bytes", {k: tuple(v.shape) for k, v in out.items() if k.startswith("d4_out")})
