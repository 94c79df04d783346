"""TEST INFRASTRUCTURE ONLY (never imported by the product path).

CPU restatement of the step BEFORE the render path, SURVEY 8f rank 2: the per-pixel Gaussian producer
``GaussianAdapter.forward`` (/root/reference/src/model/encoder/common/gaussian_adapter.py:50-98) with
``build_covariance`` / ``quaternion_to_matrix`` (.../common/gaussians.py:7-43), ``get_world_rays``
(/root/reference/src/geometry/projection.py:74-113) and the application of precomputed Wigner-D matrices
``rotate_sh_with_precomputed`` (/root/reference/src/misc/sh_rotation_optimized.py:56-101).

Parity status: PINNED for everything restated here - ``load_reference_adapter`` imports the UNMODIFIED reference
files in the authoring container and tests/golden/make_golden_adapter.py records their outputs and autograd
gradients (tests/golden/ref_adapter_small.npz); tests/test_adapter_cpu.py holds this restatement to that fixture.
NOT restated: the Wigner-D matrices themselves (e3nn.o3.wigner_D, an external dependency absent from this image):
they are an INPUT here and in the CUDA kernel, exactly as in the reference's own precomputed path
(sh_rotation_optimized.py:27-53).
"""
from __future__ import annotations

import importlib.util
import sys
import types

import torch


def sh_mask(sh_degree: int) -> torch.Tensor:
    """gaussian_adapter.py:40-46"""
    m = torch.ones(((sh_degree + 1) ** 2,), dtype=torch.float32)
    for degree in range(1, sh_degree + 1):
        m[degree ** 2:(degree + 1) ** 2] = 0.1 * 0.25 ** degree
    return m


def adapter_restated(extrinsics, intrinsics, coordinates, depths, opacities, raw, wigner, image_shape, scale_min,
                     scale_max, sh_degree, eps: float = 1e-8):
    """extrinsics [N,4,4], intrinsics [N,3,3], coordinates [N,n,2], depths [N,n], opacities [N,n], raw [N,n,7+3*d_sh],
    wigner: list over degree of [N,2l+1,2l+1].  Returns means [N,n,3], covariances [N,n,3,3], harmonics [N,n,3,d_sh],
    opacities, scales [N,n,3], rotations [N,n,4] - plain torch, differentiable."""
    d_sh = (sh_degree + 1) ** 2
    h, w = image_shape
    scales, rotations, sh = raw.split((3, 4, 3 * d_sh), dim=-1)
    scales = scale_min + (scale_max - scale_min) * scales.sigmoid()                      # :62-65
    pixel_size = 1 / torch.tensor((w, h), dtype=torch.float32)                           # :67
    mult = (0.1 * torch.einsum("nij,j->ni", intrinsics[:, :2, :2].inverse(), pixel_size)).sum(-1)  # :106-110
    scales = scales * depths[..., None] * mult[:, None, None]                            # :69
    rotations = rotations / (rotations.norm(dim=-1, keepdim=True) + eps)                 # :72
    sh = sh.reshape(*sh.shape[:-1], 3, d_sh) * sh_mask(sh_degree).to(sh.dtype)           # :75-76
    i, j, k, r = rotations.unbind(-1)                                                    # gaussians.py:12-29
    two_s = 2 / ((rotations * rotations).sum(-1) + 1e-8)
    R = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    R = R.reshape(*R.shape[:-1], 3, 3)
    S = scales.diag_embed()
    cov = R @ S @ S.transpose(-1, -2) @ R.transpose(-1, -2)                              # gaussians.py:36-43
    C = extrinsics[:, None, :3, :3]
    cov = C @ cov @ C.transpose(-1, -2)                                                  # :80-81
    hom = torch.cat((coordinates, torch.ones_like(coordinates[..., :1])), -1)            # projection.py:82
    d = torch.einsum("nij,nkj->nki", intrinsics.inverse(), hom)                          # :83-85
    d = d / d.norm(dim=-1, keepdim=True)                                                 # :105
    d = torch.einsum("nij,nkj->nki", extrinsics[:, :3, :3], d)                           # :108-109 (w = 0)
    means = extrinsics[:, None, :3, 3] + d * depths[..., None]                           # :112, gaussian_adapter.py:85
    out = []
    for degree in range(sh_degree + 1):                                                  # sh_rotation_optimized.py:80-99
        D = wigner[degree][:, None, None]
        out.append(torch.einsum("...ij,...j->...i", D, sh[..., degree ** 2:(degree + 1) ** 2]))
    return means, cov, torch.cat(out, -1), opacities, scales, rotations


def load_reference_adapter(ref_root: str = "/root/reference"):
    """Import the UNMODIFIED reference gaussian_adapter.py (+ gaussians.py, projection.py, sh_rotation_optimized.py)
    in the authoring container.  Absent packages are stubbed: jaxtyping (annotations only) and e3nn.o3 (only
    wigner_D / matrix_to_angles live there; the fixture feeds precomputed matrices instead).  ``rotate_sh`` as the
    adapter imports it (sh_rotation.py) is routed to the reference's own rotate_sh_with_precomputed with the
    matrices set in ``mod.WIGNER`` (list over degree of [b,v,2l+1,2l+1])."""
    saved = {k: v for k, v in sys.modules.items() if k == "src" or k.startswith("src.") or k in ("jaxtyping", "e3nn", "e3nn.o3")}
    for k in list(saved):
        del sys.modules[k]
    try:
        jt = types.ModuleType("jaxtyping")

        class _Sub:
            def __class_getitem__(cls, item):
                return cls
        for nm in ('Float', 'Int', 'Int64', 'Int32', 'Bool', 'Shaped', 'UInt8', 'Float32', 'Float64'):
            setattr(jt, nm, _Sub)
        sys.modules["jaxtyping"] = jt
        e3 = types.ModuleType("e3nn")
        o3 = types.ModuleType("e3nn.o3")
        o3.matrix_to_angles = o3.wigner_D = None
        e3.o3 = o3
        sys.modules["e3nn"], sys.modules["e3nn.o3"] = e3, o3
        for name in ("src", "src.model", "src.model.encoder", "src.model.encoder.common", "src.geometry", "src.misc"):
            m = types.ModuleType(name)
            m.__path__ = [f"{ref_root}/{name.replace('.', '/')}"]
            sys.modules[name] = m

        def load(name):
            spec = importlib.util.spec_from_file_location(name, f"{ref_root}/{name.replace('.', '/')}.py")
            mod = importlib.util.module_from_spec(spec)
            sys.modules[name] = mod
            spec.loader.exec_module(mod)
            return mod

        load("src.geometry.projection")
        opt = load("src.misc.sh_rotation_optimized")
        holder = types.ModuleType("src.misc.sh_rotation")
        holder.WIGNER = None
        holder.rotate_sh = lambda sh, rotations: opt.rotate_sh_with_precomputed(sh, holder.WIGNER, rotation_batch_dims=2)
        sys.modules["src.misc.sh_rotation"] = holder
        load("src.model.encoder.common.gaussians")
        mod = load("src.model.encoder.common.gaussian_adapter")
        mod.WIGNER_HOLDER = holder
        return mod
    finally:
        for k in list(sys.modules):
            if k == "src" or k.startswith("src.") or k in ("jaxtyping", "e3nn", "e3nn.o3"):
                del sys.modules[k]
        sys.modules.update(saved)
