"""The per-pixel Gaussian producer in front of the render path, with the reference's names and
call surface

  GaussianAdapterCfg, GaussianAdapter, Gaussians   src/model/encoder/common/gaussian_adapter.py:13-118
  precomputed Wigner-D path                        src/misc/sh_rotation_optimized.py:27-101

but ONE kernel forward and ONE backward of libtsplat.so (csrc/adapter.cu) instead of ~40 torch ops.  The Wigner-D
matrices are an input, as in the reference's own precomputed path: ``GaussianAdapter.forward(...,
wigner_matrices=precompute_wigner_d_matrices(c2w_rotations, sh_degree)[0])`` - they depend on the context cameras only
(e3nn.o3.wigner_D stays on the caller's side).  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch
from torch import Tensor, nn

from . import _lib


@dataclass
class Gaussians:
    """gaussian_adapter.py:13-20 (the adapter's own record: six fields)."""
    means: Tensor
    covariances: Tensor
    scales: Tensor
    rotations: Tensor
    harmonics: Tensor
    opacities: Tensor


@dataclass
class GaussianAdapterCfg:
    gaussian_scale_min: float
    gaussian_scale_max: float
    sh_degree: int


class TsplatAdapterProblem(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("N", "n", "sh_degree", "H", "W")] + [
        (n, C.c_float) for n in ("scale_min", "scale_max", "eps")] + [
        (n, C.c_void_p) for n in ("extrinsics", "intrinsics", "coordinates", "depths", "raw", "wigner", "sh_mask")]


def pack_wigner(wigner_matrices: Sequence[Tensor], N: int) -> Tensor:
    """list over degree of [..., 2l+1, 2l+1] (leading dims flatten to N views) -> [N, sum (2l+1)^2] float32."""
    parts = []
    for l, D in enumerate(wigner_matrices):
        m = 2 * l + 1
        if tuple(D.shape[-2:]) != (m, m):
            raise RuntimeError(f"wigner_matrices[{l}] must end in ({m}, {m}), got {tuple(D.shape)}")
        parts.append(D.reshape(-1, m * m))
        if parts[-1].shape[0] != N:
            raise RuntimeError(f"wigner_matrices[{l}] holds {parts[-1].shape[0]} views, expected {N}")
    return torch.cat(parts, dim=-1).to(torch.float32).contiguous()


def _problem(cfg, N, n, image_shape, eps, ext, intr, coords, depths, raw, wig, mask) -> TsplatAdapterProblem:
    p = TsplatAdapterProblem()
    p.N, p.n, p.sh_degree = N, n, cfg.sh_degree
    p.H, p.W = int(image_shape[0]), int(image_shape[1])
    p.scale_min, p.scale_max, p.eps = cfg.gaussian_scale_min, cfg.gaussian_scale_max, eps
    p.extrinsics, p.intrinsics, p.coordinates, p.depths, p.raw, p.wigner, p.sh_mask = (
        t.data_ptr() for t in (ext, intr, coords, depths, raw, wig, mask))
    return p


class _Adapter(torch.autograd.Function):
    @staticmethod
    def forward(ctx, coords, depths, raw, ext, intr, wig, mask, cfg, image_shape, eps):
        L = _lib.lib()
        if not raw.is_cuda:
            raise RuntimeError("transplat_b200.adapter runs on a CUDA device only (no CPU fallback)")
        coords, depths, raw = (t.contiguous() for t in (coords, depths, raw))
        N, n = depths.shape
        d_sh = (cfg.sh_degree + 1) ** 2
        dev = raw.device
        means = torch.empty((N, n, 3), dtype=torch.float32, device=dev)
        cov = torch.empty((N, n, 3, 3), dtype=torch.float32, device=dev)
        harm = torch.empty((N, n, 3, d_sh), dtype=torch.float32, device=dev)
        scales = torch.empty((N, n, 3), dtype=torch.float32, device=dev)
        rot = torch.empty((N, n, 4), dtype=torch.float32, device=dev)
        prob = _problem(cfg, N, n, image_shape, eps, ext, intr, coords, depths, raw, wig, mask)
        with torch.cuda.device(dev):
            st = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(L.tsplat_adapter_forward(C.byref(prob), C.c_void_p(means.data_ptr()), C.c_void_p(cov.data_ptr()),
                                                C.c_void_p(harm.data_ptr()), C.c_void_p(scales.data_ptr()),
                                                C.c_void_p(rot.data_ptr()), C.c_void_p(st)), "tsplat_adapter_forward")
        ctx.save_for_backward(coords, depths, raw, ext, intr, wig, mask)
        ctx.cfg, ctx.image_shape, ctx.eps = cfg, image_shape, eps
        ctx.mark_non_differentiable(scales, rot)  # export-only outputs (gaussian_adapter.py:91-94)
        return means, cov, harm, scales, rot

    @staticmethod
    def backward(ctx, g_means, g_cov, g_harm, _gs, _gr):
        L = _lib.lib()
        coords, depths, raw, ext, intr, wig, mask = ctx.saved_tensors
        N, n = depths.shape
        dev = raw.device
        zeros = lambda ref_shape: torch.zeros(ref_shape, dtype=torch.float32, device=dev)
        d_sh = (ctx.cfg.sh_degree + 1) ** 2
        g_means = g_means.contiguous() if g_means is not None else zeros((N, n, 3))
        g_cov = g_cov.contiguous() if g_cov is not None else zeros((N, n, 3, 3))
        g_harm = g_harm.contiguous() if g_harm is not None else zeros((N, n, 3, d_sh))
        g_coords, g_depths, g_raw = torch.empty_like(coords), torch.empty_like(depths), torch.empty_like(raw)
        prob = _problem(ctx.cfg, N, n, ctx.image_shape, ctx.eps, ext, intr, coords, depths, raw, wig, mask)
        with torch.cuda.device(dev):
            st = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(L.tsplat_adapter_backward(C.byref(prob), C.c_void_p(g_means.data_ptr()), C.c_void_p(g_cov.data_ptr()),
                                                 C.c_void_p(g_harm.data_ptr()), C.c_void_p(g_coords.data_ptr()),
                                                 C.c_void_p(g_depths.data_ptr()), C.c_void_p(g_raw.data_ptr()),
                                                 C.c_void_p(st)), "tsplat_adapter_backward")
        return g_coords, g_depths, g_raw, None, None, None, None, None, None, None


class GaussianAdapter(nn.Module):
    """gaussian_adapter.py:30-118: same constructor, ``sh_mask`` buffer, ``d_sh`` / ``d_in`` properties and
    ``forward`` argument order; ``forward`` takes the precomputed Wigner-D matrices as an extra keyword."""

    def __init__(self, cfg: GaussianAdapterCfg):
        super().__init__()
        self.cfg = cfg
        self.register_buffer("sh_mask", torch.ones((self.d_sh,), dtype=torch.float32), persistent=False)
        for degree in range(1, self.cfg.sh_degree + 1):
            self.sh_mask[degree ** 2:(degree + 1) ** 2] = 0.1 * 0.25 ** degree

    @property
    def d_sh(self) -> int:
        return (self.cfg.sh_degree + 1) ** 2

    @property
    def d_in(self) -> int:
        return 7 + 3 * self.d_sh

    def forward(self, extrinsics: Tensor, intrinsics: Tensor, coordinates: Tensor, depths: Tensor, opacities: Tensor,
                raw_gaussians: Tensor, image_shape: tuple, eps: float = 1e-8,
                wigner_matrices: Optional[List[Tensor]] = None) -> Gaussians:
        """Shapes of the reference's call site (encoder_trans.py:307-318): extrinsics [b,v,1,1,1,4,4], intrinsics
        [b,v,1,1,1,3,3], coordinates [b,v,r,srf,1|spp,2], depths / opacities [b,v,r,srf,spp], raw_gaussians
        [b,v,r,srf,1|spp,d_in].  Per-view tensors may also come as [b,v,4,4] / [b,v,3,3]."""
        if wigner_matrices is None:
            raise RuntimeError("pass wigner_matrices=precompute_wigner_d_matrices(extrinsics[..., :3, :3], sh_degree)[0] "
                               "(src/misc/sh_rotation_optimized.py:27-53): the Wigner-D matrices depend on the cameras only")
        lead = depths.shape                                  # [b, v, r, srf, spp]
        b, v = lead[:2]
        N = b * v
        n = 1
        for x in lead[2:]:
            n *= x
        ext = extrinsics.reshape(N, 4, 4).to(torch.float32).contiguous()
        intr = intrinsics.reshape(N, 3, 3).to(torch.float32).contiguous()
        coords = coordinates.expand(*lead, 2).reshape(N, n, 2)
        raw = raw_gaussians.expand(*lead, self.d_in).reshape(N, n, self.d_in)
        wig = pack_wigner(wigner_matrices, N).to(raw.device)
        means, cov, harm, scales, rot = _Adapter.apply(coords, depths.reshape(N, n), raw, ext, intr, wig,
                                                       self.sh_mask.to(raw.device), self.cfg, tuple(image_shape), float(eps))
        return Gaussians(means=means.reshape(*lead, 3), covariances=cov.reshape(*lead, 3, 3),
                         scales=scales.reshape(*lead, 3), rotations=rot.reshape(*lead, 4),
                         harmonics=harm.reshape(*lead, 3, self.d_sh), opacities=opacities)

    def get_scale_multiplier(self, intrinsics: Tensor, pixel_size: Tensor, multiplier: float = 0.1) -> Tensor:
        """gaussian_adapter.py:100-111 (kept for callers that use it on its own; the kernel computes it per view)."""
        xy = multiplier * torch.einsum("...ij,j->...i", torch.linalg.inv_ex(intrinsics[..., :2, :2])[0], pixel_size)
        return xy.sum(dim=-1)
