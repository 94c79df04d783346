"""Level-1 boundary: the exact call surface of the reference's external extension
``diff_gaussian_rasterization`` as TranSplat uses it
(/root/reference/src/model/decoder/cuda_splatting.py:6-9,108-133,201-226,323-348):

    settings = GaussianRasterizationSettings(image_height=..., ..., debug=False)   # 12 fields
    image, radii = GaussianRasterizer(settings)(means3D=, means2D=, shs=, colors_precomp=,
                                                opacities=, cov3D_precomp=)

The top-level package ``diff_gaussian_rasterization/`` of this repository re-exports these names,
so the UNMODIFIED reference file imports and runs against libtsplat.so.
"""
from __future__ import annotations

import ctypes as C
from typing import NamedTuple

import torch
from torch import nn

from . import _lib
from .cameras import pack_camera_from_settings
from .rasterize import RenderConfig, rasterize


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _scalar(x) -> float:
    """tanfov arrives as a float (cuda_splatting.py:111-112) or a 0-d / 1-element tensor (:204-205)."""
    if torch.is_tensor(x):
        return float(x.reshape(-1)[0].item()) if x.numel() == 1 else float(x.item())
    return float(x)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings: GaussianRasterizationSettings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions: torch.Tensor) -> torch.Tensor:
        s = self.raster_settings
        with torch.no_grad():
            pos = positions.contiguous().float()
            if not pos.is_cuda:
                raise RuntimeError("markVisible needs CUDA tensors (no CPU fallback)")
            out = torch.empty(pos.shape[0], dtype=torch.uint8, device=pos.device)
            vm = s.viewmatrix.contiguous().float()
            pm = s.projmatrix.contiguous().float()
            stream = torch.cuda.current_stream(pos.device)
            _lib.check(_lib.lib().tsplat_mark_visible(pos.shape[0], C.c_void_p(pos.data_ptr()),
                                                      C.c_void_p(vm.data_ptr()), C.c_void_p(pm.data_ptr()),
                                                      C.c_void_p(out.data_ptr()), C.c_void_p(stream.cuda_stream)),
                       "tsplat_mark_visible")
            return out.bool()

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        s = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception("Please provide excatly one of either SHs or precomputed colors!")
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception(
                "Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        if means3D.dim() != 2 or means3D.shape[1] != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")
        if s.prefiltered:
            raise RuntimeError("prefiltered=True is not supported (the reference traps the kernel on a culled point)")
        if cov3D_precomp is None:
            cov3D_precomp = _cov3d_from_scale_rotation(scales, rotations, s.scale_modifier)
        feat = shs if shs is not None else colors_precomp
        cam = pack_camera_from_settings(s.viewmatrix, s.projmatrix, s.campos, _scalar(s.tanfovx),
                                        _scalar(s.tanfovy), s.bg)
        cfg = RenderConfig(H=int(s.image_height), W=int(s.image_width), V=1, sh_degree=int(s.sh_degree),
                           want_radii=True, want_means2d_grad=means2D is not None and means2D.requires_grad,
                           debug=bool(s.debug))
        color, _, radii = rasterize(means3D[None], cov3D_precomp[None], feat[None], opacities.reshape(1, -1),
                                    cam, cfg, means2d=means2D if cfg.want_means2d_grad else None)
        return color[0], radii[0]


class _Cov3D(torch.autograd.Function):
    """computeCov3D of the extension (the scales / rotations input mode; unused by TranSplat) and its backward, one
    kernel each (csrc/cov3d.cu): Sigma = R diag(s^2) R^T, s = scale_modifier * scale, quaternion (r, x, y, z) used as
    given - the extension does not normalise it - upper triangle."""

    @staticmethod
    def forward(ctx, scales, rotations, mod):
        if not scales.is_cuda:
            raise RuntimeError("transplat_b200 renders on a CUDA device only (no CPU fallback)")
        scales, rotations = scales.contiguous().float(), rotations.contiguous().float()
        G = scales.shape[0]
        cov6 = torch.empty((G, 6), dtype=torch.float32, device=scales.device)
        with torch.cuda.device(scales.device):
            st = torch.cuda.current_stream(scales.device).cuda_stream
            _lib.check(_lib.lib().tsplat_cov3d_forward(G, scales.data_ptr(), rotations.data_ptr(), float(mod),
                                                       cov6.data_ptr(), C.c_void_p(st)), "tsplat_cov3d_forward")
        ctx.save_for_backward(scales, rotations)
        ctx.mod = float(mod)
        return cov6

    @staticmethod
    def backward(ctx, g_cov6):
        scales, rotations = ctx.saved_tensors
        G = scales.shape[0]
        g_cov6 = g_cov6.contiguous().float()
        g_s, g_q = torch.empty_like(scales), torch.empty_like(rotations)
        with torch.cuda.device(scales.device):
            st = torch.cuda.current_stream(scales.device).cuda_stream
            _lib.check(_lib.lib().tsplat_cov3d_backward(G, scales.data_ptr(), rotations.data_ptr(), ctx.mod,
                                                        g_cov6.data_ptr(), g_s.data_ptr(), g_q.data_ptr(),
                                                        C.c_void_p(st)), "tsplat_cov3d_backward")
        return g_s, g_q, None


def _cov3d_from_scale_rotation(scales, rotations, mod):
    if scales.dim() != 2 or scales.shape[1] != 3 or rotations.dim() != 2 or rotations.shape[1] != 4:
        raise RuntimeError("scales must have dimensions (num_points, 3) and rotations (num_points, 4)")
    return _Cov3D.apply(scales, rotations, float(mod))
