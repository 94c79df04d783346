"""CPU stand-in for the ``diff_gaussian_rasterization`` module, on top of the C oracle.
TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

It exposes the call surface the reference uses (cuda_splatting.py:6-9,108-133):
``GaussianRasterizationSettings`` (12-field NamedTuple) and ``GaussianRasterizer`` (nn.Module with
an autograd.Function underneath), so that the UNMODIFIED reference file
/root/reference/src/model/decoder/cuda_splatting.py can be executed on the CPU against the oracle
(``load_reference_cuda_splatting``) to produce the golden fixtures under tests/golden/.
"""
from __future__ import annotations

import importlib.util
import sys
import types
from typing import NamedTuple

import numpy as np
import torch
from torch import nn

from . import GradAccum, Scene, backward, forward, make_camera, mark_visible


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


def _np(t):
    return t.detach().cpu().contiguous().numpy()


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, cov3D_precomp, settings):
        s = settings
        use_sh = sh is not None and sh.numel() > 0
        feat = sh if use_sh else colors_precomp
        scene = Scene(_np(means3D), _np(cov3D_precomp), _np(feat), _np(opacities), s.sh_degree)
        cam = make_camera(_np(s.viewmatrix), _np(s.projmatrix), _np(s.campos), float(s.tanfovx),
                          float(s.tanfovy), _np(s.bg))
        st = forward(scene, cam, s.image_height, s.image_width)
        ctx.scene, ctx.cam, ctx.st, ctx.use_sh = scene, cam, st, use_sh
        ctx.last_state = st
        _RasterizeGaussians.last_state = st
        return torch.from_numpy(st.color.copy()), torch.from_numpy(st.radii.copy())

    @staticmethod
    def backward(ctx, grad_color, _grad_radii):
        acc = backward(ctx.scene, ctx.cam, ctx.st, _np(grad_color), acc=GradAccum(ctx.scene))
        f32 = lambda a: torch.from_numpy(a.astype(np.float32))
        g_feat = f32(acc.sh)
        return (f32(acc.means), torch.from_numpy(acc.mean2d.copy()),
                g_feat if ctx.use_sh else None, None if ctx.use_sh else g_feat,
                f32(acc.opacity)[:, None], f32(acc.cov), None)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            return torch.from_numpy(mark_visible(_np(positions), _np(self.raster_settings.viewmatrix)))

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception("Please provide excatly one of either SHs or precomputed colors!")
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        if cov3D_precomp is None:
            raise NotImplementedError("the oracle covers the cov3D_precomp path the reference uses")
        return _RasterizeGaussians.apply(means3D, means2D, shs, colors_precomp, opacities,
                                         cov3D_precomp, self.raster_settings)


def load_reference_cuda_splatting(rasterizer_module, ref_root: str = "/root/reference"):
    """Import the UNMODIFIED reference ``src/model/decoder/cuda_splatting.py`` with
    ``rasterizer_module`` standing in for ``diff_gaussian_rasterization``.  Parent
    packages are replaced by path-only stubs so their __init__ chains (dacite, lightning, mmcv)
    are skipped.  Works only where /root/reference exists (the authoring container)."""
    saved = {k: sys.modules.get(k) for k in list(sys.modules) if k == "src" or k.startswith("src.")}
    saved["diff_gaussian_rasterization"] = sys.modules.get("diff_gaussian_rasterization")
    for k in list(sys.modules):
        if k == "src" or k.startswith("src."):
            del sys.modules[k]
    try:
        sys.modules["diff_gaussian_rasterization"] = rasterizer_module
        for name in ("src", "src.model", "src.model.decoder", "src.model.encoder",
                     "src.model.encoder.matching", "src.geometry"):
            m = types.ModuleType(name)
            m.__path__ = [f"{ref_root}/{name.replace('.', '/')}"]
            sys.modules[name] = m

        def load(name):
            spec = importlib.util.spec_from_file_location(
                name, f"{ref_root}/{name.replace('.', '/')}.py")
            mod = importlib.util.module_from_spec(spec)
            sys.modules[name] = mod
            spec.loader.exec_module(mod)
            return mod

        load("src.geometry.projection")
        load("src.model.encoder.matching.conversions")
        return load("src.model.decoder.cuda_splatting")
    finally:
        for k in list(sys.modules):
            if k == "src" or k.startswith("src."):
                del sys.modules[k]
        for k, v in saved.items():
            if v is not None:
                sys.modules[k] = v
            else:
                sys.modules.pop(k, None)
