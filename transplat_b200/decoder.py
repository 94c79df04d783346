"""Level-2 boundary: the decoder of TranSplat's render path with the reference's names,
signatures and error behaviour

  Gaussians                         src/model/types.py:7-12
  DecoderOutput, DepthRenderingMode src/model/decoder/decoder.py:12-22
  render_cuda                       src/model/decoder/cuda_splatting.py:56-136
  render_depth_cuda                 src/model/decoder/cuda_splatting.py:375-417
  render_cuda_orthographic          src/model/decoder/cuda_splatting.py:139-229
  render_cuda_with_stats            src/model/decoder/cuda_splatting.py:235-372
  DecoderSplattingCUDA              src/model/decoder/decoder_splatting_cuda.py:26-146
  DECODERS / get_decoder            src/model/decoder/__init__.py:5-13

but batched: one library call renders every (scene, view) of the batch, colour and depth come
from ONE pass, the Gaussians are never ``repeat``-ed per view (decoder_splatting_cuda.py:59-62
materialises v copies), SH is read in the encoder's [G,3,d_sh] layout (no transpose copy,
cuda_splatting.py:84), covariances in their [G,3,3] layout (no triu gather, :124), and nothing
synchronises the host (the reference calls .item() twice per view, :111-112).
"""
from __future__ import annotations

from dataclasses import dataclass
from math import isqrt
from typing import Literal, Optional

import torch
from torch import Tensor, nn

from ._lib import COV_FULL9, DEPTH_MODES, SH_G3M
from .cameras import _inverse, get_projection_matrix, pack_cameras_device, pack_cameras_explicit
from .rasterize import RenderConfig, rasterize, rasterize_mse

DepthRenderingMode = Literal["depth", "log", "disparity", "relative_disparity"]


@dataclass
class Gaussians:
    means: Tensor        # [batch, gaussian, 3]
    covariances: Tensor  # [batch, gaussian, 3, 3]
    harmonics: Tensor    # [batch, gaussian, 3, d_sh]
    opacities: Tensor    # [batch, gaussian]


@dataclass
class DecoderOutput:
    color: Tensor            # [batch, view, 3, height, width]
    depth: Optional[Tensor]  # [batch, view, height, width]


@dataclass
class GaussianContributionStats:
    """cuda_splatting.py:235-254 (same field names)."""
    total_gaussians: int
    visible_gaussians: int            # radii > 0
    high_contribution_gaussians: int  # visible and opacity >= threshold
    visible_ratio: float
    high_contribution_ratio: float
    opacity_threshold: float
    radii: Optional[Tensor] = None


def render_views(gaussians: Gaussians, extrinsics: Tensor, intrinsics: Tensor, near: Tensor, far: Tensor,
                 image_shape: tuple, background_color: Tensor, depth_mode: Optional[str] = None,
                 scale_invariant: bool = True, want_radii: bool = False, check: Optional[str] = None,
                 use_sh: bool = True, cams: Optional[Tensor] = None, mse_target: Optional[Tensor] = None,
                 mse_weight: float = 1.0, with_psnr: bool = False, grad_reduce=None):
    """The batched entry point.  With ``mse_target`` [b,v,3,h,w] a fourth value is returned: the reference's MSE colour
    loss ``mse_weight * ((color - target) ** 2).mean()`` (loss_mse.py:23-31), computed inside the render kernels.
  extrinsics [b,v,4,4], intrinsics [b,v,3,3], near/far [b,v],
    background_color [3] or [b*v,3].  Returns colour [b,v,3,h,w], depth [b,v,h,w]|None, radii [b,v,G]|None."""
    b, v = extrinsics.shape[:2]
    h, w = image_shape
    if depth_mode is not None and depth_mode not in DEPTH_MODES:
        raise ValueError(f"unknown depth_mode {depth_mode!r}")
    d_sh = gaussians.harmonics.shape[-1]
    degree = isqrt(d_sh) - 1
    assert use_sh or d_sh == 1  # cuda_splatting.py:70
    if cams is None:
        cams = pack_cameras_device(extrinsics.reshape(b * v, 4, 4), intrinsics.reshape(b * v, 3, 3),
                                   near.reshape(b * v), far.reshape(b * v), background_color,
                                   scale_invariant=scale_invariant)
    cfg = RenderConfig(H=h, W=w, V=v, sh_degree=degree, sh_layout=SH_G3M, cov_layout=COV_FULL9,
                       depth_mode=DEPTH_MODES[depth_mode], want_radii=want_radii, check=check, grad_reduce=grad_reduce)
    # use_sh=False: the single coefficient is the colour itself (colors_precomp = shs[:, 0, :], cuda_splatting.py:130)
    feat = gaussians.harmonics if use_sh else gaussians.harmonics[..., 0]
    loss = psnr = None
    if mse_target is not None:
        out = rasterize_mse(gaussians.means, gaussians.covariances, feat, gaussians.opacities, cams, cfg,
                            mse_target.reshape(b * v, 3, h, w), mse_weight, with_psnr=with_psnr)
        color, depth, radii, loss = out[:4]
        psnr = out[4] if with_psnr else None  # [b*v], the batch axis of compute_psnr's "(b v) c h w" input
    else:
        color, depth, radii = rasterize(gaussians.means, gaussians.covariances, feat, gaussians.opacities, cams, cfg)
    color = color.reshape(b, v, 3, h, w)
    depth = depth.reshape(b, v, h, w) if depth is not None else None
    radii = radii.reshape(b, v, -1) if radii is not None else None
    if mse_target is not None:
        return (color, depth, radii, loss, psnr) if with_psnr else (color, depth, radii, loss)
    return color, depth, radii


def render_cuda(extrinsics, intrinsics, near, far, image_shape, background_color, gaussian_means,
                gaussian_covariances, gaussian_sh_coefficients, gaussian_opacities,
                scale_invariant: bool = True, use_sh: bool = True) -> Tensor:
    """Signature of cuda_splatting.py:56-69: every argument has a leading "batch" axis that is
    (scene x view) flattened, Gaussians already repeated per view.  Each batch entry is rendered as
    its own one-view scene, in one library call."""
    assert use_sh or gaussian_sh_coefficients.shape[-1] == 1
    b = extrinsics.shape[0]
    g = Gaussians(gaussian_means, gaussian_covariances, gaussian_sh_coefficients, gaussian_opacities)
    color, _, _ = render_views(g, extrinsics[:, None], intrinsics[:, None], near[:, None], far[:, None],
                               image_shape, background_color, None, scale_invariant, use_sh=use_sh)
    return color.reshape(b, 3, *image_shape)


def render_cuda_orthographic(extrinsics, width, height, near, far, image_shape, background_color, gaussian_means,
                             gaussian_covariances, gaussian_sh_coefficients, gaussian_opacities,
                             fov_degrees: float = 0.1, use_sh: bool = True, dump: Optional[dict] = None) -> Tensor:
    """Signature and camera arithmetic of cuda_splatting.py:139-229: a "fake orthographic" camera pulled back
    until ``fov_degrees`` spans ``width``.  The reference writes a [batch] distance into ONE 4x4 ``move_back``
    (:169-170), which only works for batch 1; here each batch entry gets its own."""
    b = extrinsics.shape[0]
    assert use_sh or gaussian_sh_coefficients.shape[-1] == 1
    dev = extrinsics.device
    fov_x = torch.tensor(fov_degrees, device=dev).deg2rad()
    tan_fov_x = (0.5 * fov_x).tan()
    distance_to_near = (0.5 * width) / tan_fov_x
    tan_fov_y = 0.5 * height / distance_to_near
    fov_y = (2 * tan_fov_y).atan()  # sic (:166)
    near = near + distance_to_near
    far = far + distance_to_near
    move_back = torch.eye(4, dtype=torch.float32, device=dev).repeat(b, 1, 1)
    move_back[:, 2, 3] = -distance_to_near
    extrinsics = extrinsics @ move_back
    if dump is not None:
        dump["extrinsics"] = extrinsics
        dump["fov_x"] = fov_x
        dump["fov_y"] = fov_y
        dump["near"] = near
        dump["far"] = far
    projection_matrix = get_projection_matrix(near, far, fov_x.expand(b), fov_y).transpose(1, 2)
    view_matrix = _inverse(extrinsics).transpose(1, 2)
    full_projection = view_matrix @ projection_matrix
    cams = pack_cameras_explicit(view_matrix, full_projection, extrinsics[:, :3, 3], tan_fov_x.expand(b), tan_fov_y,
                                 background_color)
    g = Gaussians(gaussian_means, gaussian_covariances, gaussian_sh_coefficients, gaussian_opacities)
    color, _, _ = render_views(g, extrinsics[:, None], None, near[:, None], far[:, None], image_shape,
                               background_color, None, use_sh=use_sh, cams=cams)
    return color.reshape(b, 3, *image_shape)


def _contribution_stats(radii: Tensor, opacities: Tensor, opacity_threshold: float) -> list:
    """cuda_splatting.py:353-369 for every (scene, view): radii [b,v,G], opacities [b,G]; one host read."""
    b, v, G = radii.shape
    visible = radii > 0
    high = visible & (opacities >= opacity_threshold)[:, None, :]
    counts = torch.stack((visible.sum(dim=-1), high.sum(dim=-1)), dim=-1).cpu()
    stats = []
    for i in range(b):
        for j in range(v):
            nv, nh = int(counts[i, j, 0]), int(counts[i, j, 1])
            stats.append(GaussianContributionStats(
                total_gaussians=G, visible_gaussians=nv, high_contribution_gaussians=nh,
                visible_ratio=nv / G if G > 0 else 0.0, high_contribution_ratio=nh / G if G > 0 else 0.0,
                opacity_threshold=opacity_threshold, radii=radii[i, j].clone()))
    return stats


def render_cuda_with_stats(extrinsics, intrinsics, near, far, image_shape, background_color, gaussian_means,
                           gaussian_covariances, gaussian_sh_coefficients, gaussian_opacities,
                           scale_invariant: bool = True, use_sh: bool = True, opacity_threshold: float = 0.01):
    """Signature of cuda_splatting.py:256-372: images [batch,3,h,w] + one GaussianContributionStats per batch
    entry (visible = radii > 0; high contribution = visible and opacity >= threshold)."""
    b = extrinsics.shape[0]
    g = Gaussians(gaussian_means, gaussian_covariances, gaussian_sh_coefficients, gaussian_opacities)
    color, _, radii = render_views(g, extrinsics[:, None], intrinsics[:, None], near[:, None], far[:, None],
                                   image_shape, background_color, None, scale_invariant, want_radii=True,
                                   use_sh=use_sh)
    return color.reshape(b, 3, *image_shape), _contribution_stats(radii, gaussian_opacities, opacity_threshold)


def render_depth_cuda(extrinsics, intrinsics, near, far, image_shape, gaussian_means, gaussian_covariances,
                      gaussian_opacities, scale_invariant: bool = True, mode: DepthRenderingMode = "depth") -> Tensor:
    """Signature of cuda_splatting.py:375-386.  The depth channel is composited by the same kernel
    as colour; the colour written alongside it here comes from zero harmonics and is discarded."""
    b, G = gaussian_means.shape[:2]
    zeros = torch.zeros((b, G, 3, 1), dtype=torch.float32, device=gaussian_means.device)
    g = Gaussians(gaussian_means, gaussian_covariances, zeros, gaussian_opacities)
    bg = torch.zeros((b, 3), dtype=torch.float32, device=gaussian_means.device)
    _, depth, _ = render_views(g, extrinsics[:, None], intrinsics[:, None], near[:, None], far[:, None],
                               image_shape, bg, mode, scale_invariant)
    return depth.reshape(b, *image_shape)


class DecoderSplattingCUDA(nn.Module):
    """decoder_splatting_cuda.py:26-146 — same constructor shape (cfg, dataset_cfg) where
    ``dataset_cfg.background_color`` is read, same forward/render_depth/forward_with_gaussian_stats."""

    def __init__(self, cfg=None, dataset_cfg=None, background_color=(0.0, 0.0, 0.0)) -> None:
        super().__init__()
        self.cfg = cfg
        self.dataset_cfg = dataset_cfg
        if dataset_cfg is not None and hasattr(dataset_cfg, "background_color"):
            background_color = dataset_cfg.background_color
        self.register_buffer("background_color", torch.tensor(background_color, dtype=torch.float32),
                             persistent=False)

    def forward(self, gaussians: Gaussians, extrinsics: Tensor, intrinsics: Tensor, near: Tensor, far: Tensor,
                image_shape: tuple, depth_mode: Optional[DepthRenderingMode] = None, grad_reduce=None) -> DecoderOutput:
        """decoder_splatting_cuda.py:41-73.  ``grad_reduce`` = (process group, chunks) is this library's addition for
        view sharding (parallel.forward_view_sharded): the backward sums the gradients over the group itself."""
        color, depth, _ = render_views(gaussians, extrinsics, intrinsics, near, far, image_shape,
                                       self.background_color, depth_mode, grad_reduce=grad_reduce)
        return DecoderOutput(color, depth)

    def forward_mse(self, gaussians: Gaussians, extrinsics: Tensor, intrinsics: Tensor, near: Tensor, far: Tensor,
                    image_shape: tuple, target_image: Tensor, depth_mode: Optional[DepthRenderingMode] = None,
                    weight: float = 1.0, with_psnr: bool = False):
        """forward() + LossMse (src/loss/loss_mse.py:23-31: ``weight * ((prediction.color - batch["target"]["image"])
        ** 2).mean()``) fused into the render kernels: returns (DecoderOutput, loss).  Backpropagating the loss never
        materialises the colour difference or its gradient.  ``with_psnr`` adds a third value: the
        per-view ``compute_psnr`` [b*v] the training step logs as train/psnr_probabilistic (model_wrapper.py:149-153,
        src/evaluation/metrics.py:12-19), reduced in the same epilogue."""
        out = render_views(gaussians, extrinsics, intrinsics, near, far, image_shape, self.background_color,
                           depth_mode, mse_target=target_image, mse_weight=weight, with_psnr=with_psnr)
        if with_psnr:
            return DecoderOutput(out[0], out[1]), out[3], out[4]
        return DecoderOutput(out[0], out[1]), out[3]

    def render_depth(self, gaussians: Gaussians, extrinsics: Tensor, intrinsics: Tensor, near: Tensor, far: Tensor,
                     image_shape: tuple, mode: DepthRenderingMode = "depth") -> Tensor:
        _, depth, _ = render_views(gaussians, extrinsics, intrinsics, near, far, image_shape,
                                   torch.zeros_like(self.background_color), mode)
        return depth

    def forward_with_gaussian_stats(self, gaussians: Gaussians, extrinsics, intrinsics, near, far, image_shape,
                                    depth_mode: Optional[DepthRenderingMode] = None,
                                    opacity_threshold: float = 0.01):
        """decoder_splatting_cuda.py:99-146 / cuda_splatting.py:353-369: per-view contribution
        statistics from ``radii`` (the only consumer of radii in the reference)."""
        b, v = extrinsics.shape[:2]
        color, depth, radii = render_views(gaussians, extrinsics, intrinsics, near, far, image_shape,
                                           self.background_color, depth_mode, want_radii=True)
        return DecoderOutput(color, depth), _contribution_stats(radii, gaussians.opacities, opacity_threshold)


@dataclass
class DecoderSplattingCUDACfg:
    name: Literal["splatting_cuda"]


DECODERS = {"splatting_cuda": DecoderSplattingCUDA}


def get_decoder(decoder_cfg, dataset_cfg):
    return DECODERS[decoder_cfg.name](decoder_cfg, dataset_cfg)
