"""Deterministic synthetic scenes shaped like the TranSplat encoder's output.

Restates the encoder's *output contract*, not the encoder:
  - one Gaussian per context pixel, order (view, row, col)  (encoder_trans.py:336-353)
  - means = origin + dir * depth                            (gaussian_adapter.py:84-85)
  - scales = (0.5 + 14.5 sigmoid) * depth * 0.1 (1/(fx w) + 1/(fy h))
                                                            (gaussian_adapter.py:62-69,98-109)
  - covariance = C2W R diag(s^2) R^T C2W^T                  (gaussians.py:32-43, gaussian_adapter.py:79-81)
  - SH = N(0,1) * mask, mask 1 for l=0 and 0.1*0.25^l above (gaussian_adapter.py:45-46)
  - opacity = sigmoid(N(0, 1.5^2))                          (encoder_trans.py:139-152)
Everything is generated on the CPU with a seeded torch.Generator and float32.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
import torch.nn.functional as F


@dataclass
class SyntheticScene:
    means: torch.Tensor        # [G,3]
    covariances: torch.Tensor  # [G,3,3]
    harmonics: torch.Tensor    # [G,3,d_sh]
    opacities: torch.Tensor    # [G]
    extrinsics: torch.Tensor   # [V,4,4] target cameras, camera-to-world (dataset_re10k.py:230-233)
    intrinsics: torch.Tensor   # [V,3,3] normalised (dataset_re10k.py:221-228)
    near: torch.Tensor         # [V]
    far: torch.Tensor          # [V]
    image_shape: tuple         # (h, w)

    def to(self, device):
        return SyntheticScene(*[t.to(device) if torch.is_tensor(t) else t for t in (
            self.means, self.covariances, self.harmonics, self.opacities, self.extrinsics,
            self.intrinsics, self.near, self.far, self.image_shape)])


def _yaw(angle: float) -> torch.Tensor:
    c, s = math.cos(angle), math.sin(angle)
    return torch.tensor([[c, 0.0, s], [0.0, 1.0, 0.0], [-s, 0.0, c]], dtype=torch.float32)


def _camera(pos_x: float, yaw: float) -> torch.Tensor:
    e = torch.eye(4, dtype=torch.float32)
    e[:3, :3] = _yaw(yaw)
    e[0, 3] = pos_x
    return e


def _quat_to_matrix(q: torch.Tensor, eps: float = 1e-8) -> torch.Tensor:
    i, j, k, r = torch.unbind(q, dim=-1)
    two_s = 2 / ((q * q).sum(dim=-1) + eps)
    o = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(*q.shape[:-1], 3, 3)


def make_scene(n_ctx: int = 2, h: int = 256, w: int = 256, n_views: int = 4, seed: int = 1236,
               sh_degree: int = 4, near: float = 1.0, far: float = 100.0, d0: float = 2.0,
               d1: float = 12.0, fx: float = 0.9, fy: float = 0.9, render_hw=None) -> SyntheticScene:
    """C2 defaults: 2 context views of 256x256 -> G = 131072 Gaussians, 4 target views."""
    g = torch.Generator().manual_seed(seed)
    d_sh = (sh_degree + 1) ** 2
    K = torch.tensor([[fx, 0, 0.5], [0, fy, 0.5], [0, 0, 1]], dtype=torch.float32)
    Kinv = K.inverse()
    ctx = [_camera(0.3 * k, math.radians(3.0) * (k / max(n_ctx - 1, 1))) for k in range(n_ctx)]

    means, covs, shs, ops = [], [], [], []
    mask = torch.ones(d_sh)
    for l in range(1, sh_degree + 1):
        mask[l * l:(l + 1) ** 2] = 0.1 * 0.25 ** l
    ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    for c2w in ctx:
        n = h * w
        off = (torch.sigmoid(torch.randn(n, 2, generator=g)) - 0.5) * torch.tensor([1.0 / w, 1.0 / h])
        uv = torch.stack(((xs.reshape(-1) + 0.5) / w, (ys.reshape(-1) + 0.5) / h), -1) + off
        coarse = torch.rand(1, 1, 8, 8, generator=g) * (math.log(d1) - math.log(d0)) + math.log(d0)
        depth = F.interpolate(coarse, size=(h, w), mode="bicubic", align_corners=False).exp()
        depth = depth.reshape(-1).clamp(near, far)
        dirs = torch.cat((uv, torch.ones(n, 1)), -1) @ Kinv.T
        dirs = dirs / dirs.norm(dim=-1, keepdim=True)
        dirs_w = dirs @ c2w[:3, :3].T
        means.append(c2w[:3, 3] + dirs_w * depth[:, None])
        mult = 0.1 * (1.0 / (fx * w) + 1.0 / (fy * h))
        scales = (0.5 + 14.5 * torch.sigmoid(torch.randn(n, 3, generator=g))) * depth[:, None] * mult
        q = torch.randn(n, 4, generator=g)
        q = q / (q.norm(dim=-1, keepdim=True) + 1e-8)
        R = _quat_to_matrix(q)
        S = torch.diag_embed(scales)
        cov = R @ S @ S.transpose(-1, -2) @ R.transpose(-1, -2)
        Rc = c2w[:3, :3]
        covs.append(Rc @ cov @ Rc.T)
        shs.append(torch.randn(n, 3, d_sh, generator=g) * mask)
        ops.append(torch.sigmoid(torch.randn(n, generator=g) * 1.5))

    ext, t_all = [], [(i + 1) / (n_views + 1) for i in range(n_views)]
    last = 0.3 * (n_ctx - 1)
    for t in t_all:
        ext.append(_camera(last * t, math.radians(3.0) * t))
    hw = render_hw or (h, w)
    return SyntheticScene(
        means=torch.cat(means).contiguous(), covariances=torch.cat(covs).contiguous(),
        harmonics=torch.cat(shs).contiguous(), opacities=torch.cat(ops).contiguous(),
        extrinsics=torch.stack(ext), intrinsics=K.expand(n_views, 3, 3).contiguous(),
        near=torch.full((n_views,), near), far=torch.full((n_views,), far), image_shape=hw)


def upstream_grads(n_views: int, h: int, w: int, seed: int = 99):
    """dL/dcolor for an MSE loss against a U(0,1) target is produced by the caller from the
    rendered colour; this returns the random target and dL/ddepth = N(0,1)/numel."""
    g = torch.Generator().manual_seed(seed)
    target = torch.rand(n_views, 3, h, w, generator=g)
    dd = torch.randn(n_views, h, w, generator=g) / (n_views * h * w)
    return target, dd
