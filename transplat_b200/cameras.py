"""Host-side camera preparation for the render path: the torch arithmetic of
/root/reference/src/model/decoder/cuda_splatting.py:73-96 restated op for op (so the matrices are
bit-identical to what the reference hands the rasterizer), but emitted as one packed
``[N, 64]`` float32 tensor of ``TsplatCamera`` records (include/tsplat.h) that stays on the
device: no ``.item()`` (cuda_splatting.py:111-112 syncs twice per view), no per-view Python loop.
"""
from __future__ import annotations

import torch
from torch import Tensor

CAM_FLOATS = 64
O_VIEW, O_PROJ, O_CAMPOS, O_TANX, O_TANY, O_BG, O_MSCALE, O_CSCALE, O_DROW, O_NEAR, O_FAR = (
    0, 16, 32, 35, 36, 37, 40, 41, 42, 46, 47)


def _inverse(m: Tensor) -> Tensor:
    """Same LU kernels and values as Tensor.inverse() (what the reference calls,
    cuda_splatting.py:95, projection.py:234) without its error check, which reads a flag back to
    the host and so synchronises the device on every call."""
    return torch.linalg.inv_ex(m)[0]


def get_fov(intrinsics: Tensor) -> Tensor:
    """src/geometry/projection.py:233-247 (angle between the rays through the edge midpoints)."""
    intrinsics_inv = _inverse(intrinsics)

    def process_vector(vector):
        vector = torch.tensor(vector, dtype=torch.float32, device=intrinsics.device)
        vector = torch.einsum("bij,j->bi", intrinsics_inv, vector)
        return vector / vector.norm(dim=-1, keepdim=True)

    left = process_vector([0, 0.5, 1])
    right = process_vector([1, 0.5, 1])
    top = process_vector([0.5, 0, 1])
    bottom = process_vector([0.5, 1, 1])
    fov_x = (left * right).sum(dim=-1).acos()
    fov_y = (top * bottom).sum(dim=-1).acos()
    return torch.stack((fov_x, fov_y), dim=-1)


def get_projection_matrix(near: Tensor, far: Tensor, fov_x: Tensor, fov_y: Tensor) -> Tensor:
    """cuda_splatting.py:26-53: symmetric frustum, z in (0,1), P[3,2] = 1."""
    tan_fov_x = (0.5 * fov_x).tan()
    tan_fov_y = (0.5 * fov_y).tan()
    top = tan_fov_y * near
    bottom = -top
    right = tan_fov_x * near
    left = -right
    (b,) = near.shape
    result = torch.zeros((b, 4, 4), dtype=torch.float32, device=near.device)
    result[:, 0, 0] = 2 * near / (right - left)
    result[:, 1, 1] = 2 * near / (top - bottom)
    result[:, 0, 2] = (right + left) / (right - left)
    result[:, 1, 2] = (top + bottom) / (top - bottom)
    result[:, 3, 2] = 1
    result[:, 2, 2] = far / (far - near)
    result[:, 2, 3] = -(far * near) / (far - near)
    return result


def pack_cameras(extrinsics: Tensor, intrinsics: Tensor, near: Tensor, far: Tensor,
                 background_color: Tensor, scale_invariant: bool = True) -> Tensor:
    """[N,4,4] c2w extrinsics, [N,3,3] normalised intrinsics, [N] near/far, [N,3] bg -> [N,64].

    Follows render_cuda (cuda_splatting.py:73-96,108-118) and the depth-feature row of
    render_depth_cuda (cuda_splatting.py:388-391)."""
    n = extrinsics.shape[0]
    dev = extrinsics.device
    inv_unscaled = _inverse(extrinsics)  # render_depth_cuda uses the unscaled inverse (:389)
    near_u, far_u = near, far
    if scale_invariant:
        scale = 1 / near
        extrinsics = extrinsics.clone()
        extrinsics[..., :3, 3] = extrinsics[..., :3, 3] * scale[:, None]
        near = near * scale
        far = far * scale
    else:
        scale = torch.ones_like(near)
    fov_x, fov_y = get_fov(intrinsics).unbind(dim=-1)
    tan_fov_x = (0.5 * fov_x).tan()
    tan_fov_y = (0.5 * fov_y).tan()
    projection_matrix = get_projection_matrix(near, far, fov_x, fov_y).transpose(1, 2)
    view_matrix = _inverse(extrinsics).transpose(1, 2)
    full_projection = view_matrix @ projection_matrix

    cams = torch.zeros((n, CAM_FLOATS), dtype=torch.float32, device=dev)
    cams[:, O_VIEW:O_VIEW + 16] = view_matrix.reshape(n, 16)
    cams[:, O_PROJ:O_PROJ + 16] = full_projection.reshape(n, 16)
    cams[:, O_CAMPOS:O_CAMPOS + 3] = extrinsics[:, :3, 3]
    cams[:, O_TANX] = tan_fov_x
    cams[:, O_TANY] = tan_fov_y
    cams[:, O_BG:O_BG + 3] = background_color
    cams[:, O_MSCALE] = scale
    cams[:, O_CSCALE] = scale ** 2
    cams[:, O_DROW:O_DROW + 4] = inv_unscaled[:, 2, :]
    cams[:, O_NEAR] = near_u
    cams[:, O_FAR] = far_u
    return cams


def pack_cameras_device(extrinsics: Tensor, intrinsics: Tensor, near: Tensor, far: Tensor,
                        background_color: Tensor, scale_invariant: bool = True) -> Tensor:
    """Same contract as :func:`pack_cameras`, computed by ONE kernel of libtsplat.so
    (tsplat_pack_cameras, csrc/cameras.cu) on the current stream: no torch op chain (the torch
    restatement above is ~100 tiny launches per call), no host synchronisation.
    ``background_color`` is [3] (shared) or [N,3]."""
    import ctypes as C

    from . import _lib

    if not extrinsics.is_cuda:
        raise RuntimeError("transplat_b200 renders on a CUDA device only (no CPU fallback): cameras must be CUDA tensors")
    n = extrinsics.shape[0]
    dev = extrinsics.device
    f32 = lambda t: t.to(dtype=torch.float32).contiguous()
    extrinsics, intrinsics, near, far, bg = (f32(t) for t in (extrinsics, intrinsics, near, far, background_color))
    if tuple(extrinsics.shape) != (n, 4, 4) or tuple(intrinsics.shape) != (n, 3, 3) or near.numel() != n or far.numel() != n:
        raise RuntimeError("pack_cameras_device: extrinsics [N,4,4], intrinsics [N,3,3], near/far [N] expected")
    if bg.numel() == 3:
        stride = 0
    elif tuple(bg.shape) == (n, 3):
        stride = 3
    else:
        raise RuntimeError("background_color must be [3] or [N,3]")
    cams = torch.empty((n, CAM_FLOATS), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(_lib.lib().tsplat_pack_cameras(n, extrinsics.data_ptr(), intrinsics.data_ptr(), near.data_ptr(),
                                                  far.data_ptr(), bg.data_ptr(), stride, 1 if scale_invariant else 0,
                                                  cams.data_ptr(), C.c_void_p(stream)), "tsplat_pack_cameras")
    return cams


def pack_camera_from_settings(viewmatrix: Tensor, projmatrix: Tensor, campos: Tensor, tanfovx,
                              tanfovy, bg: Tensor) -> Tensor:
    """Level 1: one record from the fields of GaussianRasterizationSettings
    (cuda_splatting.py:108-121).  tanfov may be a float or a 0-d/1-element tensor
    (cuda_splatting.py:204-205 passes tensors)."""
    dev = viewmatrix.device
    cam = torch.zeros((1, CAM_FLOATS), dtype=torch.float32, device=dev)
    cam[0, O_VIEW:O_VIEW + 16] = viewmatrix.reshape(16)
    cam[0, O_PROJ:O_PROJ + 16] = projmatrix.reshape(16)
    cam[0, O_CAMPOS:O_CAMPOS + 3] = campos.reshape(3)
    tail = torch.zeros(CAM_FLOATS - O_TANX, dtype=torch.float32)
    tail[0] = float(tanfovx)
    tail[1] = float(tanfovy)
    tail[O_MSCALE - O_TANX] = 1.0
    tail[O_CSCALE - O_TANX] = 1.0
    tail[O_DROW - O_TANX + 2] = 1.0
    tail[O_NEAR - O_TANX] = 1.0
    tail[O_FAR - O_TANX] = 100.0
    cam[0, O_TANX:] = tail.to(dev, non_blocking=True)
    cam[0, O_BG:O_BG + 3] = bg.reshape(3)
    return cam


def pack_cameras_explicit(view_matrix: Tensor, full_projection: Tensor, campos: Tensor, tan_fov_x: Tensor,
                          tan_fov_y: Tensor, background_color: Tensor) -> Tensor:
    """[N] records from explicit matrices in the reference's transposed convention (what it passes as
    ``viewmatrix`` / ``projmatrix``), no scale invariance: the camera set-up of render_cuda_orthographic
    (cuda_splatting.py:165-214).  Everything stays on the device (tan_fov as tensors, no ``.item()``)."""
    n = view_matrix.shape[0]
    dev = view_matrix.device
    cams = torch.zeros((n, CAM_FLOATS), dtype=torch.float32, device=dev)
    cams[:, O_VIEW:O_VIEW + 16] = view_matrix.reshape(n, 16)
    cams[:, O_PROJ:O_PROJ + 16] = full_projection.reshape(n, 16)
    cams[:, O_CAMPOS:O_CAMPOS + 3] = campos.reshape(n, 3)
    cams[:, O_TANX] = tan_fov_x
    cams[:, O_TANY] = tan_fov_y
    cams[:, O_BG:O_BG + 3] = background_color
    cams[:, O_MSCALE] = 1.0
    cams[:, O_CSCALE] = 1.0
    cams[:, O_DROW + 2] = 1.0
    cams[:, O_NEAR] = 1.0
    cams[:, O_FAR] = 100.0
    return cams
