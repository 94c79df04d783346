"""Dense float64 autograd formulation of one rasterizer view.  TEST INFRASTRUCTURE ONLY.

Independent check of the C oracle's hand-written backward: the forward math of
SURVEY A.3/A.5/A.7 is written with differentiable torch ops in binary64 over a dense
[pixels, Gaussians] matrix, and torch.autograd supplies the gradients.  The discrete decisions
(visibility, radius, tile rectangle, depth order) are computed HERE, in binary64, from the same
definitions (``geometry`` below); tests/test_oracle_grad.py compares them with the C oracle's
integer outputs instead of consuming those (they may legitimately differ only where a binary32
quantity sits within rounding distance of a decision boundary, which the test bounds).
Two reference quirks are reproduced on purpose:
  * the alpha cap min(0.99, .) is NOT masked in the reference's gradient -> straight-through;
  * colour clamp max(.,0) masks the gradient (``clamped`` flags) -> plain autograd of clamp.
Small scenes only (memory is pixels x Gaussians).
"""
from __future__ import annotations

import numpy as np
import torch

from . import O_BG, O_CAMPOS, O_CSCALE, O_DROW, O_FAR, O_MSCALE, O_NEAR, O_PROJ, O_TANX, O_TANY, O_VIEW

SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792,
         0.5462742152960396]
SH_C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
         -0.4570457994644658, 1.445305721320277, -0.5900435899266435]
SH_C4 = [2.5033429417967046, -1.7701307697799304, 0.9461746957575601, -0.6690465435572892,
         0.10578554691520431, -0.6690465435572892, 0.47308734787878004, -1.7701307697799304,
         0.6258357354491761]


def sh_basis(deg: int, d: torch.Tensor, sh_l4: bool = True) -> torch.Tensor:
    """[G,3] unit directions -> [G,n] basis values (closed-form real SH polynomials)."""
    x, y, z = d.unbind(-1)
    one = torch.ones_like(x)
    b = [SH_C0 * one]
    if deg > 0:
        b += [-SH_C1 * y, SH_C1 * z, -SH_C1 * x]
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        b += [SH_C2[0] * xy, SH_C2[1] * yz, SH_C2[2] * (2 * zz - xx - yy), SH_C2[3] * xz,
              SH_C2[4] * (xx - yy)]
    if deg > 2:
        b += [SH_C3[0] * y * (3 * xx - yy), SH_C3[1] * xy * z, SH_C3[2] * y * (4 * zz - xx - yy),
              SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy), SH_C3[4] * x * (4 * zz - xx - yy),
              SH_C3[5] * z * (xx - yy), SH_C3[6] * x * (xx - 3 * yy)]
    if deg > 3 and sh_l4:
        b += [SH_C4[0] * xy * (xx - yy), SH_C4[1] * yz * (3 * xx - yy), SH_C4[2] * xy * (7 * zz - 1),
              SH_C4[3] * yz * (7 * zz - 3), SH_C4[4] * (zz * (35 * zz - 30) + 3),
              SH_C4[5] * xz * (7 * zz - 3), SH_C4[6] * (xx - yy) * (7 * zz - 1),
              SH_C4[7] * xz * (xx - 3 * yy), SH_C4[8] * (xx * (xx - 3 * yy) - yy * (3 * xx - yy))]
    return torch.stack(b, -1)


def render_dense(means, cov6, sh, opacity, cam: np.ndarray, H: int, W: int, sh_degree: int,
                 radii: np.ndarray | None = None, depths: np.ndarray | None = None, depth_mode: int = 0,
                 sh_l4: bool = True, geometry_out: dict | None = None):
    """means [G,3], cov6 [G,6], sh [G,M,3], opacity [G] (float64 tensors, may require grad).
    With ``radii``/``depths`` None (the default) visibility, radius, tile rectangle and depth order are this
    function's own binary64 evaluation of SURVEY A.2/A.3; ``geometry_out`` receives them (radii, radius_real,
    visible, depths, order, rect) for comparison with the oracle.  Returns colour [3,H,W] and the un-averaged
    depth channel [H,W] (or None)."""
    f64 = torch.float64
    c = torch.from_numpy(np.asarray(cam, np.float64))
    vm, pj = c[O_VIEW:O_VIEW + 16], c[O_PROJ:O_PROJ + 16]
    campos, tanx, tany, bg = c[O_CAMPOS:O_CAMPOS + 3], c[O_TANX], c[O_TANY], c[O_BG:O_BG + 3]
    s, s2 = c[O_MSCALE], c[O_CSCALE]
    R = vm.reshape(4, 4).T[:3, :3]          # conventional rotation: R[r][c] = vm[4c+r]
    tvec = vm.reshape(4, 4)[3, :3]
    Pm = pj.reshape(4, 4).T                  # conventional full projection
    fx, fy = W / (2 * tanx), H / (2 * tany)

    m = means * s
    S6 = cov6 * s2
    t = m @ R.T + tvec
    ph = torch.cat((m, torch.ones_like(m[:, :1])), -1) @ Pm.T
    pw = 1 / (ph[:, 3] + 1e-7)
    ndc = ph[:, :2] * pw[:, None]
    px = ((ndc[:, 0] + 1) * W - 1) * 0.5
    py = ((ndc[:, 1] + 1) * H - 1) * 0.5

    limx, limy = 1.3 * tanx, 1.3 * tany
    tz = t[:, 2]
    tx = torch.clamp(t[:, 0] / tz, -limx, limx) * tz
    ty = torch.clamp(t[:, 1] / tz, -limy, limy) * tz
    zero = torch.zeros_like(tz)
    J = torch.stack((torch.stack((fx / tz, zero, -(fx * tx) / (tz * tz)), -1),
                     torch.stack((zero, fy / tz, -(fy * ty) / (tz * tz)), -1)), 1)  # [G,2,3]
    T = J @ R
    Sig = torch.stack((S6[:, 0], S6[:, 1], S6[:, 2], S6[:, 1], S6[:, 3], S6[:, 4], S6[:, 2], S6[:, 4],
                       S6[:, 5]), -1).reshape(-1, 3, 3)
    c2 = T @ Sig @ T.transpose(1, 2)
    a, b, cc = c2[:, 0, 0] + 0.3, c2[:, 0, 1], c2[:, 1, 1] + 0.3
    det = a * cc - b * b
    conA, conB, conC = cc / det, -b / det, a / det

    if sh.dim() == 3:
        d = m - campos
        d = d / d.norm(dim=-1, keepdim=True)
        B = sh_basis(sh_degree, d, sh_l4)
        rgb = torch.einsum("gk,gkc->gc", B, sh[:, :B.shape[1], :]) + 0.5
        rgb = torch.clamp(rgb, min=0.0)
    else:
        rgb = sh
    feats = [rgb]
    if depth_mode:
        drow = c[O_DROW:O_DROW + 4]
        z = means @ drow[:3] + drow[3]
        near, far = c[O_NEAR], c[O_FAR]
        if depth_mode == 1:
            f = z
        elif depth_mode == 2:
            f = 1 / z
        elif depth_mode == 3:
            eps = 1e-10
            dn, df = 1 / (near + eps), 1 / (far + eps)
            f = 1 - (1 / (z + eps) - df) / (dn - df + eps)
        else:
            f = torch.minimum(z, near).maximum(far).log()
        feats.append(torch.clamp(SH_C0 * f + 0.5, min=0.0)[:, None])
    feat = torch.cat(feats, -1)  # [G,CH]

    gx, gy = (W + 15) // 16, (H + 15) // 16
    if radii is None:
        # own discrete decisions, binary64: near cull, det == 0, radius = ceil(3 sqrt(lambda_max)) with
        # the eigenvalue floor, getRect with C truncation, zero-area rectangles culled
        a_n, b_n, c_n, det_n = (v.detach().numpy() for v in (a, b, cc, det))
        tz_n, px_n, py_n = tz.detach().numpy(), px.detach().numpy(), py.detach().numpy()
        mid = 0.5 * (a_n + c_n)
        lam = mid + np.sqrt(np.maximum(0.1, mid * mid - det_n))
        radius_real = 3.0 * np.sqrt(np.maximum(lam, 0.0))
        own_r = np.ceil(radius_real)
        front = (tz_n > 0.2) & (det_n != 0.0)
        r0 = np.where(front, own_r, 0.0)
        tr = lambda v, hi: np.clip(np.trunc(v), 0, hi)
        x0, y0 = tr((px_n - r0) / 16.0, gx), tr((py_n - r0) / 16.0, gy)
        x1, y1 = tr((px_n + r0 + 15.0) / 16.0, gx), tr((py_n + r0 + 15.0) / 16.0, gy)
        visible = front & ((x1 - x0) * (y1 - y0) > 0)
        radii = np.where(visible, own_r, 0).astype(np.int32)
        depths = np.where(visible, tz_n, 0.0).astype(np.float32)  # the sort key is the binary32 bit pattern
        if geometry_out is not None:
            geometry_out.update(radii=radii, radius_real=radius_real, visible=visible, depths=depths,
                                rect=np.stack((x0, y0, x1, y1), -1).astype(np.int64), px=px_n, py=py_n, tz=tz_n)
    # depth order (stable: ties by index) over visible Gaussians
    idx = np.nonzero(radii > 0)[0]
    order = idx[np.argsort(depths[idx].view(np.uint32), kind="stable")]
    if geometry_out is not None:
        geometry_out["order"] = order
    order_t = torch.from_numpy(order)
    # tile-rectangle mask (getRect)
    pxf = px.detach().numpy().astype(np.float32)[order]
    pyf = py.detach().numpy().astype(np.float32)[order]
    r = radii[order].astype(np.float32)
    f32 = np.float32
    rx0 = np.clip(((pxf - r) / f32(16)).astype(np.int64), 0, gx)
    ry0 = np.clip(((pyf - r) / f32(16)).astype(np.int64), 0, gy)
    rx1 = np.clip(((((pxf + r) + f32(16)) - f32(1)) / f32(16)).astype(np.int64), 0, gx)
    ry1 = np.clip(((((pyf + r) + f32(16)) - f32(1)) / f32(16)).astype(np.int64), 0, gy)
    ys, xs = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    tpx, tpy = (xs.reshape(-1) // 16)[:, None], (ys.reshape(-1) // 16)[:, None]
    rect = torch.from_numpy((tpx >= rx0) & (tpx < rx1) & (tpy >= ry0) & (tpy < ry1))  # [P,N]

    pixx = torch.from_numpy(xs.reshape(-1).astype(np.float64))[:, None]
    pixy = torch.from_numpy(ys.reshape(-1).astype(np.float64))[:, None]
    dx = px[order_t][None, :] - pixx
    dy = py[order_t][None, :] - pixy
    A, Bc, Cc = conA[order_t][None], conB[order_t][None], conC[order_t][None]
    power = -0.5 * (A * dx * dx + Cc * dy * dy) - Bc * dx * dy
    a_raw = opacity[order_t][None] * torch.exp(power)
    alpha = a_raw + (torch.clamp(a_raw, max=0.99) - a_raw).detach()  # straight-through cap
    keep = rect & (power.detach() <= 0) & (alpha.detach() >= 1.0 / 255.0)
    alpha = torch.where(keep, alpha, torch.zeros_like(alpha))
    one_m = 1 - alpha
    Tincl = torch.cumprod(one_m, dim=1)
    Texcl = torch.cat((torch.ones_like(Tincl[:, :1]), Tincl[:, :-1]), 1)
    # termination: first contributing entry whose test_T < 1e-4 stops the pixel (not composited)
    stop = keep & (Tincl.detach() < 1e-4)
    stopped = torch.cumsum(stop.to(torch.int64), dim=1) > 0
    w = torch.where(stopped, torch.zeros_like(alpha), alpha * Texcl)
    live = torch.where(stopped, torch.ones_like(one_m), one_m)
    Tfinal = torch.prod(live, dim=1)
    out = w @ feat[order_t]  # [P,CH]
    color = (out[:, :3] + Tfinal[:, None] * bg[None, :]).T.reshape(3, H, W)
    depth = out[:, 3].reshape(H, W) if depth_mode else None
    return color, depth
