"""Shared helpers of the parity tests: run the CPU oracle and the CUDA library on the same inputs
and compare stage by stage.  The oracle is the checker; the CUDA path is the thing checked."""
from __future__ import annotations

import numpy as np
import torch

import oracle
from transplat_b200.rasterize import RenderConfig, backward_raw, forward_raw, inspect_workspace

COLOR_ATOL = 1e-5   # north_star: rendered colour/depth within 1e-5 abs
GRAD_RTOL = 1e-4    # north_star: Gaussian gradients within 1e-4 relative (to the tensor's max-abs)
# Pixels on which some alpha sits within rounding distance of 1/255 (or T(1-alpha) of 1e-4, or the exponent of 0) are
# identified by the oracle (oracle.borderline_pixels): there the device's ex2-based exp and glibc's expf may take
# different branches - one contribution of alpha ~ 1/255 more or less.  Those pixels (a few per million), and the
# Gaussians of their tile lists in the gradient comparison, are held to the explicit looser bounds below; EVERYTHING
# else is held to the north_star tolerances, and n_contrib must match exactly outside them.
BORDERLINE_FRAC = 1e-6          # borderline pixels per (pair, pixel) evaluation (256 x num_rendered): the oracle flags ~3e-7
BORDERLINE_COLOR_ATOL = 2e-2    # one alpha = 1/255 contribution of a colour in [0, ~4]
BORDERLINE_GRAD_RTOL = 5e-2     # of the tensor's max-abs, on the Gaussians of the borderline pixels' tile lists


def oracle_batch(means, cov, sh, opacity, cams, cfg: RenderConfig, dL_dcolor=None, dL_ddepth=None):
    """means [B,G,3] ... numpy/torch CPU; cams [B*V,64].  Returns (list of ViewState, list of GradAccum|None)."""
    to_np = lambda t: t.detach().cpu().numpy() if torch.is_tensor(t) else np.asarray(t)
    means, cov, sh, opacity, cams = map(to_np, (means, cov, sh, opacity, cams))
    B = means.shape[0]
    states, accs = [], []
    for b in range(B):
        scene = oracle.Scene(means[b], cov[b], sh[b], opacity[b], cfg.sh_degree, cfg.sh_layout,
                             cfg.cov_layout, cfg.sh_l4)
        acc = oracle.GradAccum(scene) if dL_dcolor is not None else None
        for v in range(cfg.V):
            n = b * cfg.V + v
            st = oracle.forward(scene, cams[n], cfg.H, cfg.W, depth_mode=cfg.depth_mode)
            states.append(st)
            if acc is not None:
                oracle.backward(scene, cams[n], st, to_np(dL_dcolor)[n],
                                to_np(dL_ddepth)[n] if cfg.depth_mode else None, acc=acc,
                                depth_mode=cfg.depth_mode)
                st.extra["mean2d"] = acc.mean2d
                st.extra["screen"] = acc.screen
        accs.append(acc)
    return states, accs


def compare_forward(ins: dict, color, depth, states, cfg: RenderConfig) -> dict:
    """Bit-exact checks on integer/bit quantities, tolerance checks on colour/depth."""
    T = ins["T"]
    rep = {"views": len(states)}
    cpu = {k: (v.cpu().numpy() if torch.is_tensor(v) else v) for k, v in ins.items()}
    color, depth = color.cpu().numpy(), (depth.cpu().numpy() if depth is not None else None)
    D_total, off = 0, 0
    rep.update(radii_mismatch=0, depth_bits_mismatch=0, tiles_touched_mismatch=0, xy_bits_mismatch=0,
               clamped_mismatch=0, keys_mismatch=0, vals_mismatch=0, ranges_mismatch=0, n_contrib_mismatch=0,
               conic_max_abs=0.0, rgbd_max_abs=0.0, color_max_abs=0.0, depth_max_abs=0.0, final_T_max_abs=0.0,
               borderline_pixels=0, borderline_n_contrib_mismatch=0, borderline_color_max_abs=0.0,
               borderline_depth_max_abs=0.0)
    for n, st in enumerate(states):
        rep["radii_mismatch"] += int((cpu["radii"][n] != st.radii).sum())
        rep["depth_bits_mismatch"] += int((cpu["depths"][n].view(np.uint32) != st.depths.view(np.uint32)).sum())
        rep["tiles_touched_mismatch"] += int((cpu["tiles_touched"][n].astype(np.uint32) != st.tiles_touched).sum())
        rep["xy_bits_mismatch"] += int((cpu["xy"][n].view(np.uint32) != st.xy.view(np.uint32)).sum())
        rep["clamped_mismatch"] += int((cpu["clamped"][n] != st.clamped).sum())
        rep["conic_max_abs"] = max(rep["conic_max_abs"], float(np.abs(cpu["conic_opacity"][n] - st.conic_opacity).max(initial=0)))
        rep["rgbd_max_abs"] = max(rep["rgbd_max_abs"], float(np.abs(cpu["rgbd"][n] - st.rgbd).max(initial=0)))
        D = st.num_rendered
        gk = cpu["keys"][off:off + D].view(np.uint64)
        gv = cpu["vals"][off:off + D].view(np.uint32)
        if gk.shape[0] != D:
            rep["keys_mismatch"] += D
        else:
            ok = st.keys + (np.uint64(n * T) << np.uint64(32))
            rep["keys_mismatch"] += int((gk != ok).sum())
            rep["vals_mismatch"] += int((gv != st.vals).sum())
        r = cpu["ranges"][n].astype(np.int64)
        o = st.ranges.astype(np.int64).copy()
        nonempty = o[:, 1] > o[:, 0]
        o[nonempty] += off
        rep["ranges_mismatch"] += int((r[nonempty] != o[nonempty]).sum()) + int(
            ((r[~nonempty, 1] - r[~nonempty, 0]) != 0).sum())
        soft = oracle.borderline_pixels(st)
        st.extra["borderline"] = soft
        firm = ~soft
        mx = lambda a, m: float(a[m].max(initial=0.0))
        nc_bad = cpu["n_contrib"][n].astype(np.uint32) != st.n_contrib
        rep["borderline_pixels"] += int(soft.sum())
        rep["n_contrib_mismatch"] += int((nc_bad & firm).sum())
        rep["borderline_n_contrib_mismatch"] += int((nc_bad & soft).sum())
        rep["final_T_max_abs"] = max(rep["final_T_max_abs"], mx(np.abs(cpu["final_T"][n] - st.final_T), firm))
        dcol = np.abs(color[n] - st.color).max(0)
        rep["color_max_abs"] = max(rep["color_max_abs"], mx(dcol, firm))
        rep["borderline_color_max_abs"] = max(rep["borderline_color_max_abs"], mx(dcol, soft))
        if depth is not None:
            ddep = np.abs(depth[n] - st.depth)
            rep["depth_max_abs"] = max(rep["depth_max_abs"], mx(ddep, firm))
            rep["borderline_depth_max_abs"] = max(rep["borderline_depth_max_abs"], mx(ddep, soft))
        off += D
        D_total += D
    rep["pixels"] = int(color.shape[0] * color.shape[-2] * color.shape[-1])
    rep["evaluations"] = 256 * D_total
    rep["num_rendered_gpu"] = cpu["num_rendered"]
    rep["num_rendered_oracle"] = D_total
    return rep


def rel_err(a: np.ndarray, ref: np.ndarray) -> float:
    m = float(np.abs(ref).max())
    return float(np.abs(a - ref).max()) / m if m > 0 else float(np.abs(a).max())


def soft_gaussians(states, B: int, V: int) -> np.ndarray:
    """bool [B,G]: Gaussians in the tile list of a borderline pixel of any view of their scene."""
    G = states[0].G
    out = np.zeros((B, G), bool)
    for n, st in enumerate(states):
        soft = st.extra.get("borderline")
        if soft is not None and soft.any():
            out[n // V] |= oracle.gaussians_of_pixels(st, soft)
    return out


def compare_backward(grads, accs, soft=None) -> dict:
    """x_rel: max error over the Gaussians NOT touched by a borderline pixel, relative to the tensor's max-abs;
    x_rel_borderline: the same over the touched ones (held to BORDERLINE_GRAD_RTOL)."""
    g_means, g_cov, g_sh, g_op = (g.cpu().numpy().astype(np.float64) for g in grads[:4])
    rep = {}
    for name, g, pick in (("means", g_means, lambda a: a.means), ("cov", g_cov, lambda a: a.cov),
                          ("sh", g_sh, lambda a: a.sh), ("opacity", g_op, lambda a: a.opacity)):
        ref = np.stack([pick(a) for a in accs])
        m = float(np.abs(ref).max())
        err = np.abs(g - ref).reshape(ref.shape[0], ref.shape[1], -1).max(-1)
        scale = m if m > 0 else 1.0
        if soft is None:
            soft = np.zeros(err.shape, bool)
        rep[name + "_rel"] = float(err[~soft].max(initial=0.0)) / scale
        rep[name + "_rel_borderline"] = float(err[soft].max(initial=0.0)) / scale
        rep[name + "_maxabs_ref"] = m
    rep["borderline_gaussians"] = int(soft.sum())
    return rep


def run_both(means, cov, sh, opacity, cams, cfg: RenderConfig, with_backward=True, seed=7, oracle_cache=None):
    """Everything on the same seeded inputs.  Returns (forward report, backward report|None).  ``oracle_cache`` (a dict)
    keeps the oracle's result between calls that differ only in the CUDA binning path."""
    dev = torch.device("cuda:0")
    cfg.check = "sync"
    gm, gc, gs, go, gcam = (t.to(dev) for t in (means, cov, sh, opacity, cams))
    color, depth, radii, state = forward_raw(cfg, gm, gc, gs, go, gcam)
    ins = inspect_workspace(state)
    N = cams.shape[0]
    g = torch.Generator().manual_seed(seed)
    dLc = torch.randn(N, 3, cfg.H, cfg.W, generator=g)
    dLd = torch.randn(N, cfg.H, cfg.W, generator=g)
    if oracle_cache is not None and "states" in oracle_cache:
        states, accs = oracle_cache["states"], oracle_cache["accs"]
    else:
        states, accs = oracle_batch(means, cov, sh, opacity, cams, cfg, dLc if with_backward else None, dLd)
        if oracle_cache is not None:
            oracle_cache.update(states=states, accs=accs)
    fwd = compare_forward(ins, color, depth, states, cfg)
    if radii is not None:
        fwd["out_radii_mismatch"] = int(sum((radii[n].cpu().numpy() != st.radii).sum() for n, st in enumerate(states)))
    bwd = None
    if with_backward:
        grads = backward_raw(cfg, state, dLc.to(dev), dLd.to(dev) if cfg.depth_mode else None)
        torch.cuda.synchronize()
        B = means.shape[0]
        softg = soft_gaussians(states, B, cfg.V)
        bwd = compare_backward(grads, accs, softg)
        softv = np.stack([oracle.gaussians_of_pixels(st, st.extra["borderline"]) for st in states])  # per view [N,G]
        if grads[4] is not None:
            ref2d = np.stack([st.extra["mean2d"] for st in states])
            e2d = np.abs(grads[4].cpu().numpy().astype(np.float64) - ref2d).max(-1)
            bwd["mean2d_rel"] = float(e2d[~softv].max(initial=0.0)) / max(float(np.abs(ref2d).max()), 1e-30)
        ins = inspect_workspace(state)
        acc_gpu = ins["grad_acc"].cpu().numpy().astype(np.float64)
        scr = np.stack([st.extra["screen"] for st in states])
        # accumulators of culled Gaussians (radius 0) are never initialised nor read (preprocess.cu / preprocess_bwd.cu)
        vis = (ins["radii"].cpu().numpy() > 0)[..., None]
        acc_gpu, scr = np.where(vis, acc_gpu, 0.0), np.where(vis, scr[..., :acc_gpu.shape[-1]], 0.0)
        # slots 0..4 hold the moments S of u = dL/dG * G over (dx, dy, dx^2, dx dy, dy^2); preprocess_backward turns them
        # into the reference's quantities with the per-Gaussian conic (A, B, C): do the same here
        con = ins["conic_opacity"].cpu().numpy().astype(np.float64)
        A, B, Cc = con[..., 0], con[..., 1], con[..., 2]
        sx, sy = acc_gpu[..., 0].copy(), acc_gpu[..., 1].copy()
        acc_gpu[..., 0] = -(A * sx + B * sy) * 0.5 * cfg.W
        acc_gpu[..., 1] = -(Cc * sy + B * sx) * 0.5 * cfg.H
        acc_gpu[..., 2:5] *= -0.5
        acc_gpu = np.where(vis, acc_gpu, 0.0)
        firmv = ~softv
        bwd["screen_rel"] = max(
            float(np.abs(acc_gpu[..., k] - scr[..., k])[firmv].max(initial=0.0)) / max(float(np.abs(scr[..., k]).max()), 1e-30)
            for k in range(10 if cfg.depth_mode else 9))
    return fwd, bwd
