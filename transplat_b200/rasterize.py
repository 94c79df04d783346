"""torch.autograd.Function over the C-ABI of libtsplat.so: one call renders B scenes x V views
(colour and, optionally, the fused depth channel) and one backward call returns the gradients
summed over views.  PyTorch is plumbing here: device memory, the current stream, autograd.

Replaces ``_RasterizeGaussians`` of the reference's external extension; the
level-1 wrapper with the extension's exact call surface is in ``rasterizer.py``.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import torch

from . import _lib
from ._lib import COV_FULL9, COV_TRIU6, DEPTH_MODES, SH_G3M, SH_GM3, TsplatLayout, TsplatMse, TsplatProblem, TsplatStatus

# capacity of the (tile, Gaussian) pair list, in pairs per (view, Gaussian); grows when a call
# reports that it came close (see _Capacity)
_DEFAULT_PAIRS_PER_GAUSSIAN = float(os.environ.get("TSPLAT_PAIRS_PER_GAUSSIAN", "8"))
# "auto": the first call of a problem shape waits for the device's pair count (and retries with a larger
# list if it overflowed); afterwards the capacity tracks 2x the last observed count and calls are
# asynchronous.  "sync": every call waits.  "deferred": never wait (overflow surfaces at backward / next call).
_CHECK_MODE = os.environ.get("TSPLAT_CHECK", "auto")
# binning path: "bucket" (count per tile -> scan -> emit -> per-tile shared-memory sort; default) or
# "onesweep" (global LSD radix sort over the full 64-bit key); identical results
_SORT_MODE = os.environ.get("TSPLAT_SORT", "bucket")


class OverflowError_(RuntimeError):
    pass


class _Capacity:
    """d_cap per problem shape.  The pair count D is only known on the device; every call copies its status block back
    asynchronously and EVERY status is eventually read: a later call of any shape sweeps the completed ones, grows
    the capacity of their shape if needed and raises if one of them had overflowed (its images and gradients were
    incomplete).  Nothing is dropped unread: a forward-only loop whose host runs ahead keeps its statuses queued
    (bounded: beyond ``max_inflight`` the oldest one is waited for)."""

    max_inflight = 64

    def __init__(self):
        self.cap = {}
        self.inflight = []       # (key, pinned host block, event, d_cap, adaptive) of calls whose status has not been read
        self.calibrated = set()  # keys whose pair count has been observed at least once
        self.last_num_rendered = None  # pair count of the most recent call whose status has been read
        self.last_num_block_entries = None

    def sweep(self, wait: bool = False):
        """Read every status block that has arrived (all of them with ``wait``); raise for an overflowed call."""
        keep, bad = [], None
        for i, (key, host, event, d_cap, adaptive) in enumerate(self.inflight):
            overdue = len(self.inflight) - i > self.max_inflight
            if not (wait or overdue or event.query()):
                keep.append((key, host, event, d_cap, adaptive))
                continue
            event.synchronize()
            st = TsplatStatus.from_buffer_copy(host.numpy().tobytes())
            if adaptive:
                self.observe(key, st.num_rendered, st.num_block_entries)
            else:
                self.last_num_rendered = int(st.num_rendered)
            if st.overflow:
                bad = (int(st.num_rendered), d_cap)
        self.inflight = keep
        if bad is not None:
            raise OverflowError_(
                f"pair list overflow in an EARLIER render call: D={bad[0]} > d_cap={bad[1]}; its images and "
                "gradients were incomplete (the capacity has been raised for the next call).  Raise "
                "TSPLAT_PAIRS_PER_GAUSSIAN or run with TSPLAT_CHECK=sync.")

    def forget(self, event):
        """The caller has read this call's status itself."""
        self.inflight = [e for e in self.inflight if e[2] is not event]

    def get(self, key, n_items):
        if key not in self.cap:
            self.cap[key] = min(max(int(_DEFAULT_PAIRS_PER_GAUSSIAN * n_items), 4096), (1 << 30) - 1)
        return self.cap[key]

    def observe(self, key, num_rendered, num_block_entries=0):
        """Keep the capacity at >= 2x the last observed pair count D and the per-block lists' capacity
        (BLOCK_ENTRY_RATIO x d_cap entries) at >= 2x the last observed entry count E (never shrinks)."""
        self.calibrated.add(key)
        self.last_num_rendered = int(num_rendered)
        self.last_num_block_entries = int(num_block_entries)
        want = max(int(num_rendered * 2), int(num_block_entries * 2) // _lib.BLOCK_ENTRY_RATIO + 1)
        want = min(want + 4096, (1 << 30) - 1)
        if want > self.cap.get(key, 0):
            self.cap[key] = want

    grow = observe


_capacity = _Capacity()


@dataclass
class RenderConfig:
    H: int
    W: int
    V: int                     # views per scene
    sh_degree: int
    sh_layout: int = SH_GM3
    cov_layout: int = COV_TRIU6
    depth_mode: int = 0
    sh_l4: int = 1
    want_radii: bool = False
    want_means2d_grad: bool = False
    check: str | None = None   # None -> TSPLAT_CHECK env ("auto" default)
    d_cap: int | None = None
    debug: bool = False        # reference debug=True: synchronise and check after every stage
    sort: str | None = None    # None -> TSPLAT_SORT env ("bucket" default) | "bucket" | "onesweep"
    # view sharding (parallel.py): (process group, chunks) - the backward sums the per-Gaussian gradients over the group
    # itself, one Gaussian range at a time, each all-reduce overlapping the next range's kernels - or
    # (process group, "nvls", chunks, ctas) - the kernels write a symmetric slab, csrc/exchange.cu sums the replicas
    # through the switch (two-shot: multimem.ld_reduce + multimem.st), range by range under the next range's kernel;
    # (process group, "multimem") - the per-Gaussian kernel adds its values with multimem.red into a symmetric slab
    # replicated on every GPU of the NVSwitch domain (parallel.SymmetricGradSlab): no separate collective
    grad_reduce: tuple | None = None


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _problem(cfg: RenderConfig, means, cov, sh, opacity, cams) -> TsplatProblem:
    p = TsplatProblem()
    B, G = means.shape[0], means.shape[1]
    p.B, p.G, p.V, p.H, p.W = B, G, cfg.V, cfg.H, cfg.W
    p.sh_degree = cfg.sh_degree
    if sh.dim() == 3:      # colors_precomp [B,G,3]
        p.sh_coeffs = 0
    else:
        p.sh_coeffs = sh.shape[2] if cfg.sh_layout == SH_GM3 else sh.shape[3]
    p.sh_layout, p.cov_layout, p.depth_mode, p.sh_l4 = cfg.sh_layout, cfg.cov_layout, cfg.depth_mode, cfg.sh_l4
    p.flags = (_lib.FLAG_SORT_ONESWEEP if (cfg.sort or _SORT_MODE) == "onesweep" else 0) | (_lib.FLAG_DEBUG if cfg.debug else 0)
    p.means, p.cov, p.sh, p.opacity, p.cams = (means.data_ptr(), cov.data_ptr(), sh.data_ptr(),
                                               opacity.data_ptr(), cams.data_ptr())
    return p


class _State:
    """What backward needs from forward: the opaque workspace and the problem description."""
    __slots__ = ("problem", "workspace", "d_cap", "layout", "key", "tensors", "status_host", "event", "bwd_calls",
                 "mse")  # mse: (target, color, weight/numel factor) of a fused-loss forward, else None


def _check_inputs(means, cov, sh, opacity, cams, cfg):
    if not means.is_cuda:
        raise RuntimeError("transplat_b200 renders on a CUDA device only (no CPU fallback)")
    for name, t in (("means", means), ("cov", cov), ("sh", sh), ("opacity", opacity), ("cams", cams)):
        if t.dtype != torch.float32:
            raise RuntimeError(f"{name} must be float32")
        if t.device != means.device:
            raise RuntimeError(f"{name} is on {t.device}, means on {means.device}")
    if means.dim() != 3 or means.shape[-1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    B, G = means.shape[:2]
    want_cov = (B, G, 6) if cfg.cov_layout == COV_TRIU6 else (B, G, 3, 3)
    if tuple(cov.shape) != want_cov:
        raise RuntimeError(f"covariance must have shape {want_cov}, got {tuple(cov.shape)}")
    if tuple(opacity.shape) != (B, G):
        raise RuntimeError(f"opacity must have shape {(B, G)}, got {tuple(opacity.shape)}")
    if tuple(cams.shape) != (B * cfg.V, 64):
        raise RuntimeError(f"cameras must have shape {(B * cfg.V, 64)}, got {tuple(cams.shape)}")


def forward_raw(cfg: RenderConfig, means, cov, sh, opacity, cams, mse_target=None, mse_weight: float = 1.0,
                mse_psnr: bool = False):
    """Launch the forward pipeline; returns (color, depth|None, radii|None, state).  With ``mse_target`` [N,3,H,W] the
    render kernel's epilogue also accumulates the reference's MSE loss (loss_mse.py:23-31); it is returned as
    ``state.mse[3]`` (a 1-element device tensor) and consumed by backward_raw; with ``mse_psnr`` the same epilogue also
    reduces the per-view PSNR of the clipped images (metrics.py:12-19) into ``state.mse[4]`` [N]."""
    L = _lib.lib()
    _check_inputs(means, cov, sh, opacity, cams, cfg)
    means, cov, sh, opacity, cams = (t.contiguous() for t in (means, cov, sh, opacity, cams))
    B, G = means.shape[:2]
    N = B * cfg.V
    dev = means.device
    key = (dev.index, N, G, cfg.H, cfg.W)
    check = cfg.check or _CHECK_MODE
    # Under CUDA-graph capture (transplat_b200.graphs) nothing may touch the host: the capacity must already be
    # calibrated by an eager call of the same shape, no pinned status buffer is allocated and no event is polled;
    # the caller checks the overflow flag of the replayed work with read_status_device(state).
    capturing = torch.cuda.is_current_stream_capturing()
    if capturing:
        if key not in _capacity.calibrated and not cfg.d_cap:
            raise RuntimeError("CUDA-graph capture needs a calibrated pair-list capacity: run one eager call of this "
                               "problem shape first (or pass RenderConfig.d_cap)")
        check = "graph"
    else:
        _capacity.sweep()  # statuses of earlier calls that have arrived: grow capacities, raise for an overflow
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev)
        while True:
            d_cap = cfg.d_cap or (_capacity.cap[key] if capturing else _capacity.get(key, N * max(G, 1)))
            prob = _problem(cfg, means, cov, sh, opacity, cams)
            layout = TsplatLayout()
            _lib.check(L.tsplat_workspace_layout(C.byref(prob), d_cap, C.byref(layout)), "tsplat_workspace_layout")
            ws = torch.empty(layout.total_bytes, dtype=torch.uint8, device=dev)
            # colour and depth share one flat allocation (one D2H copy for a caller that wants both on the host)
            n_col, n_dep = N * 3 * cfg.H * cfg.W, (N * cfg.H * cfg.W if cfg.depth_mode else 0)
            images = torch.empty(_pad64(n_col) + n_dep, dtype=torch.float32, device=dev)
            color = images[:n_col].view(N, 3, cfg.H, cfg.W)
            depth = images[_pad64(n_col):_pad64(n_col) + n_dep].view(N, cfg.H, cfg.W) if cfg.depth_mode else None
            radii = torch.empty((N, G), dtype=torch.int32, device=dev) if cfg.want_radii else None
            loss = psnr = None
            if mse_target is not None:
                if tuple(mse_target.shape) != (N, 3, cfg.H, cfg.W) or mse_target.dtype != torch.float32 or \
                        mse_target.device != dev:
                    raise RuntimeError(f"mse_target must be float32 {(N, 3, cfg.H, cfg.W)} on {dev}")
                mse_target = mse_target.contiguous()
                loss = torch.empty(1, dtype=torch.float32, device=dev)
                psnr = torch.empty(N, dtype=torch.float32, device=dev) if mse_psnr else None
                m = TsplatMse(target=mse_target.data_ptr(), loss_sum=loss.data_ptr(), color=0, dL_dloss=0, scale=0.0,
                              loss_scale=mse_weight / float(n_col), psnr=psnr.data_ptr() if mse_psnr else 0)
                _lib.check(L.tsplat_forward_mse(C.byref(prob), _ptr(ws), layout.total_bytes, d_cap, _ptr(color),
                                                _ptr(depth), _ptr(radii), C.byref(m), C.c_void_p(stream.cuda_stream)),
                           "tsplat_forward_mse")
            else:
                _lib.check(L.tsplat_forward(C.byref(prob), _ptr(ws), layout.total_bytes, d_cap, _ptr(color),
                                            _ptr(depth), _ptr(radii), C.c_void_p(stream.cuda_stream)),
                           "tsplat_forward")
            if capturing:
                host = event = None
                break
            host = torch.empty(32, dtype=torch.uint8, pin_memory=True)
            _lib.check(L.tsplat_status_read(_ptr(ws), C.c_void_p(host.data_ptr()),
                                            C.c_void_p(stream.cuda_stream)), "tsplat_status_read")
            event = torch.cuda.Event()
            event.record(stream)
            if check == "sync" or (check == "auto" and key not in _capacity.calibrated and not cfg.d_cap):
                event.synchronize()
                st = TsplatStatus.from_buffer_copy(host.numpy().tobytes())
                if not cfg.d_cap:
                    _capacity.observe(key, st.num_rendered, st.num_block_entries)
                if st.overflow:
                    if cfg.d_cap:
                        raise OverflowError_(f"pair list overflow: D={st.num_rendered} (block-list entries "
                                             f"{st.num_block_entries}) > d_cap={d_cap}")
                    continue
            else:
                _capacity.inflight.append((key, host, event, d_cap, not cfg.d_cap))
            break
    state = _State()
    state.problem, state.workspace, state.d_cap, state.layout, state.key = prob, ws, d_cap, layout, key
    state.tensors = (means, cov, sh, opacity, cams)
    state.status_host, state.event = host, event
    state.bwd_calls = 0
    # color.detach(): the tensor handed to autograd gets grad_fn = the node that owns this state; keeping THAT object here
    # would close a reference cycle and leave the workspace (22 GiB at C5) to the cycle collector
    state.mse = ((mse_target, color.detach(), 2.0 * mse_weight / float(color.numel()), loss, psnr)
                 if mse_target is not None else None)
    return color, depth, radii, state


def read_status(state) -> TsplatStatus:
    if state.event is None:  # captured in a CUDA graph: read the device copy
        return read_status_device(state)
    state.event.synchronize()
    _capacity.forget(state.event)  # the caller has the flag now: a later call's sweep need not raise for it again
    return TsplatStatus.from_buffer_copy(state.status_host.numpy().tobytes())


def read_status_device(state) -> TsplatStatus:
    """Blocking read of the status block straight from the workspace (after a graph replay, or for debugging)."""
    return TsplatStatus.from_buffer_copy(state.workspace[:32].cpu().numpy().tobytes())


def backward_raw(cfg: RenderConfig, state, grad_color, grad_depth, grad_loss=None):
    L = _lib.lib()
    means, cov, sh, opacity, cams = state.tensors
    if state.event is None or torch.cuda.is_current_stream_capturing():
        pass  # CUDA-graph capture: the caller checks read_status_device(state) after a replay
    elif state.event.query() or (cfg.check or _CHECK_MODE) == "sync":
        st = read_status(state)
        _capacity.forget(state.event)
        if not cfg.d_cap:
            _capacity.observe(state.key, st.num_rendered, st.num_block_entries)
        if st.overflow:
            raise OverflowError_(
                f"pair list overflow in the forward pass: D={st.num_rendered} > d_cap={state.d_cap}; the rendered "
                "images were incomplete.  Raise TSPLAT_PAIRS_PER_GAUSSIAN or run with TSPLAT_CHECK=sync.")
    # else: the forward is still in flight (training step: backward is issued right behind it).  Do not stall the
    # host: enqueue the backward now (memory-safe even after an overflow); the forward's status block stays queued
    # in _capacity.inflight and the next call of any shape reads it and raises.
    dev = means.device
    N = means.shape[0] * cfg.V
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev)
        fused = state.mse is not None and grad_loss is not None
        if grad_color is not None:
            grad_color = grad_color.contiguous()
        elif not fused:
            grad_color = torch.zeros((N, 3, cfg.H, cfg.W), dtype=torch.float32, device=dev)
        if cfg.depth_mode:
            grad_depth = grad_depth.contiguous() if grad_depth is not None else torch.zeros(
                (N, cfg.H, cfg.W), dtype=torch.float32, device=dev)
        else:
            grad_depth = None
        # the four per-Gaussian gradients are views of ONE flat slab [means | cov | sh | opacity] (segments
        # 256-B aligned): a view-sharded caller all-reduces it in place (parallel.py), a host-side caller
        # copies it back with one transfer (flat_gradients below)
        sizes = [means.numel(), cov.numel(), sh.numel(), opacity.numel()]
        offs = [0]
        for n_ in sizes[:-1]:
            offs.append(offs[-1] + _pad64(n_))
        reduce = cfg.grad_reduce
        multimem = reduce is not None and reduce[1] == "multimem"
        nvls = reduce is not None and reduce[1] == "nvls"
        nccl_fallback = False
        if multimem or nvls:
            from . import parallel
            try:
                sym = parallel.SymmetricGradSlab.get(reduce[0], dev, offs[-1] + _pad64(sizes[-1]))
            except Exception as exc:  # no multicast support here: every rank of the box fails alike
                if multimem:
                    raise
                parallel.nvls_failed(exc)
                nvls, nccl_fallback = False, True
        if multimem or nvls:
            # multimem: this rank's replica, zeroed; every rank has zeroed its own before anyone adds
            slab = sym.begin_step() if multimem else sym.next_replica()
        else:
            slab = torch.empty(offs[-1] + sizes[-1], dtype=torch.float32, device=dev)
        g_means, g_cov, g_sh, g_op = (slab[o:o + n_].view(t.shape)
                                      for o, n_, t in zip(offs, sizes, (means, cov, sh, opacity)))
        g_m2d = (torch.empty((N, means.shape[1], 3), dtype=torch.float32, device=dev)
                 if cfg.want_means2d_grad else None)
        # the forward leaves the screen-space accumulators zero; every further backward re-clears them
        if state.bwd_calls:
            state.problem.flags |= _lib.FLAG_REZERO_GRAD_ACC
        state.bwd_calls += 1
        m = None
        if fused:
            target, color, scale = state.mse[:3]
            grad_loss = grad_loss.reshape(1).to(torch.float32).contiguous()
            m = TsplatMse(target=target.data_ptr(), loss_sum=0, color=color.data_ptr(), dL_dloss=grad_loss.data_ptr(),
                          scale=scale, loss_scale=0.0, psnr=0)

        # where the kernels write the gradients: the slab itself, or - multimem - the slab's multicast alias
        outs = [_ptr(t) for t in (g_means, g_cov, g_sh, g_op)]
        if multimem:
            outs = [C.c_void_p(sym.multicast_ptr + 4 * o) for o in offs]

        def launch():
            if fused:
                _lib.check(L.tsplat_backward_mse(C.byref(state.problem), _ptr(state.workspace), state.layout.total_bytes,
                                                 state.d_cap, _ptr(grad_color), _ptr(grad_depth), *outs, _ptr(g_m2d),
                                                 C.byref(m), C.c_void_p(stream.cuda_stream)), "tsplat_backward_mse")
            else:
                _lib.check(L.tsplat_backward(C.byref(state.problem), _ptr(state.workspace), state.layout.total_bytes,
                                             state.d_cap, _ptr(grad_color), _ptr(grad_depth), *outs, _ptr(g_m2d),
                                             C.c_void_p(stream.cuda_stream)), "tsplat_backward")

        if multimem:
            if g_m2d is not None:
                raise RuntimeError("grad_reduce='multimem' does not cover the screen-space gradient (level 1 only)")
            base_flags = state.problem.flags
            try:
                state.problem.flags = base_flags | _lib.FLAG_BWD_MULTIMEM
                launch()
            finally:
                state.problem.flags = base_flags
            sym.end_step()  # every rank's kernel has finished: the replicas hold the sums over all ranks' views
        elif nvls:
            if g_m2d is not None:
                raise RuntimeError("grad_reduce='nvls' does not cover the screen-space gradient (level 1 only)")
            _backward_nvls(state, launch, sym, stream, offs, sizes, means, int(reduce[2]), int(reduce[3]))
        elif reduce is None or means.shape[0] != 1 or nccl_fallback:
            launch()
            if reduce is not None:  # several scenes per call: their Gaussian ranges are not contiguous in the slab
                import torch.distributed as dist
                dist.all_reduce(slab, op=dist.ReduceOp.SUM, group=reduce[0])
        else:
            # Staged backward: the render stage fills the screen-space accumulators of all Gaussians, then
            # the per-Gaussian stage runs range by range; the four gradient pieces of a finished range are summed over
            # the ranks (NCCL runs on its own stream, ordered after what is enqueued here so far) while the kernels
            # of the next range run.
            import torch.distributed as dist
            group, chunks = reduce
            G = means.shape[1]
            per = -(-G // max(int(chunks), 1))
            step = -(-per // 64) * 64  # ranges start at multiples of 64 (the per-Gaussian kernel's CTA)
            base_flags = state.problem.flags
            works = []
            try:
                state.problem.flags = base_flags | _lib.FLAG_BWD_RENDER_ONLY
                launch()
                for g0 in range(0, G, step):
                    g1 = min(G, g0 + step)
                    state.problem.flags = (base_flags | _lib.FLAG_BWD_GAUSSIANS_ONLY) & ~_lib.FLAG_REZERO_GRAD_ACC
                    state.problem.g_begin, state.problem.g_count = g0, g1 - g0
                    launch()
                    # (torch's coalescing manager around these four made the step 30x slower on 2 B200: measured, not used)
                    for t in (g_means, g_cov, g_sh, g_op):
                        works.append(dist.all_reduce(t[0, g0:g1], op=dist.ReduceOp.SUM, group=group, async_op=True))
            finally:  # the problem record is reused by a repeated backward over the same forward state
                state.problem.flags = base_flags
                state.problem.g_begin = state.problem.g_count = 0
                for w in works:
                    w.wait()
    return g_means, g_cov, g_sh, g_op, g_m2d


def _backward_nvls(state, launch, sym, stream, offs, sizes, means, chunks, num_ctas):
    """View sharding, two-shot exchange through the switch (parallel.py, csrc/exchange.cu): the kernels have written / write
    this rank's replica of the symmetric slab; rank r then sums the r-th part of every piece over all replicas and stores
    the sums into all of them.  With chunks > 1 (one scene per call) the per-Gaussian stage runs range by range and the
    exchange of a finished range runs on a side stream under the kernel of the next range."""
    up4 = lambda n: (n + 3) & ~3  # whole 16-byte units: a piece may run into its segment's padding
    G = means.shape[1]
    if chunks <= 1 or means.shape[0] != 1:
        launch()
        sym.barrier(0)  # every rank's replica is complete
        sym.all_reduce([(o, up4(n)) for o, n in zip(offs, sizes)], num_ctas)
        sym.barrier(0)  # every rank's part is stored everywhere
        return
    per = -(-G // chunks)
    step = -(-per // 64) * 64  # ranges start at multiples of 64 (the per-Gaussian kernel's CTA; 16-byte units too)
    num_ctas = num_ctas or 16  # link-bound from 16 CTAs up (measured): leave the SMs to the kernel it runs under
    per_g = [n // G for n in sizes]  # floats per Gaussian in each of the four pieces
    side = sym.side_stream
    base_flags = state.problem.flags
    try:
        state.problem.flags = base_flags | _lib.FLAG_BWD_RENDER_ONLY
        launch()
        for g0 in range(0, G, step):
            g1 = min(G, g0 + step)
            state.problem.flags = (base_flags | _lib.FLAG_BWD_GAUSSIANS_ONLY) & ~_lib.FLAG_REZERO_GRAD_ACC
            state.problem.g_begin, state.problem.g_count = g0, g1 - g0
            launch()
            side.wait_stream(stream)
            with torch.cuda.stream(side):
                sym.barrier(0)  # this range is complete on every rank (and the previous range's exchange has ended)
                sym.all_reduce([(o + g0 * k, up4((g1 - g0) * k)) for o, k in zip(offs, per_g)], num_ctas)
        with torch.cuda.stream(side):
            sym.barrier(0)
        stream.wait_stream(side)
    finally:  # the problem record is reused by a repeated backward over the same forward state
        state.problem.flags = base_flags
        state.problem.g_begin = state.problem.g_count = 0


class _Rasterize(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means, cov, sh, opacity, means2d, cams, cfg):
        color, depth, radii, state = forward_raw(cfg, means, cov, sh, opacity, cams)
        ctx.cfg, ctx.state = cfg, state
        ctx.has_m2d = means2d is not None
        ctx.m2d_shape = tuple(means2d.shape) if means2d is not None else None
        outs = [color]
        outs.append(depth if depth is not None else torch.empty(0, device=color.device))
        outs.append(radii if radii is not None else torch.empty(0, dtype=torch.int32, device=color.device))
        ctx.mark_non_differentiable(outs[2])
        if depth is None:
            ctx.mark_non_differentiable(outs[1])
        return tuple(outs)

    @staticmethod
    def backward(ctx, grad_color, grad_depth, _grad_radii):
        cfg = ctx.cfg
        g_means, g_cov, g_sh, g_op, g_m2d = backward_raw(cfg, ctx.state, grad_color,
                                                         grad_depth if cfg.depth_mode else None)
        if ctx.has_m2d and g_m2d is not None:
            g_m2d = g_m2d.reshape(ctx.m2d_shape)
        return g_means, g_cov, g_sh, g_op, (g_m2d if ctx.has_m2d else None), None, None


class _RasterizeMse(torch.autograd.Function):
    """Render + the reference's MSE colour loss (src/loss/loss_mse.py:23-31) in the same kernels:
    the loss value comes out of the render-forward epilogue, and the backward forms dL/dcolor = dL_dcolor +
    dL/dloss * 2 w / numel * (colour - target) in the render-backward prologue - no elementwise kernel in between."""

    @staticmethod
    def forward(ctx, means, cov, sh, opacity, cams, cfg, target, weight, with_psnr):
        color, depth, radii, state = forward_raw(cfg, means, cov, sh, opacity, cams, mse_target=target,
                                                 mse_weight=float(weight), mse_psnr=bool(with_psnr))
        ctx.cfg, ctx.state = cfg, state
        ctx.set_materialize_grads(False)
        outs = [color, depth if depth is not None else torch.empty(0, device=color.device),
                radii if radii is not None else torch.empty(0, dtype=torch.int32, device=color.device),
                state.mse[3].reshape(()),
                state.mse[4] if with_psnr else torch.empty(0, device=color.device)]
        ctx.mark_non_differentiable(outs[2], outs[4])
        if depth is None:
            ctx.mark_non_differentiable(outs[1])
        return tuple(outs)

    @staticmethod
    def backward(ctx, grad_color, grad_depth, _grad_radii, grad_loss, _grad_psnr):
        cfg = ctx.cfg
        g_means, g_cov, g_sh, g_op, _ = backward_raw(cfg, ctx.state, grad_color,
                                                     grad_depth if cfg.depth_mode else None, grad_loss=grad_loss)
        return g_means, g_cov, g_sh, g_op, None, None, None, None, None


def rasterize_mse(means, cov, sh, opacity, cams, cfg: RenderConfig, target, weight: float = 1.0,
                  with_psnr: bool = False):
    """rasterize() + loss = weight * ((color - target) ** 2).mean() computed inside the render kernels.
    target [B*V,3,H,W].  Returns color, depth|None, radii|None, loss (0-d tensor) and, with ``with_psnr``, the
    per-view PSNR [B*V] the reference logs next to the loss (metrics.py:12-19, model_wrapper.py:149-153)."""
    color, depth, radii, loss, psnr = _RasterizeMse.apply(means, cov, sh, opacity, cams, cfg, target, weight, with_psnr)
    out = (color, (depth if cfg.depth_mode else None), (radii if cfg.want_radii else None), loss)
    return out + (psnr,) if with_psnr else out


def _pad64(n: int) -> int:
    return (n + 63) // 64 * 64


def flat_gradients(*grads):
    """The flat slab the gradients returned by one backward call are views of (None if they are not)."""
    base = getattr(grads[0], "_base", None)
    if base is None or any(getattr(g, "_base", None) is not base for g in grads):
        return None
    return base


def flat_images(color, depth=None):
    """The flat slab colour (and depth) of one forward call are views of (None if they are not)."""
    base = getattr(color, "_base", None)
    while base is not None and getattr(base, "_base", None) is not None:
        base = base._base
    return base


def rasterize(means, cov, sh, opacity, cams, cfg: RenderConfig, means2d=None):
    """means [B,G,3], cov [B,G,6]|[B,G,3,3], sh [B,G,M,3]|[B,G,3,M]|[B,G,3], opacity [B,G],
    cams [B*V,64] (cameras.pack_cameras) -> color [B*V,3,H,W], depth [B*V,H,W]|None, radii|None."""
    color, depth, radii = _Rasterize.apply(means, cov, sh, opacity, means2d, cams, cfg)
    return color, (depth if cfg.depth_mode else None), (radii if cfg.want_radii else None)


def inspect_workspace(state) -> dict:
    """Typed views of the inspectable workspace sections (parity tests).  Synchronises."""
    lay, ws, p = state.layout, state.workspace, state.problem
    N, G, H, W = p.B * p.V, p.G, p.H, p.W
    T = lay.tiles_x * lay.tiles_y
    st = read_status(state)
    D = min(int(st.num_rendered), state.d_cap)

    def sec(off, nbytes, dtype, shape):
        return ws[off:off + nbytes].view(dtype).reshape(shape)

    return {
        "num_rendered": int(st.num_rendered), "overflow": int(st.overflow), "T": T,
        "num_block_entries": int(st.num_block_entries),
        "sort_bits": lay.sort_bits,
        "depths": sec(lay.depths, 4 * N * G, torch.float32, (N, G)),
        "radii": sec(lay.radii, 4 * N * G, torch.int32, (N, G)),
        # interleaved 48-B records: x, y | conic.xyz, opacity | r, g, b, depth feature | id, cull threshold
        "xy": sec(lay.xy, lay.rec_stride * N * G, torch.float32, (N, G, lay.rec_stride // 4))[..., 0:2],
        "conic_opacity": sec(lay.xy, lay.rec_stride * N * G, torch.float32, (N, G, lay.rec_stride // 4))[..., 2:6],
        "rgbd": sec(lay.xy, lay.rec_stride * N * G, torch.float32, (N, G, lay.rec_stride // 4))[..., 6:10],
        "tiles_touched": sec(lay.tiles_touched, 4 * N * G, torch.int32, (N, G)),
        "clamped": sec(lay.clamped, N * G, torch.uint8, (N, G)),
        "keys": sec(lay.keys, 8 * D, torch.int64, (D,)),
        "vals": sec(lay.vals, 4 * D, torch.int32, (D,)),
        "ranges": sec(lay.ranges, 8 * N * T, torch.int32, (N, T, 2)),
        "final_T": sec(lay.final_T, 4 * N * H * W, torch.float32, (N, H, W)),
        "n_contrib": sec(lay.n_contrib, 4 * N * H * W, torch.int32, (N, H, W)),
        "grad_acc": sec(lay.grad_acc, 48 * N * G, torch.float32, (N, G, 12)),
        # per-block lists: jobs [N*T*8, 4] = first entry, count, (view*T + tile), 1 + deepest contributor's position
        "block_jobs": sec(lay.block_jobs, 16 * N * T * 8, torch.int32, (N * T * 8, 4)),
    }
