"""Multi-GPU plumbing for the render path: one process per GPU, torch.distributed.

* scene sharding — scenes are independent: scene s -> rank s mod N, no data-path collective
  (the pattern of the reference's Lightning DDP over scenes, src/main.py:117-129).
* view sharding — one scene's target views split over ranks (view j -> rank j mod N), Gaussians
  replicated; forward needs no communication; backward sums the per-Gaussian gradients
  (means 3 + covariance 9 + SH 75 + opacity 1 floats per Gaussian) over the ranks - the only
  exchange on the path.  Two ways:
    - ONE NCCL all-reduce over a single flat slab behind the backward kernels (default), or
    - fused (``fused_reduce=True``): the kernel that PRODUCES the gradients (preprocess_backward) adds them with
      ``multimem.red.add.f32`` through the multicast alias of a symmetric slab (torch symmetric memory, NVLS): the
      NVSwitch reduces every value into every rank's replica while the kernel is still computing, so the reduction
      overlaps the math and no collective follows - only a barrier.  Exact, but every GPU receives `world` slabs:
      slower than the all-reduce already at 2 GPUs (measured), kept as an option.
    - two-shot NVLS (``nvls_reduce=True``, optionally with ``overlap_chunks``): the kernels write this rank's replica of
      a symmetric slab with ordinary stores; our own kernel (csrc/exchange.cu: multimem.ld_reduce + multimem.st) then
      sums the replicas in place through the switch, (1 + 1/world) slabs per GPU link and direction; with chunks, the
      exchange of one Gaussian range runs on a side stream under the per-Gaussian kernel of the next.
"""
from __future__ import annotations

import os
from typing import Optional, Sequence

import torch
import torch.distributed as dist

from .decoder import DecoderOutput, Gaussians


class SymmetricGradSlab:
    """Gradient slab replicated on every rank of an NVSwitch domain with a multicast alias (torch symmetric memory).
    Two replicas are cycled, so the gradients returned by one backward stay valid until the backward after the next.

    begin_step(): zero this rank's replica, then a device-side barrier (every replica is zero before any rank adds);
    end_step():   a device-side barrier (every rank's producing kernel has finished: the sums are complete everywhere).
    Both only enqueue work on the current stream."""
    _cache = {}

    def __init__(self, group, device, numel):
        import torch.distributed._symmetric_memory as symm
        self.bufs, self.hdls = [], []
        for _ in range(2):
            t = symm.empty(numel, dtype=torch.float32, device=device)
            h = symm.rendezvous(t, group if group is not None else dist.group.WORLD)
            if not h.multicast_ptr:
                raise RuntimeError("symmetric memory on this system has no multicast (NVLS) support: use the NCCL path")
            self.bufs.append(t)
            self.hdls.append(h)
        self.i = 0

    @classmethod
    def get(cls, group, device, numel):
        key = (id(group), device.index, int(numel))
        if key not in cls._cache:
            cls._cache[key] = cls(group, device, int(numel))
        return cls._cache[key]

    @property
    def multicast_ptr(self) -> int:
        return int(self.hdls[self.i].multicast_ptr)

    def begin_step(self):
        self.i ^= 1
        self.bufs[self.i].zero_()
        self.hdls[self.i].barrier(channel=0)
        return self.bufs[self.i]

    def end_step(self):
        self.hdls[self.i].barrier(channel=1)

    # ---- the two-shot exchange (csrc/exchange.cu): replicas written locally, summed through the switch afterwards ----
    def next_replica(self):
        """The replica this step's kernels write with ordinary stores (not cleared: they write every value)."""
        self.i ^= 1
        return self.bufs[self.i]

    def barrier(self, channel=0):
        self.hdls[self.i].barrier(channel=channel)

    def all_reduce(self, pieces, num_ctas=0):
        """Enqueue the in-place sum of `pieces` [(offset, count) in floats, multiples of 4, at most 4] of the current
        replica over the ranks, on the current stream.  The caller places barriers before and after."""
        import ctypes as C
        from . import _lib
        h = self.hdls[self.i]
        n = len(pieces)
        offs = (C.c_uint64 * 4)(*[int(o) for o, _ in pieces])
        cnts = (C.c_uint64 * 4)(*[int(c) for _, c in pieces])
        _lib.check(_lib.lib().tsplat_nvls_all_reduce(C.c_void_p(int(h.multicast_ptr)), offs, cnts, n, int(h.rank),
                                                     int(h.world_size), int(num_ctas),
                                                     C.c_void_p(torch.cuda.current_stream().cuda_stream)),
                   "tsplat_nvls_all_reduce")

    @property
    def side_stream(self):
        if not hasattr(self, "_side"):
            # high priority: the exchange's few CTAs must get SM slots as soon as any free up, or the block scheduler
            # keeps feeding the per-Gaussian kernel's queued CTAs and the two kernels run back to back
            self._side = torch.cuda.Stream(device=self.bufs[0].device, priority=-1)
        return self._side


# Exchange policy of forward_view_sharded() when the caller does not choose (measured on 2 / 4 / 8 B200 of one NVSwitch
# domain, profiles/r02_experiments.md)
NVLS_MIN_WORLD = int(os.environ.get("TSPLAT_NVLS_MIN_WORLD", "4"))   # ranks from which the NVLS exchange beats ncclAllReduce
NVLS_CHUNKS = int(os.environ.get("TSPLAT_NVLS_CHUNKS", "6"))         # Gaussian ranges a large slab is exchanged in
NVLS_CHUNK_MIN_BYTES = 256 << 20                                     # "large": below it the slab goes whole
_nvls_state = {"ok": None}


def nvls_available() -> bool:
    """False once setting up the symmetric slab has failed on this process (no multicast support, old driver): the
    exchange then stays with NCCL.  The first backward tries; rasterize.backward_raw reports the failure here."""
    return _nvls_state["ok"] is not False


def nvls_failed(exc) -> None:
    if _nvls_state["ok"] is not False:
        import warnings
        warnings.warn(f"NVLS gradient exchange unavailable ({exc!r}); using the NCCL all-reduce")
    _nvls_state["ok"] = False


def shard_indices(n: int, rank: int, world: int) -> list:
    """Round-robin ownership: item i belongs to rank i mod world."""
    return list(range(rank, n, world))


def shard_counts(n: int, world: int) -> list:
    return [len(range(r, n, world)) for r in range(world)]


class _SumGradsAcrossRanks(torch.autograd.Function):
    """Identity in forward; in backward the gradients of all tensors are packed into one flat
    buffer, summed over the group with a single all-reduce, and unpacked."""

    @staticmethod
    def forward(ctx, group, *tensors):
        ctx.group = group
        return tuple(t.view_as(t) for t in tensors)

    @staticmethod
    def backward(ctx, *grads):
        # the rasterizer's backward returns its four gradients as views of one flat slab: reduce it in place
        # (no packing copy); anything else is packed first
        from .rasterize import flat_gradients
        base = flat_gradients(*grads)
        if base is not None and base.is_contiguous():
            dist.all_reduce(base, op=dist.ReduceOp.SUM, group=ctx.group)
            return (None, *grads)
        shapes = [g.shape for g in grads]
        sizes = [g.numel() for g in grads]
        slab = torch.empty(sum(sizes), dtype=grads[0].dtype, device=grads[0].device)
        off = 0
        for g, n in zip(grads, sizes):
            slab[off:off + n].copy_(g.reshape(-1))
            off += n
        dist.all_reduce(slab, op=dist.ReduceOp.SUM, group=ctx.group)
        out, off = [], 0
        for shp, n in zip(shapes, sizes):
            out.append(slab[off:off + n].view(shp))
            off += n
        return (None, *out)


def replicate_with_summed_grads(gaussians: Gaussians, group=None) -> Gaussians:
    """Gaussians replicated on every rank whose gradients are all-reduced (sum) in backward."""
    m, c, h, o = _SumGradsAcrossRanks.apply(group, gaussians.means, gaussians.covariances, gaussians.harmonics,
                                            gaussians.opacities)
    return Gaussians(m, c, h, o)


def choose_exchange(world: int, scenes: int, gaussians_per_scene: int, on_cuda: bool, overlap_chunks: Optional[int],
                    fused_reduce: bool, nvls_reduce: Optional[bool], nvls_ok: bool) -> tuple:
    """How a view-sharded backward sums its per-Gaussian gradients: returns (mode, chunks) with mode one of
    "none" (one rank), "all_reduce" (one NCCL all-reduce of the flat slab), "staged_all_reduce" (per Gaussian range, under
    the next range's kernels), "nvls" (csrc/exchange.cu; chunks > 1: range by range on a side stream), "multimem" (the
    producing kernel adds with multimem.red).  ``overlap_chunks`` / ``nvls_reduce`` None = decide here, from what was
    measured on 2 / 4 / 8 B200 of one NVSwitch domain (profiles/r02_experiments.md): NCCL's peer-to-peer all-reduce wins
    at 2 GPUs, the two-shot exchange through the switch from 4 (its traffic per GPU does not grow with the number of
    GPUs); a large slab (C5: 738 MB) is exchanged range by range under the per-Gaussian kernel of the next range, for a
    small one (C2: 46 MB) the extra launches cost more than the overlap hides."""
    if world <= 1:
        return "none", 0
    auto_chunks = overlap_chunks is None
    chunks = 0 if auto_chunks else int(overlap_chunks)
    staged = chunks > 1 and scenes == 1 and on_cuda
    fused = bool(fused_reduce) and on_cuda
    if nvls_reduce is None:
        nvls_reduce = world >= NVLS_MIN_WORLD and (auto_chunks or chunks <= 1) and not fused_reduce and nvls_ok
    if auto_chunks and nvls_reduce and scenes == 1 and gaussians_per_scene * 88 * 4 >= NVLS_CHUNK_MIN_BYTES:
        chunks = NVLS_CHUNKS
    if nvls_reduce and not fused and on_cuda:
        return "nvls", chunks
    if fused:
        return "multimem", 0
    if staged:
        return "staged_all_reduce", chunks
    return "all_reduce", 0


def forward_view_sharded(decoder, gaussians: Gaussians, extrinsics, intrinsics, near, far, image_shape,
                         depth_mode: Optional[str] = None, group=None, overlap_chunks: Optional[int] = None,
                         fused_reduce: bool = False, nvls_reduce: Optional[bool] = None, nvls_ctas: int = 0) -> tuple:
    """Each rank renders views rank, rank+N, ... of every scene in the batch.  Returns
    (DecoderOutput for the local views, list of the global view indices rendered here).  Backward
    through the output all-reduces the per-Gaussian gradients once - or, with ``overlap_chunks`` > 1 (one scene per
    call, CUDA decoder), range by range inside the rasterizer's backward, each all-reduce running under the kernels of
    the next Gaussian range (713 MB at C5); with ``fused_reduce`` the
    per-Gaussian kernel itself reduces them through NVLS multicast memory (SymmetricGradSlab); with ``nvls_reduce`` (default:
    from 4 ranks up, when the system supports multicast memory) the
    kernels write this rank's replica and csrc/exchange.cu sums the replicas through the switch (``overlap_chunks`` > 1:
    range by range on a side stream, under the per-Gaussian kernel of the next range; ``nvls_ctas``: its grid)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    v = extrinsics.shape[1]
    mine = shard_indices(v, rank, world)
    mode, overlap_chunks = choose_exchange(world, gaussians.means.shape[0], gaussians.means.shape[1],
                                           gaussians.means.is_cuda, overlap_chunks, fused_reduce, nvls_reduce,
                                           nvls_available())
    nvls, fused, staged = mode == "nvls", mode == "multimem", mode == "staged_all_reduce"
    if mode == "all_reduce":
        gaussians = replicate_with_summed_grads(gaussians, group)
    if not mine:
        # a rank without views would skip backward and leave the others waiting in the all-reduce
        raise ValueError(f"view sharding needs at least one view per rank (views={v}, ranks={world})")
    idx = torch.arange(rank, v, world, device=extrinsics.device)  # = mine; built on the device (capturable in a CUDA graph)
    if nvls:
        extra = {"grad_reduce": (group, "nvls", int(overlap_chunks), int(nvls_ctas))}
    elif fused:
        extra = {"grad_reduce": (group, "multimem")}
    else:
        extra = {"grad_reduce": (group, int(overlap_chunks))} if staged else {}
    out = decoder.forward(gaussians, extrinsics[:, idx], intrinsics[:, idx], near[:, idx], far[:, idx], image_shape,
                          depth_mode=depth_mode, **extra)
    return out, mine


def forward_scene_sharded(decoder, scenes: Sequence, rank: int, world: int, render_one):
    """Scene sharding: `render_one(decoder, scene)` is called for the scenes this rank owns."""
    return {s: render_one(decoder, scenes[s]) for s in shard_indices(len(scenes), rank, world)}
