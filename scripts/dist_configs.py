"""torchrun --nproc-per-node N scripts/dist_configs.py [c3] [c5]: the two multi-GPU shapes of SURVEY 8(d)/8(e) that
bench.py does not quote (bench.py = C2 per rank).

  c3  training batch: 14 independent scenes (131072 Gaussians, 4 views of 256x256 each, colour+depth fwd+bwd), scene s
      on rank s mod N, no collective.  8 ranks -> 2,2,2,2,2,2,1,1 scenes: the batch takes as long as the busiest rank.
  c5  stress: ONE scene of 2097152 Gaussians, 64 views of 1024x1024, views sharded (view j on rank j mod N), Gaussians
      replicated, fwd+bwd, then ONE in-place NCCL all-reduce of the flat per-Gaussian gradient slab (713 MB).

Device times (CUDA events), barrier + synchronize on both sides, max over ranks.  One JSON line per config from rank 0.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from transplat_b200.decoder import DecoderSplattingCUDA, Gaussians  # noqa: E402
from transplat_b200.parallel import forward_view_sharded, shard_indices  # noqa: E402
from transplat_b200.synthetic import make_scene, upstream_grads  # noqa: E402

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
dec = DecoderSplattingCUDA().to(dev)
one = torch.ones((), device=dev)


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    dist.barrier()
    t = torch.tensor([a.elapsed_time(b) / steps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def c3():
    n_scenes, V, HW = 14, 4, 256
    mine = shard_indices(n_scenes, rank, world)
    target, dLd = (t[None].to(dev) for t in upstream_grads(V, HW, HW, seed=99))
    S = []
    for s in mine:
        sc = make_scene(n_ctx=2, h=HW, w=HW, n_views=V, seed=2000 + s)
        leaf = lambda t: t[None].to(dev).requires_grad_(True)
        S.append((leaf(sc.means), leaf(sc.covariances), leaf(sc.harmonics), leaf(sc.opacities),
                  [t[None].to(dev) for t in (sc.extrinsics, sc.intrinsics, sc.near, sc.far)]))

    def batch():
        for m, c, h, o, cam in S:
            out, loss = dec.forward_mse(Gaussians(m, c, h, o), *cam, (HW, HW), target, depth_mode="depth")
            torch.autograd.grad([loss, out.depth], [m, c, h, o], [one, dLd])

    ms = timed(batch, steps=20, warmup=5)
    if rank == 0:
        print(json.dumps({"config": "c3", "n_gpus": world, "scenes": n_scenes, "views": n_scenes * V,
                          "scenes_per_rank": [len(shard_indices(n_scenes, r, world)) for r in range(world)],
                          "ms_per_batch": round(ms, 4), "scenes_per_s": round(n_scenes / ms * 1e3, 1),
                          "views_per_s": round(n_scenes * V / ms * 1e3, 1), "collective": "none"}), flush=True)


def c5():
    V, HW = 64, 1024
    sc = make_scene(n_ctx=2, h=HW, w=HW, n_views=V, seed=1240)
    mine = shard_indices(V, rank, world)
    idx = torch.tensor(mine)
    tgt, dld = upstream_grads(V, HW, HW, seed=99)
    target, dLd = tgt[idx][None].to(dev), dld[idx][None].to(dev)
    leaves = [t[None].to(dev).requires_grad_(True) for t in (sc.means, sc.covariances, sc.harmonics, sc.opacities)]
    cam = [t[None].to(dev) for t in (sc.extrinsics, sc.intrinsics, sc.near, sc.far)]
    G = sc.means.shape[0]
    del sc, tgt, dld
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    split = []

    def step():
        ev[0].record()
        out, _ = forward_view_sharded(dec, Gaussians(*leaves), *cam, (HW, HW), depth_mode="depth",
                                      overlap_chunks=int(os.environ.get("OVERLAP", "0")))
        loss = ((out.color - target) ** 2).mean() + (out.depth * dLd).sum()
        g = torch.autograd.grad(loss, leaves)   # backward kernels, then the in-place all-reduce of the flat slab
        ev[1].record()
        return g

    ms = timed(step, steps=3, warmup=2)
    # the collective alone, on a slab of the same size
    slab = torch.empty(G * (3 + 9 + 75 + 1), dtype=torch.float32, device=dev)
    ar = timed(lambda: dist.all_reduce(slab), steps=5, warmup=2)
    if rank == 0:
        print(json.dumps({"config": "c5", "n_gpus": world, "gaussians": G, "views": V, "views_per_rank": len(mine),
                          "ms_per_step": round(ms, 3), "views_per_s": round(V / ms * 1e3, 1),
                          "allreduce_bytes": slab.numel() * 4, "allreduce_ms_alone": round(ar, 3),
                          "allreduce_busbw_gbs": round(slab.numel() * 4 * 2 * (world - 1) / world / (ar * 1e-3) / 1e9, 1),
                          "overlap_chunks": int(os.environ.get("OVERLAP", "0")),
                          "collective": "in-place NCCL all-reduce of the gradient slab: one after the backward kernels "
                                        "(overlap_chunks 0) or per Gaussian range under the next range's kernels"}),
              flush=True)


for name in (sys.argv[1:] or ["c3", "c5"]):
    {"c3": c3, "c5": c5}[name]()
    torch.cuda.empty_cache()
dist.destroy_process_group()
