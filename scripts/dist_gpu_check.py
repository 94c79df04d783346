"""torchrun --nproc-per-node N scripts/dist_gpu_check.py: view sharding over N GPUs against the same scene rendered whole
on each rank, for every way of summing the per-Gaussian gradients (transplat_b200/parallel.py): the default choice, one
NCCL all-reduce, chunked NCCL, multimem.red fused into the producing kernel, the library's two-shot NVLS exchange whole
and range by range.  Each twice in a row (the symmetric-memory paths cycle two replicas)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from transplat_b200.decoder import DecoderSplattingCUDA, Gaussians
from transplat_b200.parallel import forward_view_sharded
from transplat_b200.synthetic import make_scene, upstream_grads

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
V, HW = 8, 128
sc = make_scene(n_ctx=2, h=64, w=64, n_views=V, render_hw=(HW, HW), seed=5)
target, dLd = (t.to(dev) for t in upstream_grads(V, HW, HW, seed=9))
dec = DecoderSplattingCUDA().to(dev)
cam = [t[None].to(dev) for t in (sc.extrinsics, sc.intrinsics, sc.near, sc.far)]


def leaves():
    return [t[None].to(dev).requires_grad_(True) for t in (sc.means, sc.covariances, sc.harmonics, sc.opacities)]


# whole scene on this rank
L0 = leaves()
out = dec.forward(Gaussians(*L0), *cam, (HW, HW), depth_mode="depth")
loss = ((out.color[0] - target) ** 2).sum() + (out.depth[0] * dLd).sum()
g_full = torch.autograd.grad(loss, L0)
# view-sharded: this rank renders views rank, rank+world, ... - every way of summing the gradients, twice each
# (the symmetric-memory paths cycle two replicas of the slab)
MODES = {"default": {}, "one_all_reduce": {"nvls_reduce": False}, "chunked_overlap": {"overlap_chunks": 3, "nvls_reduce": False},
         "fused_multimem": {"fused_reduce": True},
         "nvls_two_shot": {"nvls_reduce": True}, "nvls_two_shot_chunked": {"nvls_reduce": True, "overlap_chunks": 3}}
ok = True
for name, kw in MODES.items():
    for rep in range(2):
        L1 = leaves()
        try:
            out_s, mine = forward_view_sharded(dec, Gaussians(*L1), *cam, (HW, HW), depth_mode="depth", **kw)
        except Exception as exc:
            if rank == 0:
                print(f"{name}: unavailable ({exc!r})")
            break
        idx = torch.tensor(mine, device=dev)
        loss_s = ((out_s.color[0] - target[idx]) ** 2).sum() + (out_s.depth[0] * dLd[idx]).sum()
        g_shard = torch.autograd.grad(loss_s, L1)
        err = max(float((a - b).abs().max() / a.abs().max()) for a, b in zip(g_full, g_shard))
        img_err = float((out_s.color[0] - out.color[0, idx]).abs().max())
        t = torch.tensor([err, img_err], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        good = bool(t[0] < 1e-4 and t[1] < 1e-5)
        ok = ok and good
        if rank == 0:
            print(f"view sharding over {world} GPUs [{name} #{rep}]: views/rank {len(mine)}, max rel gradient error vs "
                  f"whole-scene {t[0]:.3e}, image error {t[1]:.3e}", "OK" if good else "FAIL")
# the default step captured in a CUDA graph (exchange kernels and barriers included) and replayed three times: the
# gradients each replay leaves behind against the whole-scene gradients
try:
    from transplat_b200.graphs import GraphedStep
    L2 = leaves()
    idx2 = torch.arange(rank, V, world, device=dev)

    def graphed_step():
        o, _ = forward_view_sharded(dec, Gaussians(*L2), *cam, (HW, HW), depth_mode="depth")
        l = ((o.color[0] - target[idx2]) ** 2).sum() + (o.depth[0] * dLd[idx2]).sum()
        return torch.autograd.grad(l, L2)
    gs = GraphedStep(graphed_step)
    for rep in range(3):
        g_shard = gs.replay()
        torch.cuda.synchronize()
        err = max(float((a - b).abs().max() / a.abs().max()) for a, b in zip(g_full, g_shard))
        t = torch.tensor([err], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        good = bool(t[0] < 1e-4)
        ok = ok and good
        if rank == 0:
            print(f"view sharding over {world} GPUs [default, CUDA-graph replay #{rep}]: max rel gradient error vs whole-scene "
                  f"{t[0]:.3e}", "OK" if good else "FAIL")
except Exception as exc:
    ok = False
    if rank == 0:
        print(f"graphed default step: FAILED ({exc!r})")
if rank == 0:
    print("ALL OK" if ok else "SOME FAILED")
dist.destroy_process_group()
