"""world_size-2 gloo test of the view-sharding host logic (transplat_b200/parallel.py): the
per-Gaussian gradients of the two ranks are summed by ONE all-reduce over a flat slab."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from transplat_b200.parallel import shard_counts, shard_indices


def test_round_robin_sharding():
    assert shard_indices(14, 1, 8) == [1, 9]
    assert shard_counts(14, 8) == [2, 2, 2, 2, 2, 2, 1, 1]      # 14 scenes on 8 GPUs
    assert sorted(sum((shard_indices(67, r, 4) for r in range(4)), [])) == list(range(67))


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from transplat_b200.decoder import Gaussians
    from transplat_b200.parallel import replicate_with_summed_grads
    torch.manual_seed(0)
    G = 50
    g = Gaussians(*(t.requires_grad_(True) for t in (torch.randn(1, G, 3), torch.randn(1, G, 3, 3),
                                                      torch.randn(1, G, 3, 25), torch.randn(1, G))))
    r = replicate_with_summed_grads(g)
    # a stand-in "render" whose gradient differs per rank
    loss = (rank + 1) * (r.means.sum() + 2 * r.covariances.sum() + 3 * r.harmonics.sum() + 4 * r.opacities.sum())
    loss.backward()
    tot = sum(range(1, world + 1))
    ok = (torch.allclose(g.means.grad, torch.full_like(g.means, 1.0 * tot)) and
          torch.allclose(g.covariances.grad, torch.full_like(g.covariances, 2.0 * tot)) and
          torch.allclose(g.harmonics.grad, torch.full_like(g.harmonics, 3.0 * tot)) and
          torch.allclose(g.opacities.grad, torch.full_like(g.opacities, 4.0 * tot)))
    # gradients that arrive as views of ONE flat slab (what the rasterizer's backward returns) take the in-place path
    class _SlabGrads(torch.autograd.Function):
        @staticmethod
        def forward(ctx, m, c, h, o):
            ctx.shapes = [t.shape for t in (m, c, h, o)]
            return m.sum() + c.sum() + h.sum() + o.sum()

        @staticmethod
        def backward(ctx, gout):
            sizes = [int(torch.Size(s).numel()) for s in ctx.shapes]
            slab = torch.empty(sum(sizes) + 64 * 3)
            outs, off = [], 0
            for k, (n, shp) in enumerate(zip(sizes, ctx.shapes)):
                v = slab[off:off + n].view(shp)
                v.fill_(float(k + 1) * (rank + 1))
                outs.append(v)
                off += n + 64
            return tuple(outs)

    g2 = Gaussians(*(t.detach().clone().requires_grad_(True) for t in (g.means, g.covariances, g.harmonics, g.opacities)))
    r2 = replicate_with_summed_grads(g2)
    _SlabGrads.apply(r2.means, r2.covariances, r2.harmonics, r2.opacities).backward()
    for k, t in enumerate((g2.means, g2.covariances, g2.harmonics, g2.opacities)):
        ok = ok and torch.allclose(t.grad, torch.full_like(t, float(k + 1) * tot))
    ret[rank] = bool(ok)
    dist.destroy_process_group()


def test_gradient_allreduce_world2():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    assert dict(ret) == {0: True, 1: True}
