"""Generates tests/golden/ref_host_small.npz IN THE AUTHORING CONTAINER (needs /root/reference).

The UNMODIFIED reference host code (src/model/decoder/cuda_splatting.py: render_cuda and
render_depth_cuda) is imported with oracle/rasterizer_cpu.py standing in for the absent external
extension `diff_gaussian_rasterization`, and run on a small seeded scene.  The fixture therefore
pins (reference host code) o (oracle rasterizer); the GPU tests replay it through libtsplat.so.
    python tests/golden/make_golden.py
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import rasterizer_cpu as rc  # noqa: E402
from transplat_b200.synthetic import make_scene  # noqa: E402

mod = types.ModuleType("diff_gaussian_rasterization")
mod.GaussianRasterizationSettings = rc.GaussianRasterizationSettings
mod.GaussianRasterizer = rc.GaussianRasterizer
ref = rc.load_reference_cuda_splatting(mod)

H, W, V = 56, 72, 2
sc = make_scene(n_ctx=1, h=32, w=32, n_views=V, render_hw=(H, W), seed=77, near=1.25, far=60.0, d0=2.5, d1=9.0)
G = sc.means.shape[0]
cov = sc.covariances * 6.0
sh = sc.harmonics.clone()
sh[:, :, 0] -= 0.5
bg = torch.tensor([0.3, 0.1, 0.6])
rep = lambda t: t[None].expand(V, *t.shape).clone().requires_grad_(True)
means_r, cov_r, sh_r, op_r = rep(sc.means), rep(cov), rep(sh), rep(sc.opacities)
color = ref.render_cuda(sc.extrinsics, sc.intrinsics, sc.near, sc.far, (H, W), bg[None].expand(V, 3), means_r, cov_r,
                        sh_r, op_r)
g = torch.Generator().manual_seed(5)
w_c = torch.randn(color.shape, generator=g)
(color * w_c).sum().backward()
out = dict(extrinsics=sc.extrinsics, intrinsics=sc.intrinsics, near=sc.near, far=sc.far, bg=bg, means=sc.means, cov=cov,
           harmonics=sh, opacities=sc.opacities, color=color.detach(), w_color=w_c,
           # the reference sums the per-view copies through autograd of `repeat` (decoder_splatting_cuda.py:59-62)
           g_means=means_r.grad.sum(0), g_cov=cov_r.grad.sum(0), g_sh=sh_r.grad.sum(0), g_op=op_r.grad.sum(0))
w_d = torch.randn(V, H, W, generator=g)
out["w_depth"] = w_d
for mode in ("depth", "disparity", "relative_disparity", "log"):
    m2, c2, o2 = rep(sc.means), rep(cov), rep(sc.opacities)
    d = ref.render_depth_cuda(sc.extrinsics, sc.intrinsics, sc.near, sc.far, (H, W), m2, c2, o2, mode=mode)
    (d * w_d).sum().backward()
    out[f"depth_{mode}"] = d.detach()
    out[f"depth_{mode}_g_means"] = m2.grad.sum(0)
    out[f"depth_{mode}_g_cov"] = c2.grad.sum(0)
    out[f"depth_{mode}_g_op"] = o2.grad.sum(0)
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_host_small.npz")
np.savez_compressed(path, **{k: v.numpy() for k, v in out.items()})
print("wrote", path, os.path.getsize(path), "bytes; G =", G, "mean colour", float(color.mean()))
