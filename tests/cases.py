"""Seeded test cases shared by the GPU parity tests and scripts/gpu_check.py."""
from __future__ import annotations

import torch

from transplat_b200._lib import COV_FULL9, COV_TRIU6, SH_G3M, SH_GM3
from transplat_b200.cameras import pack_cameras
from transplat_b200.rasterize import RenderConfig
from transplat_b200.synthetic import make_scene


def case(name: str):
    """Returns (means[B,G,3], cov, sh, opacity, cams[B*V,64], cfg)."""
    if name == "tiny_l1":          # level-1 layouts, 1 view, partial tiles (40x48), scale != 1, bg != 0
        sc = make_scene(n_ctx=1, h=24, w=20, n_views=1, render_hw=(40, 48), seed=5, near=1.5, far=50.0, d0=3.0, d1=8.0)
        return _pack(sc, level=1, B=1, depth_mode=0, cov_mul=9.0, bg=(0.2, 0.5, 0.7))
    if name == "small_l2_depth":   # level-2 layouts, 2 scenes x 3 views, fused depth
        sc = make_scene(n_ctx=2, h=32, w=32, n_views=3, render_hw=(64, 80), seed=11)
        return _pack(sc, level=2, B=2, depth_mode=1, cov_mul=4.0, bg=(0.0, 0.0, 0.0))
    if name == "ragged_l2":        # G not a multiple of 4/128 (no TMA path), odd image size, disparity depth
        sc = make_scene(n_ctx=1, h=31, w=29, n_views=2, render_hw=(50, 70), seed=3, near=2.125, far=4.525, d0=2.2, d1=4.4)
        return _pack(sc, level=2, B=2, depth_mode=2, cov_mul=2.0, bg=(1.0, 1.0, 1.0), drop=5)
    if name == "mid_l2":           # 8192 Gaussians x 4 views at 128x128, relative disparity
        sc = make_scene(n_ctx=2, h=64, w=64, n_views=4, render_hw=(128, 128), seed=21)
        return _pack(sc, level=2, B=1, depth_mode=3, cov_mul=1.0, bg=(0.0, 0.0, 0.0))
    if name == "deg2_l1":          # SH degree 2 with 9 coefficients, level 1
        sc = make_scene(n_ctx=1, h=32, w=32, n_views=1, render_hw=(64, 64), seed=8, sh_degree=2)
        return _pack(sc, level=1, B=1, depth_mode=0, cov_mul=4.0, bg=(0.1, 0.1, 0.1))
    if name in ("deg0_l2", "deg1_l2", "deg3_l2"):  # the SH degrees the flagship config does not use (1, 4, 16 coefficients)
        deg = int(name[3])
        sc = make_scene(n_ctx=1, h=40, w=40, n_views=2, render_hw=(56, 72), seed=50 + deg, sh_degree=deg)
        return _pack(sc, level=2, B=1, depth_mode=1, cov_mul=5.0, bg=(0.3, 0.1, 0.6))
    if name == "dense_big":        # few tiles, very long lists: > 4096 and > 16384 pairs per tile (big-tile sort paths)
        sc = make_scene(n_ctx=1, h=144, w=144, n_views=2, render_hw=(32, 48), seed=31)
        return _pack(sc, level=2, B=1, depth_mode=1, cov_mul=120.0, bg=(0.0, 0.0, 0.0))
    if name == "c4_dtu":           # BASELINE config 4a: DTU, 3 context views -> 196608 Gaussians, 1 target view, depth
        sc = make_scene(n_ctx=3, h=256, w=256, n_views=1, seed=41, near=2.125, far=4.525, d0=2.2, d1=4.4)
        return _pack(sc, level=2, B=1, depth_mode=1, cov_mul=1.0, bg=(0.0, 0.0, 0.0))
    if name == "coplanar":         # every Gaussian at the same depth in every view: all sort keys of a tile tie, the
        # distribution sort bails out and the stable radix fallback must return Gaussian order
        sc = make_scene(n_ctx=1, h=96, w=96, n_views=2, render_hw=(48, 64), seed=17)
        sc.means[:, 2] = 4.0
        sc.extrinsics[:] = torch.eye(4)
        return _pack(sc, level=2, B=1, depth_mode=1, cov_mul=30.0, bg=(0.0, 0.0, 0.0))
    if name == "many_views_l2":    # 11 views of a scene in one call: more views than the per-Gaussian kernels have warps
        # (8 per CTA), so preprocess forward AND backward take their second round with 3 live warps
        sc = make_scene(n_ctx=1, h=36, w=36, n_views=11, render_hw=(48, 40), seed=61)
        return _pack(sc, level=2, B=2, depth_mode=1, cov_mul=3.0, bg=(0.2, 0.0, 0.4), drop=3)
    if name == "many_views_l1":    # the same with the level-1 layouts ([G, M, 3] harmonics, [G, 6] covariances), 9 views
        sc = make_scene(n_ctx=1, h=32, w=32, n_views=9, render_hw=(40, 56), seed=62, sh_degree=3)
        return _pack(sc, level=1, B=1, depth_mode=0, cov_mul=3.0, bg=(0.0, 0.3, 0.1))
    if name == "sub_tile":         # an image smaller than one 16x16 tile (9 x 7 pixels): one partial tile, partial blocks
        sc = make_scene(n_ctx=1, h=12, w=12, n_views=2, render_hw=(7, 9), seed=63)
        return _pack(sc, level=2, B=1, depth_mode=1, cov_mul=0.5, bg=(0.5, 0.5, 0.5))
    if name == "c2":               # BASELINE config 2: 131072 Gaussians, 4 views, 256x256, colour+depth
        sc = make_scene()
        return _pack(sc, level=2, B=1, depth_mode=1, cov_mul=1.0, bg=(0.0, 0.0, 0.0))
    if name == "c3_two_scenes":    # BASELINE config 3 shape: a batch of scenes at full size in ONE call (2 x 131072 x 4 views)
        sc = make_scene()
        return _pack(sc, level=2, B=2, depth_mode=1, cov_mul=1.0, bg=(0.0, 0.0, 0.0))
    if name == "c4b_video":        # BASELINE config 4b: the re10k video index, 67 target views of one scene in ONE call
        sc = make_scene(n_views=67, seed=43)
        return _pack(sc, level=2, B=1, depth_mode=1, cov_mul=1.0, bg=(0.0, 0.0, 0.0))
    if name == "c5_stress":        # BASELINE config 5: 2 097 152 Gaussians at 1024x1024 (2 of the 64 views)
        sc = make_scene(n_ctx=2, h=1024, w=1024, n_views=2, seed=55)
        return _pack(sc, level=2, B=1, depth_mode=1, cov_mul=1.0, bg=(0.0, 0.0, 0.0))
    raise KeyError(name)


def _pack(sc, level, B, depth_mode, cov_mul, bg, drop=0):
    G = sc.means.shape[0] - drop
    V = sc.extrinsics.shape[0]
    H, W = sc.image_shape
    means = sc.means[:G]
    cov = sc.covariances[:G] * cov_mul
    sh = sc.harmonics[:G].clone()
    sh[:, :, 0] -= 0.6  # some colours clamp at 0
    op = sc.opacities[:G]
    d_sh = sh.shape[-1]
    degree = int(round(d_sh ** 0.5)) - 1
    def rep(t):
        # scene b>0 = scene 0 shifted a little, so the scenes differ
        return torch.stack([t + 0.01 * b if t is means else t for b in range(B)]).contiguous()
    means_b, cov_b, sh_b, op_b = rep(means), rep(cov), rep(sh), rep(op)
    ext = sc.extrinsics[None].expand(B, V, 4, 4).reshape(B * V, 4, 4)
    intr = sc.intrinsics[None].expand(B, V, 3, 3).reshape(B * V, 3, 3)
    near = sc.near[None].expand(B, V).reshape(-1)
    far = sc.far[None].expand(B, V).reshape(-1)
    cams = pack_cameras(ext, intr, near, far, torch.tensor([bg]).expand(B * V, 3))
    if level == 1:
        cov_b = cov_b.reshape(B, G, 9)[:, :, [0, 1, 2, 4, 5, 8]].contiguous()
        sh_b = sh_b.permute(0, 1, 3, 2).contiguous()
        cfg = RenderConfig(H=H, W=W, V=V, sh_degree=degree, sh_layout=SH_GM3, cov_layout=COV_TRIU6,
                           depth_mode=depth_mode, want_radii=True, want_means2d_grad=True)
    else:
        cfg = RenderConfig(H=H, W=W, V=V, sh_degree=degree, sh_layout=SH_G3M, cov_layout=COV_FULL9,
                           depth_mode=depth_mode, want_radii=True)
    return means_b, cov_b, sh_b, op_b, cams, cfg
