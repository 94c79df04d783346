"""BASELINE config 1: the reference's OWN encoder, unmodified, run on the CPU with random-init weights on
synthetic images -> 131 072 Gaussians -> rendered at 256x256 through the CPU path (the unmodified reference decoder
file `cuda_splatting.py` with the CPU oracle standing in for the absent extension).  Plumbing only: it proves that what
the encoder (backbone, depth-aware deformable cross-view transformer, cost volume, Gaussian adapter) emits feeds the
render path's call surface unchanged.  No GPU, no trained weights, no dataset.

    python scripts/c1_plumbing.py            # prints one JSON line: pass/fail, shapes, wall times

Needs the reference tree (TSPLAT_REFERENCE_ROOT, default /root/reference): it exists in the authoring container only, so
this never runs on the GPU box.  The reference files are imported where they lie; nothing is copied.

Stand-ins for packages that are absent from this image (none of them is on the render path):
  * jaxtyping            annotations only
  * omegaconf / dacite / pytorch_lightning / wandb / skvideo / matplotlib: names only (never called); xformers and saes
    stay absent (the reference guards both imports with try / except ImportError)
  * mmcv.utils.ext_loader.load_ext               the compiled deformable-attention op is never called on the CPU path
  * mmcv.ops.multi_scale_deform_attn.multi_scale_deformable_attn_pytorch   its published pure-PyTorch definition
                                                 (bilinear `grid_sample` per level, weighted sum): restated below
  * e3nn.o3.matrix_to_angles / wigner_D          replaced by real-SH rotation matrices solved numerically from the real
                                                 spherical harmonics themselves (D_l = Y_l(R x) pinv(Y_l(x)), exact for a
                                                 rotation); `matrix_to_angles` just carries R through
  * torch.load / load_state_dict of checkpoints/depth_anything_v2_vitb.pth and the unimatch weights: random init instead
  * torch.cuda.synchronize                       the reference calls it unconditionally (encoder_trans.py:200-319): no-op
A random-init DepthAnything head outputs a constant map, so the reference's own (da - min) / (max - min) normalisation
(encoder_trans.py:222-224) is 0/0: the harness adds 1e-3 * N(0,1) noise to that one intermediate so the encoder's output
is finite, and says so in its report.
"""
from __future__ import annotations

import importlib
import importlib.abc
import importlib.util
import json
import math
import os
import sys
import time
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF = os.environ.get("TSPLAT_REFERENCE_ROOT", "/root/reference")

import numpy as np
import torch
import torch.nn.functional as F
import torchvision  # noqa: F401  (imported before the stand-ins are installed)


# ------------------------------------------------------------------------------------------------
# stand-ins
# ------------------------------------------------------------------------------------------------
class _Anything:
    """A name that can be subscripted, called, subclassed and attribute-walked; never does anything."""

    def __init__(self, *a, **k):
        pass

    def __class_getitem__(cls, item):
        return cls

    def __call__(self, *a, **k):
        return self

    def __getattr__(self, name):
        return _Anything()


def _stub_module(name: str) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__path__ = []  # a package: submodules resolve through the finder below

    def _attr(attr):
        if attr.startswith("__"):
            raise AttributeError(attr)
        return _Anything

    m.__getattr__ = _attr  # type: ignore[assignment]
    return m


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    """Resolves the absent third-party packages to permissive stubs, and the reference's `src.*` PACKAGES to path-only
    namespaces so that their __init__ chains (hydra, lightning, dataset registries) are skipped while every MODULE file
    is executed unmodified from where it lies."""
    ABSENT = ("jaxtyping", "omegaconf", "dacite", "pytorch_lightning", "lightning", "wandb", "skvideo", "matplotlib",
              "mmcv", "e3nn", "lpips", "colorama", "hydra", "beartype", "moviepy", "plyfile", "cv2",
              "svg", "colorspacious", "tabulate")

    # packages whose __init__ pulls in the training stack (hydra, lightning, dataset / loss registries): path-only
    SKIP_INIT = {"src", "src.model", "src.model.encoder", "src.model.decoder", "src.dataset", "src.dataset.shims", "src.misc",
                 "src.geometry", "src.visualization", "src.loss", "src.evaluation", "src.model.encoder.visualization"}

    def __init__(self):
        # decided once, before this finder is installed: which of the candidates really are absent from the image
        self.absent = set()
        for top in self.ABSENT:
            try:
                if importlib.util.find_spec(top) is None or top in ("mmcv", "e3nn"):
                    self.absent.add(top)
            except (ImportError, ValueError):
                self.absent.add(top)

    def find_spec(self, fullname, path=None, target=None):
        top = fullname.split(".")[0]
        if top in self.absent:
            return importlib.util.spec_from_loader(fullname, self, is_package=True)
        if top == "src":
            d = os.path.join(REF, *fullname.split("."))
            if os.path.isdir(d):
                init = os.path.join(d, "__init__.py")
                if fullname not in self.SKIP_INIT and os.path.exists(init):  # a leaf package: its __init__ runs unmodified
                    return importlib.util.spec_from_file_location(fullname, init, submodule_search_locations=[d])
                return importlib.util.spec_from_loader(fullname, self, is_package=True)
            f = d + ".py"
            if os.path.exists(f):
                return importlib.util.spec_from_file_location(fullname, f)
        return None

    def create_module(self, spec):
        if spec.name.split(".")[0] == "src":
            m = types.ModuleType(spec.name)
            m.__path__ = [os.path.join(REF, *spec.name.split("."))]
            return m
        return _stub_module(spec.name)

    def exec_module(self, module):
        pass


def multi_scale_deformable_attn_pytorch(value, value_spatial_shapes, sampling_locations, attention_weights):
    """The published pure-PyTorch definition of multi-scale deformable attention (Deformable DETR / mmcv.ops):
    value [bs, sum(H_l W_l), heads, dim], sampling_locations [bs, queries, heads, levels, points, 2] in [0,1],
    attention_weights [bs, queries, heads, levels, points] -> [bs, queries, heads * dim]."""
    bs, _, num_heads, embed_dims = value.shape
    _, num_queries, num_heads, num_levels, num_points, _ = sampling_locations.shape
    value_list = value.split([int(H_) * int(W_) for H_, W_ in value_spatial_shapes], dim=1)
    sampling_grids = 2 * sampling_locations - 1
    sampling_value_list = []
    for level, (H_, W_) in enumerate(value_spatial_shapes):
        H_, W_ = int(H_), int(W_)
        value_l_ = value_list[level].flatten(2).transpose(1, 2).reshape(bs * num_heads, embed_dims, H_, W_)
        sampling_grid_l_ = sampling_grids[:, :, :, level].transpose(1, 2).flatten(0, 1)
        sampling_value_list.append(F.grid_sample(value_l_, sampling_grid_l_, mode="bilinear", padding_mode="zeros",
                                                 align_corners=False))
    attention_weights = attention_weights.transpose(1, 2).reshape(bs * num_heads, 1, num_queries, num_levels * num_points)
    output = (torch.stack(sampling_value_list, dim=-2).flatten(-2) * attention_weights).sum(-1)
    return output.view(bs, num_heads * embed_dims, num_queries).transpose(1, 2).contiguous()


def _real_sh_matrix(l: int, pts: np.ndarray) -> np.ndarray:
    from kat import real_sh
    return np.array([[real_sh(l, m, p) for m in range(-l, l + 1)] for p in pts])  # [n, 2l+1]


_SH_PTS = None


def real_sh_rotation(l: int, R: np.ndarray) -> np.ndarray:
    """D with Y_l(R x) = D Y_l(x) for the real spherical harmonics of the rasterizer's table (tests/kat.py: real_sh)."""
    global _SH_PTS
    if _SH_PTS is None:
        rng = np.random.default_rng(0)
        p = rng.standard_normal((64, 3))
        _SH_PTS = p / np.linalg.norm(p, axis=1, keepdims=True)
    Y0 = _real_sh_matrix(l, _SH_PTS)
    Y1 = _real_sh_matrix(l, _SH_PTS @ R.T)
    return (np.linalg.pinv(Y0) @ Y1).T


def install_stand_ins():
    sys.meta_path.insert(0, _StubFinder())
    sys.path.insert(0, REF)
    import mmcv.ops.multi_scale_deform_attn as msda  # noqa: E402  (stub package)
    msda.multi_scale_deformable_attn_pytorch = multi_scale_deformable_attn_pytorch
    import mmcv.utils as mu
    mu.ext_loader = types.SimpleNamespace(load_ext=lambda *a, **k: _Anything())
    import e3nn.o3 as o3
    # matrix_to_angles carries the rotation through; wigner_D builds the real-SH rotation of that degree
    o3.matrix_to_angles = lambda R: (R, None, None)

    def wigner_D(degree, alpha, beta, gamma):
        R = alpha.detach().cpu().double().numpy()
        flat = R.reshape(-1, 3, 3)
        out = np.stack([real_sh_rotation(degree, r) for r in flat])
        return torch.from_numpy(out.reshape(*R.shape[:-2], 2 * degree + 1, 2 * degree + 1)).to(alpha.dtype)

    o3.wigner_D = wigner_D
    torch.cuda.synchronize = lambda *a, **k: None
    # global config the encoder reads (encoder_trans.py:85,129)
    gc = importlib.import_module("src.global_cfg")
    cfg = types.SimpleNamespace(mode="test", dataset=types.SimpleNamespace(view_sampler=types.SimpleNamespace(num_context_views=2)))
    gc.get_cfg = lambda: cfg
    gc.cfg = cfg


def build_encoder():
    """EncoderTrans with the reference's own config values (config/model/encoder/trans.yaml, config/experiment/re10k.yaml)."""
    import yaml
    enc_cfg = yaml.safe_load(open(os.path.join(REF, "config", "model", "encoder", "trans.yaml")))
    exp_path = os.path.join(REF, "config", "experiment", "re10k.yaml")
    if os.path.exists(exp_path):
        exp = yaml.safe_load(open(exp_path)) or {}
        for k, v in ((exp.get("model") or {}).get("encoder") or {}).items():
            if isinstance(v, dict) and isinstance(enc_cfg.get(k), dict):
                enc_cfg[k].update(v)
            else:
                enc_cfg[k] = v

    def ns(d):
        return types.SimpleNamespace(**{k: ns(v) if isinstance(v, dict) else v for k, v in d.items()})

    orig_load, orig_lsd = torch.load, torch.nn.Module.load_state_dict
    torch.load = lambda *a, **k: {}
    torch.nn.Module.load_state_dict = lambda self, *a, **k: None  # checkpoints are absent: random init
    try:
        mod = importlib.import_module("src.model.encoder.encoder_trans")
        cfg = ns(enc_cfg)
        cfg.unimatch_weights_path = None
        torch.manual_seed(0)
        enc = mod.EncoderTrans(cfg).eval()
    finally:
        torch.load, torch.nn.Module.load_state_dict = orig_load, orig_lsd
    return enc, cfg


def synthetic_context(h=256, w=256, v=2, seed=3):
    g = torch.Generator().manual_seed(seed)
    low = torch.rand(1, v, 3, 16, 16, generator=g)
    img = F.interpolate(low.flatten(0, 1), size=(h, w), mode="bicubic", align_corners=False).clamp(0, 1).view(1, v, 3, h, w)
    ext = torch.eye(4).repeat(1, v, 1, 1)
    for k in range(v):
        ext[0, k, 0, 3] = 0.3 * k
    K = torch.tensor([[0.9, 0, 0.5], [0, 0.9, 0.5], [0, 0, 1.0]]).repeat(1, v, 1, 1)
    return {"image": img, "extrinsics": ext, "intrinsics": K, "near": torch.full((1, v), 1.0), "far": torch.full((1, v), 100.0),
            "index": torch.arange(v)[None]}


def run(render_views: int = 1) -> dict:
    report = {"config": "C1 plumbing: reference EncoderTrans on CPU (random init) -> 131072 Gaussians -> CPU render 256x256"}
    if not os.path.exists(os.path.join(REF, "src", "model", "encoder", "encoder_trans.py")):
        report.update(status="skipped", why="reference tree not present")
        return report
    install_stand_ins()
    t0 = time.perf_counter()
    enc, cfg = build_encoder()
    report["encoder_parameters"] = int(sum(p.numel() for p in enc.parameters()))
    report["build_s"] = round(time.perf_counter() - t0, 2)
    # the one perturbation: a random-init DepthAnything head is constant => the reference's min/max normalisation is 0/0
    da_forward = enc.da_model.forward

    def da_noisy(x):
        depth, feat = da_forward(x)
        g = torch.Generator().manual_seed(11)
        return depth + 1e-3 * torch.randn(depth.shape, generator=g), feat

    enc.da_model.forward = da_noisy
    ctx = synthetic_context()
    t0 = time.perf_counter()
    with torch.no_grad():
        gaussians = enc(ctx, global_step=0, deterministic=True)
    report["encoder_forward_s"] = round(time.perf_counter() - t0, 2)
    G = gaussians.means.shape[1]
    report["gaussians"] = int(G)
    report["shapes"] = {k: list(getattr(gaussians, k).shape) for k in ("means", "covariances", "harmonics", "opacities")}
    finite = all(bool(torch.isfinite(getattr(gaussians, k)).all()) for k in ("means", "covariances", "harmonics", "opacities"))
    report["finite"] = finite

    # ---- render through the UNMODIFIED reference decoder file on the CPU (oracle as the extension)
    from oracle import rasterizer_cpu
    ref_cs = rasterizer_cpu.load_reference_cuda_splatting(rasterizer_cpu, REF)
    tgt = synthetic_context(v=render_views, seed=5)
    tgt["extrinsics"][0, :, 0, 3] = 0.15
    t0 = time.perf_counter()
    rep = lambda t: t[0][None].expand(render_views, *t.shape[1:])
    color = ref_cs.render_cuda(tgt["extrinsics"][0], tgt["intrinsics"][0], tgt["near"][0], tgt["far"][0], (256, 256),
                               torch.zeros(render_views, 3), rep(gaussians.means), rep(gaussians.covariances),
                               rep(gaussians.harmonics), rep(gaussians.opacities))
    depth = ref_cs.render_depth_cuda(tgt["extrinsics"][0], tgt["intrinsics"][0], tgt["near"][0], tgt["far"][0], (256, 256),
                                     rep(gaussians.means), rep(gaussians.covariances), rep(gaussians.opacities), mode="depth")
    color, depth = color.detach(), depth.detach()
    report["render_s"] = round(time.perf_counter() - t0, 2)
    st = rasterizer_cpu._RasterizeGaussians.last_state
    report["num_rendered_last_pass"] = int(st.num_rendered)
    report["visible_gaussians_last_pass"] = int((st.radii > 0).sum())
    report["color"] = {"shape": list(color.shape), "finite": bool(torch.isfinite(color).all()), "mean": round(float(color.mean()), 5)}
    report["depth"] = {"shape": list(depth.shape), "finite": bool(torch.isfinite(depth).all()), "mean": round(float(depth.mean()), 5)}
    # the same Gaussians through the batched level-2 contract of this repository (layouts only; CPU oracle underneath)
    import oracle
    from transplat_b200.cameras import pack_cameras
    cams = pack_cameras(tgt["extrinsics"][0], tgt["intrinsics"][0], tgt["near"][0], tgt["far"][0], torch.zeros(render_views, 3)).numpy()
    sc = oracle.Scene(gaussians.means[0].numpy(), gaussians.covariances[0].numpy(), gaussians.harmonics[0].numpy(),
                      gaussians.opacities[0].numpy(), 4, oracle.SH_G3M, oracle.COV_FULL9)
    st2 = oracle.forward(sc, cams[0], 256, 256, depth_mode="depth")
    report["batched_contract_max_abs_vs_reference_file"] = {
        "color": float(np.abs(st2.color - color[0].numpy()).max()), "depth": float(np.abs(st2.depth - depth[0].numpy()).max())}
    ok = (finite and G == 131072 and report["color"]["finite"] and report["depth"]["finite"] and st.num_rendered > 0
          and report["batched_contract_max_abs_vs_reference_file"]["color"] <= 1e-5)
    report["status"] = "pass" if ok else "fail"
    report["perturbation"] = "1e-3 * N(0,1) added to the random-init DepthAnything output (constant otherwise: 0/0 in the reference's own normalisation)"
    return report


if __name__ == "__main__":
    print(json.dumps(run()), flush=True)
