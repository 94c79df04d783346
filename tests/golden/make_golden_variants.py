"""Generates tests/golden/ref_host_variants.npz IN THE AUTHORING CONTAINER (needs /root/reference): the remaining
call-surface variants of the UNMODIFIED reference host code (src/model/decoder/cuda_splatting.py) -
render_cuda_orthographic (:139-229), render_cuda_with_stats (:256-372) and render_cuda(use_sh=False) (:130) -
with oracle/rasterizer_cpu.py standing in for the absent external extension, on the small seeded scene of
make_golden.py.  The GPU tests replay it through libtsplat.so.
    python tests/golden/make_golden_variants.py
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
cov = sc.covariances * 6.0
sh = sc.harmonics.clone()
sh[:, :, 0] -= 0.5
bg = torch.tensor([0.3, 0.1, 0.6])
rep = lambda t: t[None].expand(V, *t.shape).clone().requires_grad_(True)
g = torch.Generator().manual_seed(11)
out = dict(bg=bg)

# --- render_cuda_with_stats: images + (visible, high-contribution) counts and radii per view
means_r, cov_r, sh_r, op_r = rep(sc.means), rep(cov), rep(sh), rep(sc.opacities)
img, stats = ref.render_cuda_with_stats(sc.extrinsics, sc.intrinsics, sc.near, sc.far, (H, W), bg[None].expand(V, 3),
                                        means_r, cov_r, sh_r, op_r, opacity_threshold=0.3)
out["stats_color"] = img.detach()
out["stats_visible"] = torch.tensor([s.visible_gaussians for s in stats])
out["stats_high"] = torch.tensor([s.high_contribution_gaussians for s in stats])
out["stats_radii"] = torch.stack([s.radii for s in stats])

# --- use_sh=False: the single coefficient is the colour (colors_precomp)
col = torch.rand(sc.means.shape[0], 3, 1, generator=g)
means_r, cov_r, col_r, op_r = rep(sc.means), rep(cov), rep(col), rep(sc.opacities)
img = ref.render_cuda(sc.extrinsics, sc.intrinsics, sc.near, sc.far, (H, W), bg[None].expand(V, 3), means_r, cov_r,
                      col_r, op_r, use_sh=False)
w = torch.randn(img.shape, generator=g)
(img * w).sum().backward()
out.update(precomp_colors=col, precomp_color=img.detach(), precomp_w=w, precomp_g_means=means_r.grad.sum(0),
           precomp_g_cov=cov_r.grad.sum(0), precomp_g_col=col_r.grad.sum(0), precomp_g_op=op_r.grad.sum(0))

# --- render_cuda_orthographic: batch 1 per call (the reference writes the distance into one 4x4, :169-170)
width, height = torch.tensor([3.0]), torch.tensor([2.5])
near_o, far_o = torch.tensor([0.5]), torch.tensor([40.0])
ortho, ortho_g = [], []
w_o = torch.randn(V, 3, H, W, generator=g)
for i in range(V):
    m1, c1, s1, o1 = (t[None].clone().requires_grad_(True) for t in (sc.means, cov, sh, sc.opacities))
    dump = {}
    img = ref.render_cuda_orthographic(sc.extrinsics[i:i + 1], width, height, near_o, far_o, (H, W), bg[None], m1, c1, s1, o1,
                                       fov_degrees=0.1, dump=dump)
    (img * w_o[i:i + 1]).sum().backward()
    ortho.append(img[0].detach())
    ortho_g.append([m1.grad[0], c1.grad[0], s1.grad[0], o1.grad[0]])
out.update(ortho_width=width, ortho_height=height, ortho_near=near_o, ortho_far=far_o, ortho_color=torch.stack(ortho),
           ortho_w=w_o, ortho_dump_extrinsics=dump["extrinsics"], ortho_dump_fov_y=dump["fov_y"])
for k, name in enumerate(("means", "cov", "sh", "op")):
    out[f"ortho_g_{name}"] = torch.stack([gg[k] for gg in ortho_g])

# --- markVisible (GaussianRasterizer.markVisible)
from oracle import mark_visible  # noqa: E402
view0 = sc.extrinsics[0].inverse().T.contiguous()
out["visible_view0"] = torch.from_numpy(mark_visible(sc.means.numpy(), view0.numpy()).astype(np.uint8))
out["visible_viewmatrix"] = view0

# --- the step after the path: the UNMODIFIED LossMse (src/loss/loss_mse.py:21-31) on the reference's render_cuda output
import importlib.util  # noqa: E402


def load_loss_mse(ref_root="/root/reference"):
    saved = {k: v for k, v in sys.modules.items() if k == "src" or k.startswith("src.") or k == "jaxtyping"}
    for k in saved:
        del sys.modules[k]
    try:
        jt = types.ModuleType("jaxtyping")
        jt.Float = type("Float", (), {"__class_getitem__": classmethod(lambda cls, item: cls)})
        sys.modules["jaxtyping"] = jt
        for name, attrs in (("src", ()), ("src.loss", ()), ("src.dataset", ()), ("src.model", ()), ("src.model.decoder", ()),
                            ("src.dataset.types", ("BatchedExample",)), ("src.model.decoder.decoder", ("DecoderOutput",)),
                            ("src.model.types", ("Gaussians",))):
            m = types.ModuleType(name)
            m.__path__ = [f"{ref_root}/{name.replace('.', '/')}"]
            for a in attrs:
                setattr(m, a, object)
            sys.modules[name] = m
        mods = {}
        for name in ("src.loss.loss", "src.loss.loss_mse"):
            spec = importlib.util.spec_from_file_location(name, f"{ref_root}/{name.replace('.', '/')}.py")
            mods[name] = importlib.util.module_from_spec(spec)
            sys.modules[name] = mods[name]
            spec.loader.exec_module(mods[name])
        return mods["src.loss.loss_mse"]
    finally:
        for k in list(sys.modules):
            if k == "src" or k.startswith("src.") or k == "jaxtyping":
                del sys.modules[k]
        sys.modules.update(saved)


lm = load_loss_mse()
loss_fn = lm.LossMse(lm.LossMseCfgWrapper(lm.LossMseCfg(weight=0.7)))
means_r, cov_r, sh_r, op_r = rep(sc.means), rep(cov), rep(sh), rep(sc.opacities)
img = ref.render_cuda(sc.extrinsics, sc.intrinsics, sc.near, sc.far, (H, W), bg[None].expand(V, 3), means_r, cov_r, sh_r, op_r)
target = torch.rand(1, V, 3, H, W, generator=g)
pred = types.SimpleNamespace(color=img[None])
loss = loss_fn.forward(pred, {"target": {"image": target}}, None, 0)
loss.backward()
out.update(mse_target=target[0], mse_weight=torch.tensor(0.7), mse_loss=loss.detach(), mse_g_means=means_r.grad.sum(0),
           mse_g_cov=cov_r.grad.sum(0), mse_g_sh=sh_r.grad.sum(0), mse_g_op=op_r.grad.sum(0))



# --- the UNMODIFIED compute_psnr (src/evaluation/metrics.py:11-19) the training step logs next to the loss
# (model_wrapper.py:149-153); lpips / skimage (other metrics of that file) are absent here and stubbed
def load_metrics(ref_root="/root/reference"):
    stubs = {"jaxtyping": {"Float": type("Float", (), {"__class_getitem__": classmethod(lambda cls, item: cls)})},
             "lpips": {"LPIPS": object}, "skimage": {}, "skimage.metrics": {"structural_similarity": None}}
    saved = {k: sys.modules.get(k) for k in stubs}
    try:
        for k, attrs in stubs.items():
            m = types.ModuleType(k)
            for a, v in attrs.items():
                setattr(m, a, v)
            sys.modules[k] = m
        spec = importlib.util.spec_from_file_location("_ref_metrics", f"{ref_root}/src/evaluation/metrics.py")
        mm = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mm)
        return mm
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


out["mse_psnr"] = load_metrics().compute_psnr(target[0], img.detach())

path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_host_variants.npz")
np.savez_compressed(path, **{k: (v.numpy() if torch.is_tensor(v) else v) for k, v in out.items()})
print("wrote", path, os.path.getsize(path), "bytes;", "visible", out["stats_visible"].tolist(), "high", out["stats_high"].tolist(),
      "ortho mean", float(out["ortho_color"].mean()), "ortho nonbg px", int((out["ortho_color"][0] - bg[:, None, None]).abs().sum(0).gt(1e-6).sum()))
