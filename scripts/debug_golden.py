import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from transplat_b200.decoder import DecoderSplattingCUDA, Gaussians
from transplat_b200.cameras import pack_cameras, pack_cameras_device
from transplat_b200.rasterize import RenderConfig, rasterize
from transplat_b200._lib import SH_G3M, COV_FULL9
z = np.load(os.path.join(ROOT, "tests/golden/ref_host_small.npz"))
t = lambda k: torch.from_numpy(z[k]).cuda()
H, W = z["color"].shape[-2:]
V = z["extrinsics"].shape[0]
bg = t("bg")
cam_h = pack_cameras(torch.from_numpy(z["extrinsics"]), torch.from_numpy(z["intrinsics"]), torch.from_numpy(z["near"]),
                     torch.from_numpy(z["far"]), torch.from_numpy(z["bg"])[None].expand(V, 3))
cam_d = pack_cameras_device(t("extrinsics"), t("intrinsics"), t("near"), t("far"), bg).cpu()
print("cam max abs diff", (cam_h - cam_d).abs().max().item())
d = (cam_h - cam_d).abs()
print("worst fields", d.max(0).values.nonzero().flatten().tolist())
print(cam_h[0, :48]); print(cam_d[0, :48])
for name, cams in (("host", cam_h.cuda()), ("device", cam_d.cuda())):
    cfg = RenderConfig(H=H, W=W, V=V, sh_degree=4, sh_layout=SH_G3M, cov_layout=COV_FULL9, depth_mode=0)
    color, _, _ = rasterize(t("means")[None], t("cov")[None], t("harmonics")[None], t("opacities")[None], cams, cfg)
    err = (color.cpu().numpy() - z["color"])
    print(name, "max err", np.abs(err).max(), "per-row max", np.abs(err).max(axis=(0, 1, 3)).round(4).tolist())
