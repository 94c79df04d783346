"""Generates tests/golden/ref_ply_small.npz IN THE AUTHORING CONTAINER: the vertex table the UNMODIFIED reference
export_ply (src/model/ply_export.py:28-103) hands to plyfile, captured by a stand-in ``plyfile`` module (the real
package is absent; only its two entry points are stubbed, the table is produced by the reference code).
    python tests/golden/make_golden_ply.py
"""
import importlib.util, os, sys, types
from pathlib import Path
import numpy as np
import torch

captured = {}
pf = types.ModuleType("plyfile")
class PlyElement:
    @staticmethod
    def describe(elements, name):
        captured["elements"], captured["name"] = elements, name
        return elements
class PlyData:
    def __init__(self, els): pass
    def write(self, path): captured["path"] = str(path)
pf.PlyElement, pf.PlyData = PlyElement, PlyData
sys.modules["plyfile"] = pf
jt = types.ModuleType("jaxtyping")
jt.Float = type("Float", (), {"__class_getitem__": classmethod(lambda cls, item: cls)})
sys.modules["jaxtyping"] = jt
spec = importlib.util.spec_from_file_location("ref_ply_export", "/root/reference/src/model/ply_export.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

g = torch.Generator().manual_seed(9)
n = 500
q, _ = torch.linalg.qr(torch.randn(3, 3, generator=g))
ext = torch.eye(4); ext[:3, :3] = q * torch.linalg.det(q).sign(); ext[:3, 3] = torch.randn(3, generator=g)
means = torch.randn(n, 3, generator=g) * 3
scales = torch.rand(n, 3, generator=g) * 0.2 + 0.01
rot = torch.randn(n, 4, generator=g); rot = rot / rot.norm(dim=-1, keepdim=True)
harm = torch.randn(n, 3, 25, generator=g)
opac = torch.rand(n, generator=g)
ref.export_ply(ext, means, scales, rot, harm, opac, Path("/tmp/ref_ply/out.ply"))
el = captured["elements"]
table = np.stack([el[name] for name in el.dtype.names], axis=1)
out = dict(extrinsics=ext.numpy(), means=means.numpy(), scales=scales.numpy(), rotations=rot.numpy(), harmonics=harm.numpy(),
           opacities=opac.numpy(), table=table, names=np.array(el.dtype.names))
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_ply_small.npz")
np.savez_compressed(path, **out)
print("wrote", path, table.shape, el.dtype.names[:4], captured["name"])
