"""Export of the Gaussians to a .ply file (the data format on the far side of the path), same
signature, scene normalisation, axis convention and attribute table as the reference's
``export_ply`` (/root/reference/src/model/ply_export.py:28-103).  The reference writes through the ``plyfile``
package (``PlyData([PlyElement.describe(elements, "vertex")]).write(path)``: binary, native byte order); that
package is not a dependency here - the same header and the same packed float32 rows are written directly.
Host code: ``scales`` / ``rotations`` are what ``GaussianAdapter.forward`` returns next to the render inputs.
"""
from __future__ import annotations

from pathlib import Path

import numpy as np
import torch
from scipy.spatial.transform import Rotation as R
from torch import Tensor


def construct_list_of_attributes(num_rest: int) -> list:
    """ply_export.py:12-25"""
    attributes = ["x", "y", "z", "nx", "ny", "nz"]
    for i in range(3):
        attributes.append(f"f_dc_{i}")
    for i in range(num_rest):
        attributes.append(f"f_rest_{i}")
    attributes.append("opacity")
    for i in range(3):
        attributes.append(f"scale_{i}")
    for i in range(4):
        attributes.append(f"rot_{i}")
    return attributes


def ply_table(extrinsics: Tensor, means: Tensor, scales: Tensor, rotations: Tensor, harmonics: Tensor,
              opacities: Tensor) -> np.ndarray:
    """The [gaussian, 17] float32 table export_ply writes (ply_export.py:37-101)."""
    means = means - means.median(dim=0).values                       # :38
    scale_factor = means.abs().quantile(0.95, dim=0).max()           # :41
    means = means / scale_factor
    scales = scales / scale_factor
    rotation = torch.tensor([[0, 0, 1], [-1, 0, 0], [0, -1, 0]], dtype=torch.float32, device=means.device)   # :46-51
    adjustment = torch.tensor(R.from_rotvec([0, 0, -45], True).as_matrix(), dtype=torch.float32, device=means.device)
    rotation = adjustment @ rotation                                  # :56-61
    rotation = rotation @ extrinsics[:3, :3].inverse()               # :65
    means = torch.einsum("ij,...j->...i", rotation, means)           # :68
    rot = R.from_quat(rotations.detach().cpu().numpy()).as_matrix()  # :71-75 (xyzw in, wxyz out)
    rot = rotation.detach().cpu().numpy() @ rot
    x, y, z, w = R.from_matrix(rot).as_quat().T
    rot = np.stack((w, x, y, z), axis=-1)
    dc = harmonics[..., 0]                                            # :79 (DC band only)
    cols = (means.detach().cpu().numpy(), torch.zeros_like(means).detach().cpu().numpy(),
            dc.detach().cpu().contiguous().numpy(), opacities[..., None].detach().cpu().numpy(),
            scales.log().detach().cpu().numpy(), rot)
    return np.concatenate(cols, axis=1).astype(np.float32)


def export_ply(extrinsics: Tensor, means: Tensor, scales: Tensor, rotations: Tensor, harmonics: Tensor, opacities: Tensor,
               path: Path) -> None:
    table = ply_table(extrinsics, means, scales, rotations, harmonics, opacities)
    names = construct_list_of_attributes(0)
    assert table.shape[1] == len(names)
    path = Path(path)
    path.parent.mkdir(exist_ok=True, parents=True)
    header = "ply\nformat binary_little_endian 1.0\n" + f"element vertex {table.shape[0]}\n" + \
             "".join(f"property float {n}\n" for n in names) + "end_header\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(np.ascontiguousarray(table, dtype="<f4").tobytes())


def read_ply_table(path: Path):
    """Minimal reader for the files written above: (attribute names, [n, len(names)] float32)."""
    raw = Path(path).read_bytes()
    end = raw.index(b"end_header\n") + len(b"end_header\n")
    lines = raw[:end].decode("ascii").splitlines()
    n = int(next(l for l in lines if l.startswith("element vertex")).split()[-1])
    names = [l.split()[-1] for l in lines if l.startswith("property float")]
    return names, np.frombuffer(raw[end:], dtype="<f4").reshape(n, len(names))
