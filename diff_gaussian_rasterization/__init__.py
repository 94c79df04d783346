"""Drop-in for the reference's external CUDA extension of the same name
(/root/reference/requirements.txt:17; imported at src/model/decoder/cuda_splatting.py:6-9).
Put this repository on PYTHONPATH and the unmodified reference decoder renders through
libtsplat.so (sm_100a)."""
from transplat_b200.rasterizer import GaussianRasterizationSettings, GaussianRasterizer

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer"]
