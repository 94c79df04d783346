"""CPU oracle for the Gaussian-splat render path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package; ``transplat_b200`` never does.

PARITY UNPINNED: the arithmetic restated in ``tsplat_oracle.c`` lives in the external extension
``diff_gaussian_rasterization`` (/root/reference/requirements.txt:17, unpinned, source absent);
see the header of ``tsplat_oracle.c`` and SURVEY.md 8c / Appendix A.

This module is a thin ctypes/numpy binding of ``oracle/_build/libtsplat_oracle.so`` plus a
stage-by-stage driver that mirrors ``CudaRasterizer::Rasterizer::forward/backward``
(SURVEY 8a X12): preprocess -> scan -> duplicateWithKeys -> sort -> identifyTileRanges -> render.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libtsplat_oracle.so")

CAM_FLOATS = 64
# float offsets inside TsplatCamera (include/tsplat.h)
O_VIEW, O_PROJ, O_CAMPOS, O_TANX, O_TANY, O_BG, O_MSCALE, O_CSCALE, O_DROW, O_NEAR, O_FAR = (
    0, 16, 32, 35, 36, 37, 40, 41, 42, 46, 47)

SH_GM3, SH_G3M = 0, 1
COV_TRIU6, COV_FULL9 = 0, 1
DEPTH_MODES = {None: 0, "depth": 1, "disparity": 2, "relative_disparity": 3, "log": 4}


def build(force: bool = False) -> str:
    """Compile the C restatement with the committed Makefile (gcc, -ffp-contract=off)."""
    src = os.path.join(_HERE, "tsplat_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


class _View(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "G", "H", "W", "sh_degree", "sh_coeffs", "sh_layout", "cov_layout", "depth_mode", "sh_l4")
    ] + [(n, C.c_void_p) for n in ("means", "cov", "sh", "opacity", "cam")]


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.oracle_scan.restype = C.c_uint64
        _lib.oracle_higher_msb.restype = C.c_int
        _lib.oracle_num_threads.restype = C.c_int
    return _lib


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def set_num_threads(n: int) -> int:
    """OpenMP team size of the oracle (torch.distributed.run exports OMP_NUM_THREADS=1 to its ranks)."""
    lib().oracle_set_num_threads(C.c_int(int(n)))
    return num_threads()


def higher_msb(n: int) -> int:
    return int(lib().oracle_higher_msb(C.c_uint32(n)))


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def make_camera(viewmatrix, projmatrix, campos, tanfovx, tanfovy, bg=(0, 0, 0), mean_scale=1.0,
                cov_scale=None, depth_row=(0, 0, 1, 0), near=1.0, far=100.0) -> np.ndarray:
    """Pack one TsplatCamera.  viewmatrix/projmatrix are the [4,4] tensors the reference passes in
    GaussianRasterizationSettings (cuda_splatting.py:115-116); they are flattened row-major, which
    yields the m[4*c+r] convention because they are transposes."""
    cam = np.zeros(CAM_FLOATS, np.float32)
    cam[O_VIEW:O_VIEW + 16] = np.asarray(viewmatrix, np.float32).reshape(16)
    cam[O_PROJ:O_PROJ + 16] = np.asarray(projmatrix, np.float32).reshape(16)
    cam[O_CAMPOS:O_CAMPOS + 3] = np.asarray(campos, np.float32)
    cam[O_TANX] = tanfovx
    cam[O_TANY] = tanfovy
    cam[O_BG:O_BG + 3] = np.asarray(bg, np.float32)
    cam[O_MSCALE] = mean_scale
    cam[O_CSCALE] = np.float32(mean_scale) * np.float32(mean_scale) if cov_scale is None else cov_scale
    cam[O_DROW:O_DROW + 4] = np.asarray(depth_row, np.float32)
    cam[O_NEAR] = near
    cam[O_FAR] = far
    return cam


@dataclass
class ViewState:
    """Everything one forward pass of one view produces (GeometryState/BinningState/ImageState)."""
    G: int
    H: int
    W: int
    CH: int
    depths: np.ndarray = None
    radii: np.ndarray = None
    xy: np.ndarray = None
    conic_opacity: np.ndarray = None
    rgbd: np.ndarray = None
    tiles_touched: np.ndarray = None
    clamped: np.ndarray = None
    offsets: np.ndarray = None
    num_rendered: int = 0
    keys_unsorted: np.ndarray = None
    keys: np.ndarray = None
    vals: np.ndarray = None
    ranges: np.ndarray = None
    color: np.ndarray = None
    depth: np.ndarray = None
    final_T: np.ndarray = None
    n_contrib: np.ndarray = None
    sort_bits: int = 0
    extra: dict = field(default_factory=dict)


class Scene:
    """One scene's Gaussians as contiguous float32 numpy arrays plus layout flags."""

    def __init__(self, means, cov, sh, opacity, sh_degree, sh_layout=SH_GM3, cov_layout=COV_TRIU6,
                 sh_l4=1):
        self.means = np.ascontiguousarray(means, np.float32)
        self.cov = np.ascontiguousarray(cov, np.float32)
        self.sh = np.ascontiguousarray(sh, np.float32)
        self.opacity = np.ascontiguousarray(opacity, np.float32).reshape(-1)
        self.G = self.means.shape[0]
        self.sh_degree = int(sh_degree)
        self.sh_layout = sh_layout
        self.cov_layout = cov_layout
        self.sh_l4 = sh_l4
        if self.sh.ndim == 2:  # colors_precomp [G,3]
            self.M = 0
        else:
            self.M = self.sh.shape[1] if sh_layout == SH_GM3 else self.sh.shape[2]

    def view(self, cam: np.ndarray, H: int, W: int, depth_mode=None) -> _View:
        v = _View()
        v.G, v.H, v.W = self.G, H, W
        v.sh_degree, v.sh_coeffs, v.sh_layout = self.sh_degree, self.M, self.sh_layout
        v.cov_layout, v.sh_l4 = self.cov_layout, self.sh_l4
        v.depth_mode = DEPTH_MODES[depth_mode] if not isinstance(depth_mode, int) else depth_mode
        v.means, v.cov, v.sh, v.opacity = (
            self.means.ctypes.data, self.cov.ctypes.data, self.sh.ctypes.data, self.opacity.ctypes.data)
        v.cam = cam.ctypes.data
        v._keep = (cam,)
        return v


def forward(scene: Scene, cam: np.ndarray, H: int, W: int, depth_mode=None, render=True) -> ViewState:
    """CudaRasterizer::Rasterizer::forward for one view."""
    L = lib()
    cam = np.ascontiguousarray(cam, np.float32)
    v = scene.view(cam, H, W, depth_mode)
    G = scene.G
    CH = 4 if v.depth_mode != 0 else 3
    st = ViewState(G, H, W, CH)
    st.depths = np.zeros(G, np.float32)
    st.radii = np.zeros(G, np.int32)
    st.xy = np.zeros((G, 2), np.float32)
    st.conic_opacity = np.zeros((G, 4), np.float32)
    st.rgbd = np.zeros((G, 4), np.float32)
    st.tiles_touched = np.zeros(G, np.uint32)
    st.clamped = np.zeros(G, np.uint8)
    L.oracle_preprocess(C.byref(v), _p(st.depths), _p(st.radii), _p(st.xy), _p(st.conic_opacity),
                        _p(st.rgbd), _p(st.tiles_touched), _p(st.clamped))
    st.offsets = np.zeros(G, np.uint32)
    D = int(L.oracle_scan(G, _p(st.tiles_touched), _p(st.offsets)))
    st.num_rendered = D
    keys = np.zeros(max(D, 1), np.uint64)
    vals = np.zeros(max(D, 1), np.uint32)
    L.oracle_duplicate_keys(G, H, W, _p(st.xy), _p(st.depths), _p(st.radii), _p(st.offsets),
                            _p(keys), _p(vals))
    st.keys_unsorted = keys[:D].copy()
    gx, gy = (W + 15) // 16, (H + 15) // 16
    T = gx * gy
    st.sort_bits = 32 + higher_msb(T)
    L.oracle_sort_pairs(C.c_uint64(D), _p(keys), _p(vals), st.sort_bits)
    st.keys, st.vals = keys[:D], vals[:D]
    st.ranges = np.zeros((T, 2), np.uint32)
    L.oracle_tile_ranges(C.c_uint64(D), _p(keys), T, _p(st.ranges))
    if render:
        st.color = np.zeros((3, H, W), np.float32)
        st.depth = np.zeros((H, W), np.float32)
        st.final_T = np.zeros((H, W), np.float32)
        st.n_contrib = np.zeros((H, W), np.uint32)
        bg = np.ascontiguousarray(cam[O_BG:O_BG + 3])
        L.oracle_render_forward(H, W, CH, _p(st.ranges), _p(vals), _p(st.xy), _p(st.conic_opacity),
                                _p(st.rgbd), _p(bg), _p(st.color), _p(st.depth), _p(st.final_T),
                                _p(st.n_contrib))
    st.extra["vals_buf"] = vals
    return st


def borderline_pixels(st: ViewState, eps: float = 4e-7) -> np.ndarray:
    """bool [H,W]: pixels of this view whose result hangs on a float comparison within rounding distance of its
    threshold (alpha vs 1/255, T(1-alpha) vs 1e-4, power vs 0) - see oracle_render_borderline."""
    flag = np.zeros((st.H, st.W), np.uint8)
    lib().oracle_render_borderline(st.H, st.W, _p(st.ranges), _p(st.extra["vals_buf"]), _p(st.xy), _p(st.conic_opacity),
                                   C.c_float(eps), _p(flag))
    return flag.astype(bool)


def gaussians_of_pixels(st: ViewState, mask: np.ndarray) -> np.ndarray:
    """bool [G]: every Gaussian in the tile list of a pixel of ``mask``, up to that pixel's last contributor + 1
    (the entry whose decision may have flipped can be the one right behind the recorded last contributor)."""
    out = np.zeros(st.G, bool)
    gx = (st.W + 15) // 16
    for y, x in np.argwhere(mask):
        t = (y // 16) * gx + (x // 16)
        lo, hi = int(st.ranges[t, 0]), int(st.ranges[t, 1])
        out[st.vals[lo:hi]] = True
    return out


class GradAccum:
    """binary64 accumulators for one scene; several views add into the same object."""

    def __init__(self, scene: Scene):
        self.means = np.zeros((scene.G, 3), np.float64)
        self.cov = np.zeros(scene.cov.shape, np.float64)
        self.sh = np.zeros(scene.sh.shape, np.float64)
        self.opacity = np.zeros(scene.G, np.float64)
        self.mean2d = None  # last view's screen-space gradient [G,3] float32
        self.screen = None  # last view's raw accumulators [G,10] float64


def backward(scene: Scene, cam: np.ndarray, st: ViewState, dL_dcolor, dL_ddepth=None,
             acc: GradAccum | None = None, depth_mode=None) -> GradAccum:
    """CudaRasterizer::Rasterizer::backward for one view; adds into ``acc``."""
    L = lib()
    cam = np.ascontiguousarray(cam, np.float32)
    v = scene.view(cam, st.H, st.W, depth_mode)
    assert (4 if v.depth_mode != 0 else 3) == st.CH
    acc = acc or GradAccum(scene)
    dL_dcolor = np.ascontiguousarray(dL_dcolor, np.float32)
    dL_ddepth = (np.zeros((st.H, st.W), np.float32) if dL_ddepth is None
                 else np.ascontiguousarray(dL_ddepth, np.float32))
    screen = np.zeros((scene.G, 10), np.float64)
    bg = np.ascontiguousarray(cam[O_BG:O_BG + 3])
    L.oracle_render_backward(st.H, st.W, st.CH, _p(st.ranges), _p(st.extra["vals_buf"]), _p(st.xy),
                             _p(st.conic_opacity), _p(st.rgbd), _p(bg), _p(st.final_T),
                             _p(st.n_contrib), _p(dL_dcolor), _p(dL_ddepth), _p(screen))
    mean2d = np.zeros((scene.G, 3), np.float32)
    L.oracle_preprocess_backward(C.byref(v), _p(st.radii), _p(st.clamped), _p(screen), _p(acc.means),
                                 _p(acc.cov), _p(acc.sh), _p(acc.opacity), _p(mean2d))
    acc.mean2d, acc.screen = mean2d, screen
    return acc


def mark_visible(means, viewmatrix) -> np.ndarray:
    means = np.ascontiguousarray(means, np.float32)
    vm = np.ascontiguousarray(np.asarray(viewmatrix, np.float32).reshape(16))
    out = np.zeros(means.shape[0], np.uint8)
    lib().oracle_mark_visible(means.shape[0], _p(means), _p(vm), _p(out))
    return out.astype(bool)
