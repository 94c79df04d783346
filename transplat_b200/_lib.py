"""ctypes binding of ``libtsplat.so`` (include/tsplat.h).  There is no CPU fallback: if the
library is missing or cannot be loaded every entry point raises."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# TSPLAT_LIB_PATH: development only (kernel experiments built with `make OUT=... EXTRA=-D...`)
LIB_PATH = os.environ.get("TSPLAT_LIB_PATH") or os.path.join(_HERE, "lib", "libtsplat.so")
CSRC = os.path.join(_HERE, "csrc")

TSPLAT_OK, TSPLAT_EINVAL, TSPLAT_ENOSPC, TSPLAT_ECUDA = 0, 1, 2, 3
NOALLOC_WORKSPACE_BYTES = 32 << 20  # include/tsplat.h: TSPLAT_NOALLOC_WORKSPACE_BYTES
SH_GM3, SH_G3M = 0, 1
COV_TRIU6, COV_FULL9 = 0, 1
FLAG_REZERO_GRAD_ACC = 1
FLAG_SORT_ONESWEEP = 2
FLAG_DEBUG = 4
FLAG_BWD_RENDER_ONLY = 8
FLAG_BWD_GAUSSIANS_ONLY = 16
FLAG_BWD_MULTIMEM = 32
DEPTH_MODES = {None: 0, "depth": 1, "disparity": 2, "relative_disparity": 3, "log": 4}

EXPORTS = ("tsplat_version", "tsplat_last_error", "tsplat_workspace_layout", "tsplat_forward",
           "tsplat_backward", "tsplat_mark_visible", "tsplat_pack_cameras", "tsplat_status_read", "tsplat_profile_enable",
           "tsplat_profile_read", "tsplat_adapter_forward", "tsplat_adapter_backward", "tsplat_forward_mse",
           "tsplat_backward_mse", "tsplat_cov3d_forward", "tsplat_cov3d_backward", "tsplat_nvls_all_reduce", "tsplat_forward_noalloc")


class TsplatProblem(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "B", "G", "V", "H", "W", "sh_degree", "sh_coeffs", "sh_layout", "cov_layout", "depth_mode",
        "sh_l4", "flags")] + [(n, C.c_void_p) for n in ("means", "cov", "sh", "opacity", "cams")] + [
        (n, C.c_int32) for n in ("g_begin", "g_count")]


class TsplatLayout(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "total_bytes", "status", "depths", "radii", "xy", "conic_opacity", "rgbd", "tiles_touched",
        "clamped", "keys", "vals", "ranges", "final_T", "n_contrib", "grad_acc")] + [
        (n, C.c_int32) for n in ("tiles_x", "tiles_y", "sort_bits", "sort_passes", "rec_stride", "reserved0")] + [
        (n, C.c_uint64) for n in ("block_jobs", "block_refs", "e_cap")]


class TsplatMse(C.Structure):
    _fields_ = [("target", C.c_void_p), ("loss_sum", C.c_void_p), ("color", C.c_void_p), ("dL_dloss", C.c_void_p),
                ("scale", C.c_float), ("loss_scale", C.c_float), ("psnr", C.c_void_p)]


class TsplatStatus(C.Structure):
    _fields_ = [("num_rendered", C.c_uint64), ("overflow", C.c_uint32), ("reserved", C.c_uint32),
                ("num_block_entries", C.c_uint64), ("reserved1", C.c_uint64)]


BLOCK_ENTRY_RATIO = 2  # TSPLAT_BLOCK_ENTRY_RATIO: capacity of the per-block lists = 2 * d_cap entries


def build(verbose: bool = False) -> str:
    """Compile every kernel for sm_100a with nvcc (cross-compiles without a GPU)."""
    out = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("building libtsplat.so failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stdout)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback for the render path)")
        L = C.CDLL(LIB_PATH)
        L.tsplat_version.restype = C.c_int
        L.tsplat_last_error.restype = C.c_char_p
        L.tsplat_workspace_layout.argtypes = [C.POINTER(TsplatProblem), C.c_uint64, C.POINTER(TsplatLayout)]
        L.tsplat_forward.argtypes = [C.POINTER(TsplatProblem), C.c_void_p, C.c_uint64, C.c_uint64,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.tsplat_backward.argtypes = [C.POINTER(TsplatProblem), C.c_void_p, C.c_uint64, C.c_uint64,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p]
        L.tsplat_mark_visible.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.tsplat_pack_cameras.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
        L.tsplat_status_read.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.tsplat_forward_mse.argtypes = [C.POINTER(TsplatProblem), C.c_void_p, C.c_uint64, C.c_uint64,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(TsplatMse), C.c_void_p]
        L.tsplat_backward_mse.argtypes = [C.POINTER(TsplatProblem), C.c_void_p, C.c_uint64, C.c_uint64,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.POINTER(TsplatMse), C.c_void_p]
        L.tsplat_adapter_forward.argtypes = [C.c_void_p] * 7
        L.tsplat_adapter_backward.argtypes = [C.c_void_p] * 8
        L.tsplat_forward_noalloc.argtypes = [C.POINTER(TsplatProblem), C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_void_p]
        L.tsplat_nvls_all_reduce.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                                             C.c_int32, C.c_void_p]
        L.tsplat_cov3d_forward.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p]
        L.tsplat_cov3d_backward.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p]
        L.tsplat_profile_enable.argtypes = [C.c_int]
        L.tsplat_profile_read.argtypes = [C.c_void_p, C.c_int]
        _lib = L
    return _lib


def check(rc: int, what: str) -> None:
    if rc != TSPLAT_OK:
        msg = lib().tsplat_last_error().decode()
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


STAGES = ("zero", "preprocess", "bin_emit", "sort", "ranges_gather", "render_forward", "zero_bwd",
          "render_backward", "preprocess_backward")


def profile_enable(on: bool) -> None:
    check(lib().tsplat_profile_enable(1 if on else 0), "tsplat_profile_enable")


def profile_read() -> dict:
    """Per-stage milliseconds averaged over the recorded calls (synchronises on their events)."""
    buf = (C.c_float * (len(STAGES) + 1))()
    check(lib().tsplat_profile_read(buf, len(STAGES) + 1), "tsplat_profile_read")
    calls = int(buf[len(STAGES)])
    out = {k: (float(buf[i]) / calls if calls else 0.0) for i, k in enumerate(STAGES)}
    out["calls"] = calls
    return out
