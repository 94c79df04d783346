"""tsplat_forward_noalloc (include/tsplat.h; csrc/noalloc.cu): the forward pass over the library's static device
workspace, for callers that cannot allocate.  No counterpart in the reference,
so the check is against the existing entry point, tsplat_forward, on the same inputs, through the C-ABI.

Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C
import dataclasses

import pytest
import torch

from cases import case
from parity import COLOR_ATOL
from transplat_b200 import _lib
from transplat_b200.rasterize import forward_raw, read_status

pytestmark = pytest.mark.gpu


def _noalloc(prob, d_cap, N, G, cfg, dev):
    """One call of the new entry point through the C-ABI; outputs are caller-owned, nothing else is passed."""
    L = _lib.lib()
    color = torch.full((N, 3, cfg.H, cfg.W), float("nan"), dtype=torch.float32, device=dev)
    depth = torch.full((N, cfg.H, cfg.W), float("nan"), dtype=torch.float32, device=dev) if cfg.depth_mode else None
    radii = torch.full((N, G), -7, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev)
    rc = L.tsplat_forward_noalloc(C.byref(prob), d_cap, C.c_void_p(color.data_ptr()),
                                  C.c_void_p(depth.data_ptr()) if depth is not None else None,
                                  C.c_void_p(radii.data_ptr()), C.c_void_p(stream.cuda_stream))
    _lib.check(rc, "tsplat_forward_noalloc")
    return color, depth, radii


@pytest.mark.parametrize("names", [("small_l2_depth", "ragged_l2"), ("ragged_l2", "small_l2_depth")])
def test_forward_noalloc_matches_forward_twice_in_a_row(names):
    """Two calls in a row on DIFFERENT problems (the second finds the first one's contents in the shared static
    workspace), each compared with tsplat_forward on a fresh caller-owned workspace; then the first problem again."""
    dev = torch.device("cuda", 0)
    with torch.cuda.device(dev):
        for name in (*names, names[0]):
            means, cov, sh, op, cams, cfg = case(name)
            cfg = dataclasses.replace(cfg, want_radii=True)
            t = [x.to(dev) for x in (means, cov, sh, op, cams)]
            # the pair-list capacity is fixed HERE from the counts of a first call: the process-wide calibration cache
            # depends on which tests ran before (the overflow tests grow it) and with it the workspace size
            _, _, _, probe = forward_raw(cfg, *t)
            st = read_status(probe)
            assert not st.overflow and st.num_rendered > 0
            d_cap = int(max(st.num_rendered, (st.num_block_entries + 1) // _lib.BLOCK_ENTRY_RATIO)) + 1024
            cfg = dataclasses.replace(cfg, d_cap=d_cap)
            color, depth, radii, state = forward_raw(cfg, *t)
            st = read_status(state)
            assert not st.overflow and state.d_cap == d_cap
            assert state.layout.total_bytes <= _lib.NOALLOC_WORKSPACE_BYTES, "pick a smaller case for this test"
            N, G = color.shape[0], t[0].shape[1]
            c2, d2, r2 = _noalloc(state.problem, state.d_cap, N, G, cfg, dev)
            torch.cuda.synchronize(dev)
            assert torch.isfinite(c2).all()
            assert float((c2 - color).abs().max()) <= COLOR_ATOL
            if depth is not None:
                assert float((d2 - depth).abs().max()) <= COLOR_ATOL
            assert torch.equal(r2, radii)


def test_forward_noalloc_rejects_what_does_not_fit():
    """A problem whose workspace exceeds the static array is refused with a clear error and nothing is launched."""
    dev = torch.device("cuda", 0)
    L = _lib.lib()
    means, cov, sh, op, cams, cfg = case("small_l2_depth")
    t = [x.to(dev) for x in (means, cov, sh, op, cams)]
    with torch.cuda.device(dev):
        _, _, _, state = forward_raw(cfg, *t)
        read_status(state)
        torch.cuda.synchronize(dev)
        N, G = state.problem.B * state.problem.V, state.problem.G
        color = torch.zeros((N, 3, cfg.H, cfg.W), dtype=torch.float32, device=dev)
        depth = torch.zeros((N, cfg.H, cfg.W), dtype=torch.float32, device=dev)
        big = 4 << 20  # pairs: 62 B each -> far beyond 32 MiB
        rc = L.tsplat_forward_noalloc(C.byref(state.problem), big, C.c_void_p(color.data_ptr()),
                                      C.c_void_p(depth.data_ptr()), None,
                                      C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
        assert rc == _lib.TSPLAT_ENOSPC
        assert b"TSPLAT_NOALLOC_WORKSPACE_BYTES" in L.tsplat_last_error()
        torch.cuda.synchronize(dev)
        assert float(color.abs().max()) == 0.0  # untouched
