"""One process, two GPUs (development check, run on a box with >= 2 GPUs): the same problem through libtsplat.so on
cuda:1 and on cuda:0 must give identical results - per-device state of the library (dynamic shared-memory attributes,
persistent grid sizes, the static workspace of tsplat_forward_noalloc) is keyed by device."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from cases import case
from transplat_b200 import _lib
from transplat_b200.rasterize import backward_raw, forward_raw, read_status

assert torch.cuda.device_count() >= 2, "needs two GPUs"
ok = True
for name in ("small_l2_depth", "mid_l2", "tiny_l1"):
    res = {}
    for d in (1, 0, 1):
        dev = torch.device("cuda", d)
        means, cov, sh, op, cams, cfg = case(name)
        t = [x.to(dev) for x in (means, cov, sh, op, cams)]
        with torch.cuda.device(dev):
            color, depth, radii, state = forward_raw(cfg, *t)
            assert not read_status(state).overflow
            g = torch.Generator(device="cpu").manual_seed(3)
            gc = torch.randn(color.shape, generator=g).to(dev)
            gd = torch.randn(depth.shape, generator=g).to(dev) if depth is not None else None
            grads = backward_raw(cfg, state, gc, gd)
            # the static-workspace entry point on this device too
            c2 = torch.empty_like(color)
            d2 = torch.empty_like(depth) if depth is not None else None
            rc = _lib.lib().tsplat_forward_noalloc(C.byref(state.problem), state.d_cap, C.c_void_p(c2.data_ptr()),
                                                  C.c_void_p(d2.data_ptr()) if d2 is not None else None, None,
                                                  C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
            if state.layout.total_bytes <= _lib.NOALLOC_WORKSPACE_BYTES:
                _lib.check(rc, "tsplat_forward_noalloc")
                torch.cuda.synchronize(dev)
                same = torch.equal(c2, color)
                ok &= same
                print(name, "cuda:%d" % d, "noalloc == forward:", same)
            torch.cuda.synchronize(dev)
        cur = [color.cpu()] + [x.cpu() for x in grads[:4]]
        if d in res:
            pass
        res.setdefault(d, cur)
    for a, b in zip(res[0], res[1]):
        # gradients are accumulated with floating-point atomics: equal up to summation order
        close = torch.allclose(a, b, rtol=1e-4, atol=1e-6 * float(a.abs().max()))
        ok &= close
    print(name, "cuda:1 vs cuda:0:", "OK" if ok else "FAIL", float((res[0][0] - res[1][0]).abs().max()))
print("ALL OK" if ok else "SOME FAILED")
