"""CPU-side checks: the C-ABI library loads and exports every declared symbol, workspace layout
arithmetic, host camera prep vs the golden fixture (reference host code), error behaviour that
needs no GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

import oracle
from transplat_b200 import _lib
from transplat_b200._lib import COV_FULL9, SH_G3M
from transplat_b200.cameras import pack_cameras
from transplat_b200.rasterize import RenderConfig

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "ref_host_small.npz")


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "tsplat.h")).read()
    declared = set(re.findall(r"^(?:int|const char\*)\s+(tsplat_\w+)\(", hdr, flags=re.M))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    L = _lib.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert L.tsplat_version() == 200


def _problem(B=1, G=131072, V=4, H=256, W=256):
    p = _lib.TsplatProblem()
    p.B, p.G, p.V, p.H, p.W, p.sh_degree, p.sh_coeffs = B, G, V, H, W, 4, 25
    return p


def test_workspace_layout_c2():
    L = _lib.lib()
    lay = _lib.TsplatLayout()
    assert L.tsplat_workspace_layout(C.byref(_problem()), 8 * 4 * 131072, C.byref(lay)) == 0
    assert (lay.tiles_x, lay.tiles_y) == (16, 16)
    assert lay.sort_bits == 32 + 10 and lay.sort_passes == 6   # 4 views x 256 tiles = 2^10 (view, tile) ids
    offs = [getattr(lay, n) for n, _ in _lib.TsplatLayout._fields_[1:15]]
    # xy / conic_opacity / rgbd are interleaved in one 48-byte record per (view, Gaussian)
    assert lay.rec_stride == 48 and lay.conic_opacity == lay.xy + 8 and lay.rgbd == lay.xy + 24
    aligned = [o for o in offs if o not in (lay.conic_opacity, lay.rgbd)]
    assert all(o % 256 == 0 for o in aligned) and len(set(offs)) == len(offs)
    assert lay.status == 0 and max(offs) < lay.total_bytes


def test_layout_rejects_bad_arguments():
    L = _lib.lib()
    lay = _lib.TsplatLayout()
    p = _problem()
    p.sh_degree = 5
    assert L.tsplat_workspace_layout(C.byref(p), 1000, C.byref(lay)) == _lib.TSPLAT_EINVAL
    assert b"sh_degree" in L.tsplat_last_error()
    p = _problem()
    p.sh_coeffs = 9  # degree 4 needs 25
    assert L.tsplat_workspace_layout(C.byref(p), 1000, C.byref(lay)) == _lib.TSPLAT_EINVAL
    assert L.tsplat_workspace_layout(C.byref(_problem()), 0, C.byref(lay)) == _lib.TSPLAT_EINVAL
    assert L.tsplat_workspace_layout(C.byref(_problem()), 1 << 30, C.byref(lay)) == _lib.TSPLAT_EINVAL


def test_nvls_exchange_rejects_bad_arguments():
    """tsplat_nvls_all_reduce (csrc/exchange.cu) validates before it touches the device: null / unaligned slab, more than
    four pieces, a rank outside the world, pieces that are not whole 16-byte units.  (Its arithmetic is checked on 2, 4
    and 8 GPUs by scripts/dist_gpu_check.py against the whole-scene gradients: profiles/*view_sharding_check*.)"""
    L = _lib.lib()
    offs, cnts = (C.c_uint64 * 4)(0, 0, 0, 0), (C.c_uint64 * 4)(16, 0, 0, 0)
    slab = C.c_void_p(0x7f0000000000)  # never dereferenced on the host
    assert L.tsplat_nvls_all_reduce(None, offs, cnts, 1, 0, 2, 0, None) == _lib.TSPLAT_EINVAL
    assert L.tsplat_nvls_all_reduce(C.c_void_p(0x7f0000000004), offs, cnts, 1, 0, 2, 0, None) == _lib.TSPLAT_EINVAL
    assert L.tsplat_nvls_all_reduce(slab, offs, cnts, 5, 0, 2, 0, None) == _lib.TSPLAT_EINVAL
    assert L.tsplat_nvls_all_reduce(slab, offs, cnts, 1, 2, 2, 0, None) == _lib.TSPLAT_EINVAL
    assert b"rank outside the world" in L.tsplat_last_error()
    cnts[0] = 18
    assert L.tsplat_nvls_all_reduce(slab, offs, cnts, 1, 0, 2, 0, None) == _lib.TSPLAT_EINVAL
    assert b"16-byte units" in L.tsplat_last_error()
    cnts[0] = 0
    assert L.tsplat_nvls_all_reduce(slab, offs, cnts, 1, 0, 2, 0, None) == _lib.TSPLAT_OK  # nothing to do: no launch


def test_no_cpu_fallback():
    from transplat_b200.decoder import DecoderSplattingCUDA, Gaussians
    dec = DecoderSplattingCUDA()
    g = Gaussians(torch.zeros(1, 4, 3), torch.zeros(1, 4, 3, 3), torch.zeros(1, 4, 3, 25), torch.zeros(1, 4))
    e = torch.eye(4)[None, None]
    k = torch.tensor([[0.9, 0, 0.5], [0, 0.9, 0.5], [0, 0, 1.0]])[None, None]
    with pytest.raises(RuntimeError, match="CUDA device only"):
        dec.forward(g, e, k, torch.ones(1, 1), torch.full((1, 1), 100.0), (32, 32))


def test_rasterizer_argument_errors_match_reference():
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    s = GaussianRasterizationSettings(32, 32, 0.5, 0.5, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 0,
                                      torch.zeros(3), False, False)
    r = GaussianRasterizer(s)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(m, m, torch.zeros(4, 1), shs=None, colors_precomp=None, cov3D_precomp=torch.zeros(4, 6))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3))
    assert GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")


def test_host_camera_prep_plus_oracle_reproduces_reference_host_code():
    """Golden: UNMODIFIED reference render_cuda/render_depth_cuda over the oracle (tests/golden/
    make_golden.py).  Here: our pack_cameras + the oracle's fused path on the encoder layouts.
    Colour must be bit-identical (same matrices, same rasterizer arithmetic)."""
    z = np.load(GOLD)
    V = z["extrinsics"].shape[0]
    H, W = z["color"].shape[-2:]
    t = lambda k: torch.from_numpy(z[k])
    cams = pack_cameras(t("extrinsics"), t("intrinsics"), t("near"), t("far"), t("bg")[None].expand(V, 3)).numpy()
    sc = oracle.Scene(z["means"], z["cov"], z["harmonics"], z["opacities"], 4, oracle.SH_G3M, oracle.COV_FULL9)
    for mode in ("depth", "disparity", "relative_disparity", "log"):
        acc = oracle.GradAccum(sc)
        for v in range(V):
            cam = cams[v].copy()
            st = oracle.forward(sc, cam, H, W, depth_mode=mode)
            assert np.array_equal(st.color, z["color"][v])
            # the reference's depth pass computes z_cam with a torch einsum; ours with fixed-order fp32
            assert np.abs(st.depth - z[f"depth_{mode}"][v]).max() <= 1e-5  # tolerance: 1e-5 abs
            oracle.backward(sc, cam, st, z["w_color"][v], z["w_depth"][v], acc=acc, depth_mode=mode)
        ref_means = z["g_means"] + z[f"depth_{mode}_g_means"]
        ref_cov = z["g_cov"] + z[f"depth_{mode}_g_cov"]
        ref_op = z["g_op"] + z[f"depth_{mode}_g_op"]
        for a, r in ((acc.means, ref_means), (acc.cov, ref_cov), (acc.opacity, ref_op), (acc.sh, z["g_sh"])):
            assert np.abs(a - r).max() <= 1e-4 * np.abs(r).max()  # tolerance: 1e-4 relative


def test_ply_export_matches_reference_table(tmp_path):
    """transplat_b200.ply_export.export_ply against the vertex table of the UNMODIFIED reference export_ply
    (tests/golden/ref_ply_small.npz, tests/golden/make_golden_ply.py): same attributes, same float32 rows."""
    import numpy as np
    import torch
    from transplat_b200.ply_export import export_ply, read_ply_table
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_ply_small.npz"))
    t = lambda k: torch.from_numpy(z[k])
    path = tmp_path / "sub" / "g.ply"
    export_ply(t("extrinsics"), t("means"), t("scales"), t("rotations"), t("harmonics"), t("opacities"), path)
    names, table = read_ply_table(path)
    assert names == [str(n) for n in z["names"]]
    assert table.shape == z["table"].shape
    # quaternions are sign-ambiguous only through scipy's canonicalisation, which both sides share
    assert np.abs(table - z["table"]).max() <= 1e-5 * np.abs(z["table"]).max()


def test_unmodified_reference_file_marshals_like_the_batched_path(monkeypatch):
    """The UNMODIFIED reference file src/model/decoder/cuda_splatting.py, imported with this repository's
    `diff_gaussian_rasterization` package in place of the extension, reaches the library call (here intercepted: no
    GPU) with exactly the per-view camera record, scaled means / covariances and SH layout that the batched level-2
    path hands to the same library call for the same inputs.  (The GPU twin,
    tests/test_gpu_parity.py::test_unmodified_reference_file_through_the_shim, renders through it.)"""
    ref_root = os.environ.get("TSPLAT_REFERENCE_ROOT", "/root/reference")
    if not os.path.exists(os.path.join(ref_root, "src", "model", "decoder", "cuda_splatting.py")):
        pytest.skip("reference tree not present on this machine")
    import diff_gaussian_rasterization as shim
    import transplat_b200.rasterizer as level1
    from oracle.rasterizer_cpu import load_reference_cuda_splatting
    calls = []

    def fake_rasterize(means, cov, feat, opacity, cam, cfg, means2d=None):
        calls.append(dict(means=means.clone(), cov=cov.clone(), feat=feat.clone(), opacity=opacity.clone(),
                          cam=cam.clone(), cfg=cfg))
        return (torch.zeros(1, 3, cfg.H, cfg.W), None, torch.zeros(1, means.shape[1], dtype=torch.int32))

    monkeypatch.setattr(level1, "rasterize", fake_rasterize)
    ref = load_reference_cuda_splatting(shim, ref_root)
    z = np.load(GOLD)
    t = lambda k: torch.from_numpy(z[k])
    V = z["extrinsics"].shape[0]
    H, W = z["color"].shape[-2:]
    rep = lambda k: t(k)[None].expand(V, *z[k].shape)
    out = ref.render_cuda(t("extrinsics"), t("intrinsics"), t("near"), t("far"), (H, W), t("bg")[None].expand(V, 3),
                          rep("means"), rep("cov"), rep("harmonics"), rep("opacities"))
    assert tuple(out.shape) == (V, 3, H, W) and len(calls) == V
    cams = pack_cameras(t("extrinsics"), t("intrinsics"), t("near"), t("far"), t("bg")[None].expand(V, 3))
    row, col = torch.triu_indices(3, 3)
    for v, c in enumerate(calls):
        cfg = c["cfg"]
        assert (cfg.H, cfg.W, cfg.V, cfg.sh_degree, cfg.depth_mode, cfg.want_radii) == (H, W, 1, 4, 0, True)
        cam = c["cam"][0]
        # view / projection / camera position / tan(fov/2) / background: the values the batched packer produces
        assert torch.equal(cam[:40], cams[v, :40])
        assert float(cam[40]) == 1.0 and float(cam[41]) == 1.0  # level 1: the reference file has already scaled the scene
        s = cams[v, 40]
        assert torch.equal(c["means"][0], t("means") * s)
        assert torch.equal(c["cov"][0], (t("cov") * s ** 2)[:, row, col])
        assert torch.equal(c["feat"][0], t("harmonics").permute(0, 2, 1))   # [G, M, 3]
        assert torch.equal(c["opacity"][0], t("opacities"))


def test_numa_local_is_best_effort():
    """streaming.numa_local: the cpulist parser, and that the context manager never raises and restores the affinity
    (here there is no GPU: it must record the error and bind nothing)."""
    import os

    from transplat_b200.streaming import _parse_cpulist, numa_local
    assert _parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert _parse_cpulist("") == set()
    before = os.sched_getaffinity(0)
    with numa_local("cuda:0") as n:
        inside = os.sched_getaffinity(0)
    assert os.sched_getaffinity(0) == before
    assert n.info["bound"] is False or inside <= before


def test_capacity_bookkeeping_reads_every_status_block():
    """rasterize._Capacity with fake events and host blocks (no GPU): every in-flight status block is read exactly once,
    an overflowed earlier call raises at the sweep AND has already raised the capacity, a block the caller read itself
    is forgotten, capacities never shrink."""
    import numpy as np

    from transplat_b200.rasterize import OverflowError_, TsplatStatus, _Capacity

    class FakeEvent:
        def __init__(self, done):
            self.done, self.synced = done, 0

        def query(self):
            return self.done

        def synchronize(self):
            self.synced += 1
            self.done = True

    class FakeHost:
        def __init__(self, d, e, overflow):
            st = TsplatStatus(num_rendered=d, overflow=overflow, reserved=0, num_block_entries=e, reserved1=0)
            self.buf = np.frombuffer(bytes(st), dtype=np.uint8).copy()

        def numpy(self):
            return self.buf

    cap = _Capacity()
    key = (0, 4, 1000, 64, 64)
    assert cap.get(key, 4000) >= 4096
    first = cap.cap[key]
    ev_done, ev_pending, ev_mine = FakeEvent(True), FakeEvent(False), FakeEvent(False)
    cap.inflight = [(key, FakeHost(10_000, 21_000, 0), ev_done, first, True),
                    (key, FakeHost(50_000, 100_000, 1), ev_pending, first, True),
                    (key, FakeHost(1, 1, 0), ev_mine, first, True)]
    cap.forget(ev_mine)                       # the caller read that one itself
    cap.sweep()                               # only the finished one is read; nothing raises
    assert len(cap.inflight) == 1 and cap.last_num_rendered == 10_000
    assert cap.cap[key] >= max(2 * 10_000, 21_000 + 1)
    grown = cap.cap[key]
    with pytest.raises(OverflowError_):
        cap.sweep(wait=True)                  # the overflowed call surfaces, once
    assert cap.inflight == [] and ev_pending.synced == 1
    assert cap.cap[key] >= max(2 * 50_000, 100_000 + 1) > grown   # and the capacity is already raised
    cap.observe(key, 10, 10)
    assert cap.cap[key] >= 100_000            # never shrinks
    cap.sweep(wait=True)                      # nothing left: no second raise


def test_view_sharding_exchange_policy():
    """parallel.choose_exchange: what forward_view_sharded does when the caller does not choose, and that explicit
    choices are honoured (no GPU, no process group)."""
    from transplat_b200.parallel import NVLS_CHUNKS, NVLS_MIN_WORLD, choose_exchange as ce
    C2, C5 = 131072, 2097152
    assert ce(1, 1, C2, True, None, False, None, True) == ("none", 0)
    # defaults: NCCL below NVLS_MIN_WORLD ranks, the NVLS exchange from there; large slabs range by range
    assert ce(2, 1, C5, True, None, False, None, True) == ("all_reduce", 0)
    assert ce(NVLS_MIN_WORLD, 1, C2, True, None, False, None, True) == ("nvls", 0)
    assert ce(8, 1, C5, True, None, False, None, True) == ("nvls", NVLS_CHUNKS)
    assert ce(8, 2, C5, True, None, False, None, True) == ("nvls", 0)            # several scenes per call: whole slab
    assert ce(8, 1, C5, True, None, False, None, False) == ("all_reduce", 0)      # no multicast memory on this system
    assert ce(8, 1, C5, False, None, False, None, True) == ("all_reduce", 0)      # CPU tensors (the gloo tests)
    # explicit choices
    assert ce(8, 1, C5, True, 0, False, False, True) == ("all_reduce", 0)
    assert ce(8, 1, C5, True, 3, False, False, True) == ("staged_all_reduce", 3)
    assert ce(8, 1, C5, True, 3, False, None, True) == ("staged_all_reduce", 3)   # chunks given, exchange not: NCCL, as before
    assert ce(2, 1, C2, True, 0, False, True, True) == ("nvls", 0)
    assert ce(2, 1, C2, True, 4, False, True, True) == ("nvls", 4)
    assert ce(8, 1, C2, True, None, True, None, True) == ("multimem", 0)
    assert ce(8, 2, C2, True, 3, False, False, True) == ("all_reduce", 0)         # staging needs one scene per call
