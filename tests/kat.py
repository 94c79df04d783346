"""Known-answer cases for the Gaussian-splat rasterizer (SURVEY.md section 4 / Appendix A.8 probes P1-P8).

Every expectation in this file is a CLOSED FORM written in binary64 numpy from the published definition of the
algorithm - none of it calls ``oracle/`` or the kernels, and the real spherical harmonics
come from associated Legendre functions (scipy), not from the polynomial table the oracle and the kernels share.
The same cases are run through the CPU oracle (tests/test_oracle_kat.py, ``-m "not gpu"``) and through libtsplat.so
(tests/test_gpu_kat.py, ``-m gpu``, both binning paths), so the oracle and the CUDA path are held to the same answers.

A *runner* is any callable
    run(means[G,3], cov6[G,6], feat, opacity[G], cam[64], H, W, sh_degree, depth_mode=0, grads=None) -> dict
with ``feat`` = SH coefficients [G,M,3] or precomputed colours [G,3], ``grads`` = (dL/dcolor[3,H,W], dL/ddepth[H,W]|None),
returning numpy arrays: color, depth, radii, tiles_touched, depths, xy, keys (uint64, one view), vals, ranges [T,2],
n_contrib, final_T, num_rendered and, with ``grads``, g_means, g_cov, g_feat, g_opacity.

Reference anchors: /root/reference/src/model/decoder/cuda_splatting.py:108-133 (the call), :375-417 (depth modes),
/root/reference/src/scripts/test_splatter.py:22-84 (the one-Gaussian-in-front-of-a-camera recipe).
"""
from __future__ import annotations

import math

import numpy as np

COLOR_ATOL = 1e-5  # north_star: colour/depth within 1e-5 abs
GRAD_RTOL = 1e-4   # north_star: gradients within 1e-4 relative
SH_C0 = 0.28209479177387814
CAM_FLOATS = 64
O_VIEW, O_PROJ, O_CAMPOS, O_TANX, O_TANY, O_BG, O_MSCALE, O_CSCALE, O_DROW, O_NEAR, O_FAR = (
    0, 16, 32, 35, 36, 37, 40, 41, 42, 46, 47)
f32 = np.float32


# ------------------------------------------------------------------------------------------------
# cameras, written from the definitions (cuda_splatting.py:26-53, 93-96), not through the product's packer
# ------------------------------------------------------------------------------------------------
def look_at(eye, target):
    """World->camera rotation R (rows right, down, forward) with the camera at ``eye`` looking at ``target``."""
    eye, target = np.asarray(eye, np.float64), np.asarray(target, np.float64)
    fwd = target - eye
    fwd /= np.linalg.norm(fwd)
    up = np.array([0.0, 1.0, 0.0]) if abs(fwd[1]) < 0.9 else np.array([1.0, 0.0, 0.0])
    right = np.cross(up, fwd)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    return np.stack((right, down, fwd))


def camera(tanx=0.5, tany=0.5, eye=(0, 0, 0), R=None, near=1.0, far=100.0, bg=(0, 0, 0), scale=1.0):
    """One TsplatCamera record (include/tsplat.h).  ``scale`` is the reference's scale-invariance factor 1/near
    (cuda_splatting.py:73-80): means are multiplied by it inside the rasterizer, the camera position and near/far
    of the projection are pre-multiplied here exactly as render_cuda does."""
    R = np.eye(3) if R is None else np.asarray(R, np.float64)
    eye = np.asarray(eye, np.float64)
    V = np.eye(4)
    V[:3, :3] = R
    V[:3, 3] = -R @ (eye * scale)
    n, f = near * scale, far * scale
    P = np.zeros((4, 4))
    P[0, 0], P[1, 1] = 1.0 / tanx, 1.0 / tany
    P[3, 2] = 1.0
    P[2, 2] = f / (f - n)
    P[2, 3] = -(f * n) / (f - n)
    Vu = np.eye(4)
    Vu[:3, :3] = R
    Vu[:3, 3] = -R @ eye
    cam = np.zeros(CAM_FLOATS, f32)
    cam[O_VIEW:O_VIEW + 16] = V.T.reshape(16)          # flat index 4*c + r
    cam[O_PROJ:O_PROJ + 16] = (P @ V).T.reshape(16)
    cam[O_CAMPOS:O_CAMPOS + 3] = eye * scale
    cam[O_TANX], cam[O_TANY] = tanx, tany
    cam[O_BG:O_BG + 3] = bg
    cam[O_MSCALE], cam[O_CSCALE] = scale, f32(scale) * f32(scale)
    cam[O_DROW:O_DROW + 4] = Vu[2]
    cam[O_NEAR], cam[O_FAR] = near, far
    return cam


def iso_cov6(s):
    s = np.atleast_1d(np.asarray(s, np.float64))
    c = np.zeros((s.shape[0], 6))
    c[:, 0] = c[:, 3] = c[:, 5] = s * s
    return c.astype(f32)


# ------------------------------------------------------------------------------------------------
# closed forms
# ------------------------------------------------------------------------------------------------
def pixel_grid(H, W):
    ys, xs = np.meshgrid(np.arange(H, dtype=np.float64), np.arange(W, dtype=np.float64), indexing="ij")
    return xs, ys


def iso_alpha(H, W, cx, cy, sigma2, opacity):
    """alpha of an isotropic screen-space Gaussian with variance sigma2 (dilation included) at every pixel, with
    the reference's rules: cap 0.99, contributions below 1/255 dropped (A.3)."""
    xs, ys = pixel_grid(H, W)
    r2 = (xs - cx) ** 2 + (ys - cy) ** 2
    a = np.minimum(0.99, opacity * np.exp(-0.5 * r2 / sigma2))
    raw = opacity * np.exp(-0.5 * r2 / sigma2)
    return np.where(a >= 1.0 / 255.0, a, 0.0), raw


def near_threshold(raw, thr=1.0 / 255.0, rel=2e-5):
    """Pixels whose binary32 alpha may fall on the other side of a decision threshold."""
    return np.abs(raw - thr) <= rel * thr


def radius_of(sigma2_a, sigma2_c=None, b=0.0):
    """ceil(3 sqrt(lambda_max)) with the eigenvalue floor max(0.1, mid^2 - det) (A.3)."""
    a = sigma2_a
    c = sigma2_a if sigma2_c is None else sigma2_c
    det = a * c - b * b
    mid = 0.5 * (a + c)
    lam = mid + math.sqrt(max(0.1, mid * mid - det))
    r = 3.0 * math.sqrt(lam)
    assert abs(r - round(r)) > 1e-3, "known-answer case sits on a rounding boundary of ceil(): move it"
    return int(math.ceil(r))


def tile_rect(px, py, radius, H, W):
    """getRect (A.2): C truncation of the float quotient, then clamped to the tile grid."""
    gx, gy = (W + 15) // 16, (H + 15) // 16
    clamp = lambda v, hi: min(hi, max(0, v))
    x0, y0 = clamp(int((px - radius) / 16.0), gx), clamp(int((py - radius) / 16.0), gy)
    x1, y1 = clamp(int((px + radius + 15.0) / 16.0), gx), clamp(int((py + radius + 15.0) / 16.0), gy)
    return x0, y0, x1, y1


def real_sh(l, m, d):
    """Real spherical harmonic of degree l, order m at unit vector d, from the associated Legendre functions
    (Condon-Shortley phase included, as in the published 3DGS tables): an independent derivation of the 25
    polynomials the rasterizer hard-codes."""
    from scipy.special import lpmv
    x, y, z = (float(v) for v in d)
    am = abs(m)
    norm = math.sqrt((2 * l + 1) / (4 * math.pi) * math.factorial(l - am) / math.factorial(l + am))
    phi = math.atan2(y, x)
    leg = float(lpmv(am, l, z))
    if m == 0:
        return norm * leg
    trig = math.cos(am * phi) if m > 0 else math.sin(am * phi)
    return math.sqrt(2.0) * norm * leg * trig


def bits_of(z):
    return int(np.asarray(z, f32).view(np.uint32))


# ------------------------------------------------------------------------------------------------
# the cases.  Each takes a runner and raises AssertionError on a wrong answer.
# ------------------------------------------------------------------------------------------------
def kat_single_isotropic(run):
    """(i) one isotropic Gaussian on the optical axis: every pixel against
    alpha = min(0.99, o exp(-r^2 / 2 sigma^2)), sigma^2 = (f s / z)^2 + 0.3; colour = c alpha + bg (1 - alpha);
    and the closed-form gradients of sum(w * colour) w.r.t. opacity and colour."""
    H = W = 48
    z, s, o = 4.0, 0.25, 0.7
    col = np.array([0.9, 0.4, 0.1])
    bg = (0.1, 0.2, 0.3)
    cam = camera(0.5, 0.5, bg=bg)
    fpix = W / (2 * 0.5)
    sigma2 = (fpix * s / z) ** 2 + 0.3
    cx, cy = (W - 1) / 2, (H - 1) / 2
    a, raw = iso_alpha(H, W, cx, cy, sigma2, o)
    rng = np.random.default_rng(3)
    w = rng.standard_normal((3, H, W))
    res = run(np.array([[0, 0, z]], f32), iso_cov6(s), col[None].astype(f32), np.array([o], f32), cam, H, W, 0,
              grads=(w.astype(f32), None))
    want = col[:, None, None] * a[None] + np.asarray(bg)[:, None, None] * (1 - a[None])
    ok = ~near_threshold(raw)
    assert ok.mean() > 0.99
    assert np.abs(res["color"] - want)[:, ok].max() <= COLOR_ATOL
    assert np.abs(res["final_T"] - (1 - a))[ok].max() <= COLOR_ATOL
    assert np.array_equal(res["n_contrib"][ok], (a > 0).astype(np.uint32)[ok])
    assert int(res["radii"][0]) == radius_of(sigma2) == 10
    assert res["xy"][0].tolist() == [cx, cy]
    assert bits_of(res["depths"][0]) == bits_of(z)
    assert int(res["tiles_touched"][0]) == 9 and res["num_rendered"] == 9
    # gradients: alpha = o g (never capped here), colour_ch = c_ch alpha + bg_ch (1 - alpha)
    g = np.where(a > 0, np.exp(-0.5 * ((pixel_grid(H, W)[0] - cx) ** 2 + (pixel_grid(H, W)[1] - cy) ** 2) / sigma2), 0.0)
    w = w.astype(f32).astype(np.float64)
    want_go = float(((col - np.asarray(bg))[:, None, None] * w * g[None]).sum())
    want_gc = (w * a[None]).sum(axis=(1, 2))
    assert abs(res["g_opacity"][0] - want_go) <= GRAD_RTOL * abs(want_go)
    assert np.abs(res["g_feat"][0] - want_gc).max() <= GRAD_RTOL * np.abs(want_gc).max()
    # symmetric weights => zero net force on the mean (every term cancels pairwise up to rounding)
    wsym = np.ones((3, H, W), f32)
    res2 = run(np.array([[0, 0, z]], f32), iso_cov6(s), col[None].astype(f32), np.array([o], f32), cam, H, W, 0,
               grads=(wsym, None))
    assert np.abs(res2["g_means"][0][:2]).max() <= 1e-4 * max(1.0, abs(res2["g_opacity"][0]))


def kat_alpha_cap_and_cutoff(run):
    """probe P4: opacity 1 => alpha capped at 0.99 near the centre (and the cap is not masked in the gradient);
    tiny covariance => sigma^2 = 0.3 exactly the dilation; cut-off where alpha < 1/255."""
    H = W = 32
    cam = camera(0.5, 0.5)
    z = 4.0
    res = run(np.array([[0, 0, z]], f32), iso_cov6(1e-6), np.array([[1.0, 1.0, 1.0]], f32), np.array([1.0], f32), cam,
              H, W, 0, grads=(np.ones((3, H, W), f32), None))
    a, raw = iso_alpha(H, W, 15.5, 15.5, 0.3 + (W * 1e-6 / z) ** 2, 1.0)
    ok = ~near_threshold(raw) & ~near_threshold(raw, 0.99)
    assert np.abs(res["color"][0] - a)[ok].max() <= COLOR_ATOL
    assert int(res["radii"][0]) == radius_of(0.3) == 3      # ceil(3 sqrt(0.3 + sqrt(0.1)))
    # an 8-pixel-sigma splat of opacity 1: o g >= 0.99 on the four pixels nearest the centre (r^2 = 0.5)
    res = run(np.array([[0, 0, z]], f32), iso_cov6(1.0), np.array([[1.0, 1.0, 1.0]], f32), np.array([1.0], f32), cam, H,
              W, 0, grads=(np.ones((3, H, W), f32), None))
    sigma2 = (W * 1.0 / z) ** 2 + 0.3
    a, raw = iso_alpha(H, W, 15.5, 15.5, sigma2, 1.0)
    assert (a == 0.99).sum() >= 4
    ok = ~near_threshold(raw) & ~near_threshold(raw, 0.99)
    assert np.abs(res["color"][0] - a)[ok].max() <= COLOR_ATOL
    # straight-through cap: dL/do = sum_pix g (bg = 0, colour 1, 3 channels), capped pixels included
    xs, ys = pixel_grid(H, W)
    g = np.where(a > 0, np.exp(-0.5 * ((xs - 15.5) ** 2 + (ys - 15.5) ** 2) / sigma2), 0.0)
    assert abs(res["g_opacity"][0] - 3 * g.sum()) <= GRAD_RTOL * 3 * g.sum()


def kat_depth_order_swap(run):
    """(ii) two overlapping Gaussians: the nearer one is composited first whatever its index; swapping the depths
    swaps the roles."""
    H = W = 32
    cam = camera(0.5, 0.5, bg=(0.05, 0.05, 0.05))
    c = np.array([[1.0, 0.0, 0.2], [0.0, 1.0, 0.4]])
    o = np.array([0.6, 0.8])
    for zs in ((3.0, 5.0), (5.0, 3.0)):
        s = [0.3 * zs[0] / 3.0, 0.2 * zs[1] / 3.0]
        means = np.array([[0, 0, zs[0]], [0, 0, zs[1]]], f32)
        res = run(means, iso_cov6(s), c.astype(f32), o.astype(f32), cam, H, W, 0)
        al, raws = [], []
        for i in range(2):
            sigma2 = (W * s[i] / zs[i]) ** 2 + 0.3
            a, raw = iso_alpha(H, W, 15.5, 15.5, sigma2, o[i])
            al.append(a)
            raws.append(raw)
        first, second = (0, 1) if zs[0] < zs[1] else (1, 0)
        T1 = 1 - al[first]
        want = (c[first][:, None, None] * al[first] + c[second][:, None, None] * al[second] * T1 +
                0.05 * (T1 * (1 - al[second]))[None])
        ok = ~near_threshold(raws[0]) & ~near_threshold(raws[1])
        assert np.abs(res["color"] - want)[:, ok].max() <= COLOR_ATOL, zs
        assert res["vals"][:1].tolist() == [first]      # one tile pair each per tile; the first tile lists the nearer one first
        pos = np.where(al[second] > 0, 2, np.where(al[first] > 0, 1, 0))
        assert np.array_equal(res["n_contrib"][ok], pos.astype(np.uint32)[ok])


def kat_equal_depth_tie(run, n_extra=0):
    """(iii) bit-identical depths: the lower Gaussian index is composited first (stable sort of the emission
    order).  With ``n_extra`` > 0 the tile also holds that many more Gaussians at the SAME depth (a depth bucket that
    overflows the distribution sort and takes the stable radix fallback in the CUDA bucket path)."""
    H = W = 32
    cam = camera(0.5, 0.5)
    z = 4.0
    cols = np.array([[1.0, 0.0, 0.0], [0.0, 1.0, 0.0]])
    o = np.array([0.7, 0.7])
    s = 0.3
    sigma2 = (W * s / z) ** 2 + 0.3
    a, raw = iso_alpha(H, W, 15.5, 15.5, sigma2, 0.7)
    ok = ~near_threshold(raw)
    for order in ((0, 1), (1, 0)):
        means = np.array([[0, 0, z]] * (2 + n_extra), f32)
        c = np.concatenate((cols[list(order)], np.zeros((n_extra, 3))))
        op = np.concatenate((o, np.full(n_extra, 0.003)))  # the extras stay below 1/255 on every pixel
        means[2:, 0] += 0.001 * np.arange(n_extra, dtype=f32)    # distinct Gaussians, same depth bits
        sc = np.concatenate(([s, s], np.full(n_extra, s)))
        res = run(means, iso_cov6(sc), c.astype(f32), op.astype(f32), cam, H, W, 0)
        first = c[0]
        want = first[:, None, None] * a[None] + c[1][:, None, None] * (a * (1 - a))[None]
        assert np.abs(res["color"] - want)[:, ok].max() <= COLOR_ATOL, order
        # every tile's list is in ascending Gaussian index (all depths tie)
        for lo, hi in res["ranges"]:
            v = res["vals"][lo:hi]
            assert np.all(v[1:] > v[:-1])
        k = res["keys"]
        assert np.all(k[1:] >= k[:-1])
        assert len({int(x) & 0xffffffff for x in k}) == 1  # one depth pattern


def kat_radius_rect_table(run):
    """(iv) radius, tile rectangle and tiles_touched for Gaussians straddling tile and image borders, at
    z = 0.2 and one ulp above, behind the camera, with det == 0, and fully off screen."""
    H, W = 48, 64
    tan = 0.5
    z = 4.0
    cam = camera(tan, tan * H / W)  # square pixels: fx = fy = 64

    def world_of(px, py, zz=z):
        ndc_x, ndc_y = (2 * px + 1) / W - 1, (2 * py + 1) / H - 1
        return [ndc_x * zz * tan, ndc_y * zz * (tan * H / W), zz]

    fpix = W / (2 * tan)
    rows = []  # (mean, s, expected (radius, rect) or None when culled)
    limx, limy = 1.3 * tan, 1.3 * tan * H / W
    for px, py, s in ((31.5, 23.5, 0.05), (15.5, 15.5, 0.1), (16.25, 31.75, 0.2), (2.3, 3.4, 0.3), (62.3, 46.2, 0.3),
                      (-6.2, 10.4, 0.4), (70.3, 50.3, 0.4), (-40.0, 10.0, 0.2), (31.5, 90.0, 0.2)):
        m = world_of(px, py)
        # off-axis: the Jacobian has a J02/J12 term, evaluated at the guard-band-clamped point (t.x/t.z, t.y/t.z
        # clamped to +-1.3 tan(fov/2)); isotropic Sigma = s^2 I => cov2D = s^2 J J^T
        tx = min(limx, max(-limx, m[0] / z)) * z
        ty = min(limy, max(-limy, m[1] / z)) * z
        J = np.array([[fpix / z, 0, -fpix * tx / z ** 2], [0, fpix / z, -fpix * ty / z ** 2]])
        c2 = s * s * (J @ J.T)
        r = radius_of(c2[0, 0] + 0.3, c2[1, 1] + 0.3, c2[0, 1])
        rect = tile_rect(px, py, r, H, W)
        area = (rect[2] - rect[0]) * (rect[3] - rect[1])
        rows.append((m, s, (r, rect) if area else None))
    zmin = f32(0.2)
    rows.append(([0, 0, float(zmin)], 0.01, None))                                   # z == 0.2: culled (<=)
    zup = float(np.nextafter(zmin, f32(1)))
    rows.append(([0, 0, zup], 0.01, (radius_of((fpix * 0.01 / zup) ** 2 + 0.3), None)))  # one ulp above: kept
    rows.append(([0, 0, -1.0], 0.1, None))                                           # behind the camera
    means = np.array([r[0] for r in rows], f32)
    cov = iso_cov6([r[1] for r in rows])
    # det == 0: rank-1 covariance along (1,1,0), so large that the 0.3 dilation is rounded away: a = b = c = 2^24
    means = np.concatenate((means, np.array([[0, 0, z]], f32)))
    k0 = 65536.0
    cov = np.concatenate((cov, np.array([[k0, k0, 0, k0, 0, 0]], f32)))
    rows.append(([0, 0, z], None, None))
    G = means.shape[0]
    res = run(means, cov, np.full((G, 3), 0.5, f32), np.full(G, 0.5, f32), cam, H, W, 0)
    for i, (m, s, want) in enumerate(rows):
        if want is None:
            assert res["radii"][i] == 0 and res["tiles_touched"][i] == 0, (i, m)
            continue
        r, rect = want
        assert int(res["radii"][i]) == r, (i, m, int(res["radii"][i]), r)
        if rect is not None:
            assert int(res["tiles_touched"][i]) == (rect[2] - rect[0]) * (rect[3] - rect[1]), (i, rect)
            tiles = sorted(int(k >> np.uint64(32)) for k, v in zip(res["keys"], res["vals"]) if v == i)
            gx = (W + 15) // 16
            assert tiles == [y * gx + x for y in range(rect[1], rect[3]) for x in range(rect[0], rect[2])], (i, tiles)
    assert res["num_rendered"] == int(res["tiles_touched"].sum())


def kat_keys_and_ranges(run):
    """(v) key = (tile << 32) | bits(z_view); list sorted by (tile, depth); ranges = [first, last) per tile;
    the sorted bit range of the reference, 32 + getHigherMsb(T), covers every tile id (T = 256 -> 41, 4096 -> 45)."""
    H, W = 32, 64
    tan = 0.5
    cam = camera(tan, tan * H / W)
    # three Gaussians over tile columns 1-2, depths out of index order
    zs = [6.0, 3.0, 4.5]
    means = np.array([[0.0, 0.0, zz] for zz in zs], f32)     # all centred at pixel (31.5, 15.5)
    s = [12.0 / 64 * zz / 3 for zz in zs]                     # ~4 px sigma each -> radius 13: tiles x in [1,3), y in [0,2)
    res = run(means, iso_cov6(s), np.full((3, 3), 0.5, f32), np.full(3, 0.5, f32), cam, H, W, 0)
    want_keys, want_vals = [], []
    for tile in (1, 2, 5, 6):
        for i in (1, 2, 0):  # ascending depth
            want_keys.append((tile << 32) | bits_of(zs[i]))
            want_vals.append(i)
    assert [int(k) for k in res["keys"]] == want_keys
    assert res["vals"].tolist() == want_vals
    want_ranges = np.zeros((8, 2), np.int64)
    for j, tile in enumerate((1, 2, 5, 6)):
        want_ranges[tile] = (3 * j, 3 * j + 3)
    got = res["ranges"].astype(np.int64)
    assert np.array_equal(got[[1, 2, 5, 6]], want_ranges[[1, 2, 5, 6]])
    assert np.all(got[[0, 3, 4, 7], 1] - got[[0, 3, 4, 7], 0] == 0)


def kat_sh_one_hot(run, degrees=(0, 1, 2, 3, 4)):
    """(vi) one-hot SH sweeps against real spherical harmonics derived from Legendre functions: colour
    = max(0, 0.5 + sum_k sh_k Y_k(dir)), the clamp mask on the gradient, and bands above sh_degree ignored."""
    H = W = 16
    z, s, o = 3.0, 0.2, 0.8
    target = np.array([0.3, -0.2, 5.0])
    rng = np.random.default_rng(11)
    dirs = rng.standard_normal((3, 3))
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    sigma2 = (W * s / z) ** 2 + 0.3
    a, raw = iso_alpha(H, W, 7.5, 7.5, sigma2, o)
    ok = ~near_threshold(raw)
    wsum = a.sum()
    for l in degrees:
        for m in range(-l, l + 1):
            k = l * l + l + m
            for d in dirs:
                eye = target - d * z                      # the Gaussian sits on the optical axis, direction = d
                cam = camera(0.5, 0.5, eye=eye, R=look_at(eye, target))
                y = real_sh(l, m, d)
                sh = np.zeros((1, 25, 3), f32)
                amp = np.array([0.9, -2.5, 0.4]) / max(abs(y), 0.2)
                sh[0, k] = amp
                sh[0, 0] = (0.3, 0.1, -0.2)
                pre = 0.5 + SH_C0 * sh[0, 0].astype(np.float64) + sh[0, k].astype(np.float64) * y if k else \
                    0.5 + SH_C0 * sh[0, 0].astype(np.float64)
                rgb = np.maximum(pre, 0.0)
                w = np.ones((3, H, W), f32)
                res = run(target[None].astype(f32), iso_cov6(s), sh, np.array([o], f32), cam, H, W, 4, grads=(w, None))
                want = rgb[:, None, None] * a[None]
                assert np.abs(res["color"] - want)[:, ok].max() <= 2 * COLOR_ATOL, (l, m, d)
                # dL/dsh[k', ch] = Y_k'(d) * sum_pix alpha for unclamped channels, 0 for clamped ones
                live = (pre >= 0).astype(np.float64)
                if np.abs(pre).min() > 1e-3:
                    gk = res["g_feat"][0, k]
                    assert np.abs(gk - y * wsum * live).max() <= GRAD_RTOL * max(abs(y) * wsum, 1e-3) + 1e-7, (l, m)
                    g0 = res["g_feat"][0, 0]
                    assert np.abs(g0 - SH_C0 * wsum * live).max() <= GRAD_RTOL * SH_C0 * wsum
                if l > 0:  # rendered with sh_degree = l - 1 the coefficient is not evaluated at all
                    res = run(target[None].astype(f32), iso_cov6(s), sh, np.array([o], f32), cam, H, W, l - 1)
                    rgb0 = np.maximum(0.5 + SH_C0 * sh[0, 0].astype(np.float64), 0.0)
                    assert np.abs(res["color"] - rgb0[:, None, None] * a[None])[:, ok].max() <= 2 * COLOR_ATOL, (l, m)


def kat_transmittance_stop(run):
    """(vii) a 40-deep stack of equal Gaussians: compositing stops at the first entry that would take T below 1e-4,
    and that entry is NOT composited; n_contrib counts the entries before it."""
    H = W = 16
    n = 40
    z0, o, s = 3.0, 0.8, 0.3
    cam = camera(0.5, 0.5, bg=(1.0, 1.0, 1.0))
    zs = z0 + 0.01 * np.arange(n)
    means = np.stack((np.zeros(n), np.zeros(n), zs), 1).astype(f32)
    sc = s * zs / z0                                   # same footprint for every layer
    cols = np.stack(((np.arange(n) % 2).astype(float), (np.arange(n) % 3) / 2.0, np.linspace(0, 1, n)), 1)
    res = run(means, iso_cov6(sc), cols.astype(f32), np.full(n, o, f32), cam, H, W, 0)
    sigma2 = (W * s / z0) ** 2 + 0.3
    a, raw = iso_alpha(H, W, 7.5, 7.5, sigma2, o)
    want = np.zeros((3, H, W))
    Tm = np.ones((H, W))
    cnt = np.zeros((H, W), np.uint32)
    live = a > 0
    border = near_threshold(raw)
    for i in range(n):
        test = Tm * (1 - a)
        border |= live & (np.abs(test - 1e-4) <= 2e-9)
        go = live & (test >= 1e-4)
        want += np.where(go, a * Tm, 0.0)[None] * cols[i][:, None, None]
        Tm = np.where(go, test, Tm)
        cnt = np.where(go, i + 1, cnt)
        live = go
    want += Tm[None]  # bg = 1
    ok = ~border
    assert ok.mean() > 0.9
    assert np.abs(res["color"] - want)[:, ok].max() <= COLOR_ATOL
    assert np.array_equal(res["n_contrib"][ok], cnt[ok])
    assert 3 <= int(cnt[7, 7]) < 12  # the centre pixel saturates after a handful of layers, long before the 40th
    assert np.abs(res["final_T"] - Tm)[ok].max() <= COLOR_ATOL


def kat_depth_modes(run):
    """(viii) the fused depth channel in all four modes: sum_i w_i max(0, 0.2820948 f(z_i) + 0.5) with the weights
    of the colour pass, z the UNSCALED camera-space depth, and the `log` mode's log(far) quirk
    (cuda_splatting.py:393-400: z.minimum(near).maximum(far).log())."""
    H = W = 32
    near, far = 1.25, 60.0
    scale = 1.0 / near                                 # scale invariance on (cuda_splatting.py:73-80)
    cam = camera(0.5, 0.5, near=near, far=far, scale=scale)
    zs = np.array([3.0, 6.0])
    o = np.array([0.5, 0.9])
    s = np.array([0.3, 0.9])
    means = np.array([[0, 0, zs[0]], [0, 0, zs[1]]], f32)
    al, raws = [], []
    for i in range(2):
        # the rasterizer sees means * scale and cov * scale^2: the footprint f s / z is unchanged
        a, raw = iso_alpha(H, W, 15.5, 15.5, (W * s[i] / zs[i]) ** 2 + 0.3, o[i])
        al.append(a)
        raws.append(raw)
    wts = [al[0], al[1] * (1 - al[0])]
    ok = ~near_threshold(raws[0]) & ~near_threshold(raws[1])

    def feat(mode, zz):
        if mode == 1:
            return zz
        if mode == 2:
            return 1 / zz
        if mode == 3:
            eps = 1e-10
            dn, df = 1 / (near + eps), 1 / (far + eps)
            return 1 - (1 / (zz + eps) - df) / (dn - df + eps)
        return math.log(max(min(zz, near), far))       # = log(far)

    for mode in (1, 2, 3, 4):
        res = run(means, iso_cov6(s), np.full((2, 3), 0.5, f32), o.astype(f32), cam, H, W, 0, depth_mode=mode)
        want = sum(wts[i] * max(0.0, SH_C0 * feat(mode, zs[i]) + 0.5) for i in range(2))
        assert np.abs(res["depth"] - want)[ok].max() <= COLOR_ATOL * max(1.0, want.max()), mode
        assert np.abs(res["color"][0] - 0.5 * (wts[0] + wts[1]))[ok].max() <= COLOR_ATOL


def kat_scale_invariance_and_near_cull(run):
    """The 0.2 near cull applies to the SCALED depth z / near (cuda_splatting.py:73-80 rescales the scene so that
    near = 1): a Gaussian at z = 0.3 survives with near = 1 and is culled with near = 2 (0.15 <= 0.2)."""
    H = W = 32
    means = np.array([[0, 0, 0.3]], f32)
    for near, visible in ((1.0, True), (2.0, False)):
        cam = camera(0.5, 0.5, near=near, far=100.0, scale=1.0 / near)
        res = run(means, iso_cov6(0.02), np.full((1, 3), 0.5, f32), np.array([0.5], f32), cam, H, W, 0)
        assert (res["radii"][0] > 0) == visible


ALL = [kat_single_isotropic, kat_alpha_cap_and_cutoff, kat_depth_order_swap, kat_equal_depth_tie, kat_radius_rect_table,
       kat_keys_and_ranges, kat_sh_one_hot, kat_transmittance_stop, kat_depth_modes, kat_scale_invariance_and_near_cull]
