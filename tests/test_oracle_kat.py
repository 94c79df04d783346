"""Known-answer tests of the CPU oracle (oracle/tsplat_oracle.c): every case of tests/kat.py - closed forms written
from the published definition of the algorithm, independent of the oracle and of the kernels - run through the oracle.
The ``-m gpu`` twin (tests/test_gpu_kat.py) runs the same cases through libtsplat.so.  CPU only.

These tests are what pins the oracle's constants, thresholds, orders and tie rules in the absence of reference golden
vectors (the extension's source is not on this machine): radius / tile-rectangle / tiles_touched table,
key bit pattern and sort order, equal-depth tie rule, alpha cap and cut-off, transmittance stop, the 25 real
spherical harmonics (against Legendre functions), the four depth modes."""
import numpy as np
import pytest

import kat
import oracle


def oracle_run(means, cov6, feat, opacity, cam, H, W, sh_degree, depth_mode=0, grads=None):
    scene = oracle.Scene(means, cov6, feat, opacity, sh_degree, oracle.SH_GM3, oracle.COV_TRIU6)
    st = oracle.forward(scene, cam, H, W, depth_mode=depth_mode)
    out = dict(color=st.color, depth=st.depth if depth_mode else None, radii=st.radii, tiles_touched=st.tiles_touched,
               depths=st.depths, xy=st.xy, keys=st.keys, vals=st.vals, ranges=st.ranges, n_contrib=st.n_contrib,
               final_T=st.final_T, num_rendered=st.num_rendered)
    if grads is not None:
        acc = oracle.backward(scene, cam, st, grads[0], grads[1], depth_mode=depth_mode)
        out.update(g_means=acc.means, g_cov=acc.cov, g_feat=acc.sh, g_opacity=acc.opacity)
    return out


@pytest.mark.parametrize("case", kat.ALL, ids=lambda f: f.__name__)
def test_known_answer(case):
    case(oracle_run)


def test_equal_depth_tie_with_a_crowded_depth_bucket():
    kat.kat_equal_depth_tie(oracle_run, n_extra=70)


def test_sorted_bit_range():
    """getHigherMsb (A.2) = bit length of the tile count: 32 + 9 = 41 sorted bits at 256x256, 32 + 13 = 45 at 1024x1024."""
    for T in (1, 2, 16, 255, 256, 257, 4096, 17152, 65536):
        assert oracle.higher_msb(T) == T.bit_length(), T
    assert oracle.higher_msb(256) == 9 and oracle.higher_msb(4096) == 13


def test_real_sh_table_is_orthonormal():
    """The independent SH table itself: orthonormal over the sphere (Monte-Carlo-free: a 64x128 Gauss-Legendre x
    uniform quadrature integrates degree-8 polynomials exactly)."""
    xs, ws = np.polynomial.legendre.leggauss(16)
    phis = (np.arange(32) + 0.5) * (2 * np.pi / 32)
    lm = [(l, m) for l in range(5) for m in range(-l, l + 1)]
    Y = np.zeros((25, xs.size, phis.size))
    for k, (l, m) in enumerate(lm):
        for i, z in enumerate(xs):
            r = np.sqrt(1 - z * z)
            for j, p in enumerate(phis):
                Y[k, i, j] = kat.real_sh(l, m, (r * np.cos(p), r * np.sin(p), z))
    gram = np.einsum("aij,bij,i->ab", Y, Y, ws) * (2 * np.pi / 32)
    assert np.abs(gram - np.eye(25)).max() < 1e-10
