"""The known-answer cases of tests/kat.py run through libtsplat.so (C-ABI, via the torch binding), on both binning
paths: the CUDA kernels are held to the same closed-form answers as the CPU oracle (tests/test_oracle_kat.py)."""
import numpy as np
import pytest
import torch

import kat

pytestmark = pytest.mark.gpu


def make_runner(sort):
    from transplat_b200._lib import COV_TRIU6, SH_GM3
    from transplat_b200.rasterize import RenderConfig, backward_raw, forward_raw, inspect_workspace

    def run(means, cov6, feat, opacity, cam, H, W, sh_degree, depth_mode=0, grads=None):
        dev = torch.device("cuda:0")
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(dev)
        cfg = RenderConfig(H=H, W=W, V=1, sh_degree=sh_degree, sh_layout=SH_GM3, cov_layout=COV_TRIU6,
                           depth_mode=depth_mode, want_radii=True, want_means2d_grad=True, check="sync", sort=sort)
        args = (t(means)[None], t(cov6)[None], t(feat)[None], t(opacity)[None], t(cam)[None])
        color, depth, radii, state = forward_raw(cfg, *args)
        ins = inspect_workspace(state)
        n = lambda x: x.detach().cpu().numpy()
        out = dict(color=n(color)[0], depth=n(depth)[0] if depth is not None else None, radii=n(radii)[0],
                   tiles_touched=n(ins["tiles_touched"])[0].astype(np.uint32), depths=n(ins["depths"])[0],
                   xy=n(ins["xy"])[0], keys=n(ins["keys"]).view(np.uint64), vals=n(ins["vals"]).view(np.uint32),
                   ranges=n(ins["ranges"])[0].astype(np.uint32), n_contrib=n(ins["n_contrib"])[0].astype(np.uint32),
                   final_T=n(ins["final_T"])[0], num_rendered=ins["num_rendered"])
        assert np.array_equal(out["radii"], n(ins["radii"])[0])
        if grads is not None:
            gc = t(grads[0])[None]
            gd = t(grads[1])[None] if (depth_mode and grads[1] is not None) else None
            g = backward_raw(cfg, state, gc, gd)
            out.update(g_means=n(g[0])[0].astype(np.float64), g_cov=n(g[1])[0].astype(np.float64),
                       g_feat=n(g[2])[0].astype(np.float64), g_opacity=n(g[3])[0].astype(np.float64))
        return out

    return run


@pytest.mark.parametrize("sort", ["bucket", "onesweep"])
@pytest.mark.parametrize("case", kat.ALL, ids=lambda f: f.__name__)
def test_known_answer(case, sort):
    case(make_runner(sort))


@pytest.mark.parametrize("sort", ["bucket", "onesweep"])
def test_equal_depth_tie_with_a_crowded_depth_bucket(sort):
    """72 Gaussians at one bit-identical depth in every tile: the distribution sort's bucket overflows and the stable
    radix fallback (bucket path) / the stable onesweep passes must return ascending Gaussian index."""
    kat.kat_equal_depth_tie(make_runner(sort), n_extra=70)
