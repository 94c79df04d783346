"""The CPU restatement of GaussianAdapter.forward (oracle/adapter_ref.py) against the fixture recorded from the
UNMODIFIED reference (tests/golden/ref_adapter_small.npz, made by tests/golden/make_golden_adapter.py): outputs and
autograd gradients, float32, tolerance 2e-6 of each tensor's magnitude (same arithmetic, different op grouping)."""
import os

import numpy as np
import torch

from oracle.adapter_ref import adapter_restated

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_adapter_small.npz")


def load(tag):
    z = np.load(GOLD)
    deg, b, v, h, w, smin, smax = z[f"{tag}_cfg"]
    t = lambda k: torch.from_numpy(z[f"{tag}_{k}"])
    wig = [t(f"wigner{l}") for l in range(int(deg) + 1)]
    return z, t, wig, int(deg), (int(h), int(w)), float(smin), float(smax)


def rel(got, want):
    want = np.asarray(want, np.float64)
    return float(np.abs(got.detach().cpu().double().numpy() - want).max() / max(np.abs(want).max(), 1e-30))


def test_restatement_matches_reference_fixture():
    for tag in ("d4", "d2"):
        z, t, wig, deg, hw, smin, smax = load(tag)
        leaves = [t(k).clone().requires_grad_(True) for k in ("coordinates", "depths", "opacities", "raw")]
        outs = adapter_restated(t("extrinsics"), t("intrinsics"), *leaves, wig, hw, smin, smax, deg)
        names = ("means", "covariances", "harmonics", "opacities", "scales", "rotations")
        for o, k in zip(outs, names):
            assert rel(o, z[f"{tag}_out_{k}"]) <= 2e-6, (tag, k)
        loss = sum((o * t(f"w_{k}")).sum() for o, k in zip(outs[:4], names[:4]))
        loss.backward()
        for leaf, k in zip(leaves, ("coordinates", "depths", "opacities", "raw")):
            assert rel(leaf.grad, z[f"{tag}_g_{k}"]) <= 5e-6, (tag, k)
