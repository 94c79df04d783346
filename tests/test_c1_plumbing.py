"""BASELINE config 1: the reference's own encoder (EncoderTrans, unmodified, random-init weights, CPU) feeds
the render path's call surface unchanged - scripts/c1_plumbing.py.  Needs the reference tree: skipped where it is absent
(the GPU box)."""
import os
import subprocess
import sys
import json

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_encoder_feeds_the_render_path():
    ref = os.environ.get("TSPLAT_REFERENCE_ROOT", "/root/reference")
    if not os.path.exists(os.path.join(ref, "src", "model", "encoder", "encoder_trans.py")):
        pytest.skip("reference tree not present on this machine")
    # a fresh interpreter: the harness installs import hooks and stand-in modules that must not leak into other tests
    out = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "c1_plumbing.py")], capture_output=True, text=True,
                         timeout=900)
    assert out.returncode == 0, out.stderr[-2000:]
    rep = json.loads(out.stdout.strip().splitlines()[-1])
    assert rep["status"] == "pass", rep
    assert rep["gaussians"] == 131072 and rep["finite"]
    assert rep["shapes"]["harmonics"] == [1, 131072, 3, 25] and rep["shapes"]["covariances"] == [1, 131072, 3, 3]
    assert rep["num_rendered_last_pass"] > 100_000
    # the batched level-2 contract (encoder layouts, device-style camera packing) renders the same image as the
    # unmodified reference decoder file fed with the same Gaussians
    assert rep["batched_contract_max_abs_vs_reference_file"]["color"] <= 1e-5
    assert rep["batched_contract_max_abs_vs_reference_file"]["depth"] <= 1e-5
