"""CPU: the C++ host mirror of the reference classes (trajoptkp_b200/host/) compiles against include/tokp.h alone,
resolves the task's index maps through the C ABI, and raises (no host fallback) when no CUDA device exists."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "trajoptkp_b200", "_build", "shim_test")
XML = os.path.join(ROOT, "assets", "models", "acrobot.xml")


def test_shim_headers_use_only_the_c_abi():
    for f in ("tokp_shim.h", "tokp_shim_classes.h", "tokp_shim_optimiser.h"):
        src = open(os.path.join(ROOT, "trajoptkp_b200", "host", f)).read()
        assert "cuda" not in src.lower().replace("cuda library", "")
        assert "oracle" not in src
        for inc in [l for l in src.splitlines() if l.startswith("#include \"")]:
            assert inc.split('"')[1] in ("tokp.h", "tokp_shim.h", "tokp_shim_classes.h"), inc


def test_shim_builds_and_maps_indices(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present: covered by test_gpu_shim.py")
    p = subprocess.run([EXE, XML, "--no-gpu"], capture_output=True, text=True, timeout=60)
    assert p.returncode == 0, p.stderr
    out = json.loads(p.stdout)
    assert (out["dof"], out["num_ctrl"], out["nq"], out["S"]) == (2, 1, 2, 30)
    assert out["traj_create_error"] == "raised"


def test_model_translator_gather_scatter_and_host_residuals(built):
    """SURVEY 8(a) a9 / a10 on the shim: Set*/Return* vectors round-trip through the 8-field snapshot, size mismatches
    are rejected as in the reference, control limits come from actuator_ctrlrange, and the host Residuals / CostFunction
    (used by harness code such as the MPC final-cost replay) give the hand-computed acrobot values."""
    p = subprocess.run([EXE, XML, "--no-gpu"], capture_output=True, text=True, timeout=60)
    out = json.loads(p.stdout)
    assert out["scatter_ok"] == 1 and out["size_mismatch_rejected"] == 1
    assert "state vector size does not match" in p.stderr
    assert out["state_roundtrip"] == [0.5, -0.25, 1.5, -2.0, -0.25, -2.0, 0.75]
    assert out["ctrl_limits"] == [-1.0, 1.0]
    # targets are zero; acrobot.yaml weights {0, 0, 0.001, 0.001, 100}, terminal {100, 100, 1, 1, 100}
    assert out["residuals"] == [0.5, -0.25, 1.5, -2.0, 0.75]
    assert abs(out["cost_running"] - (0.001 * 1.5 ** 2 + 0.001 * 2.0 ** 2 + 100 * 0.75 ** 2)) < 1e-12
    assert abs(out["cost_terminal"] - (100 * 0.25 + 100 * 0.0625 + 2.25 + 4.0 + 100 * 0.5625)) < 1e-12
