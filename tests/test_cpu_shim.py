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
