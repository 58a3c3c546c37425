"""CPU tests of the multi-GPU sweep's host logic: trial j -> rank j mod G, no data-path collective,
one all_gather of fixed-size records.  The N>1 path runs here as world_size-2 gloo processes."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from trajoptkp_b200 import sweep

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n,world", [(100, 1), (100, 2), (100, 8), (7, 4), (3, 8), (0, 2)])
def test_shard_trials_partitions_exactly(n, world):
    seen = []
    for r in range(world):
        mine = sweep.shard_trials(n, r, world)
        assert all(j % world == r for j in mine)
        assert len(mine) <= sweep.pad_count(n, world)
        seen += mine
    assert sorted(seen) == list(range(n))
    with pytest.raises(ValueError):
        sweep.shard_trials(n, world, world)


def test_pack_and_single_process_gather():
    n = 5
    mine = sweep.shard_trials(n, 0, 1)
    recs = [np.arange(sweep.RECORD_SIZE - 1) + 10 * j for j in mine]
    table = sweep.gather_records(sweep.pack_records(mine, recs, n, 1), n)
    assert table.shape == (n, sweep.RECORD_SIZE)
    assert np.array_equal(table[:, -1], np.arange(n))
    assert np.array_equal(table[3, :-1], recs[3])
    txt = sweep.summary_csv(table)
    assert txt.splitlines()[0] == sweep.SUMMARY_HEADER and len(txt.splitlines()) == n + 1


def test_gather_detects_missing_trials():
    local = sweep.pack_records([0, 2], [np.zeros(sweep.RECORD_SIZE - 1)] * 2, 3, 1)
    with pytest.raises(RuntimeError):
        sweep.gather_records(local, 3)


WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r)
import numpy as np, torch.distributed as dist
from trajoptkp_b200 import sweep
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n = 7
mine = sweep.shard_trials(n, rank, world)
recs = [np.full(sweep.RECORD_SIZE - 1, 100.0 * rank + j) for j in mine]
table = sweep.gather_records(sweep.pack_records(mine, recs, n, world), n, dist=dist)
assert table.shape == (n, sweep.RECORD_SIZE)
for j in range(n):
    assert table[j, -1] == j and table[j, 0] == 100.0 * (j %% world) + j, (j, table[j])
dist.barrier()
dist.destroy_process_group()
sys.stdout.write("rank%%d-ok\n" %% rank)
'''


def test_gather_world_size_2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(script)]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=240, env=env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "rank0-ok" in p.stdout and "rank1-ok" in p.stdout


def test_result_files_in_reference_format(built, tmp_path):
    """summary.csv / summary.yaml / directory naming as GenDataOpenloopOptimisation writes them
    (GenTestingData.cpp:234-257,827-958), parsed the way TestingData/scripts/plot_iLQR_keypoint_results.py does."""
    import time

    import pandas as pd
    import yaml

    from helpers import make
    from trajoptkp_b200 import capi, sweep
    tc = make("walker_run", kp_method=capi.KP_VELOCITY_CHANGE, min_N=1, max_N=20)
    rng = np.random.default_rng(0)
    recs = np.abs(rng.standard_normal((7, sweep.RECORD_SIZE)))
    recs[:, 2] = rng.integers(3, 10, 7)
    when = time.struct_time((2024, 9, 3, 14, 5, 0, 1, 247, -1))
    d = sweep.write_openloop_results(str(tmp_path), tc, 80, recs, when=when)
    assert d == str(tmp_path / "TestingData" / "iLQR" / "walker_run_openloop_20240903_1405")
    df = pd.read_csv(d + "/summary.csv")
    assert list(df.columns) == ["Cost reduction", "Optimisation time (ms)", "Number iterations", "Average num dofs",
                                "Average percent derivs", "Average time derivs (ms)", "Average time BP (ms)",
                                "Average time FP (ms)"]
    assert len(df) == 7
    assert np.allclose(df["Cost reduction"].to_numpy(), recs[:, 0], rtol=1e-5)       # %g keeps 6 significant digits
    assert np.array_equal(df["Number iterations"].to_numpy(), recs[:, 2].astype(int))
    y = yaml.load(open(d + "/summary.yaml"), Loader=yaml.FullLoader)
    assert list(y) == ["optimisation horizon", "controls_noise", "model timestep", "keypoint_name", "keypoint_min_N",
                       "keypoint_max_N", "velocity_change_thresholds"]
    assert y["keypoint_name"] == "*VC_1_20" and y["optimisation horizon"] == 80
    assert y["model timestep"] == tc.timestep and y["keypoint_min_N"] == 1 and y["keypoint_max_N"] == 20
    assert np.allclose(y["velocity_change_thresholds"], tc.vel_thr) and len(y["velocity_change_thresholds"]) == tc.n
    # the other labels of SaveTestSummaryData
    assert sweep.keypoint_method_name("set_interval", 5, 5) == "SI_5"
    assert sweep.keypoint_method_name("adaptive_jerk", 2, 10) == "AJ_2_10"
    assert sweep.keypoint_method_name("iterative_error", 2, 10) == "*IE_2_10"
    assert sweep.keypoint_method_name("set_interval", 2, 10, auto_adjust=True) == "AA_2_10"
    ys = yaml.safe_load(sweep.summary_yaml(40, 0.1, 0.004, "iterative_error", 2, 10, iter_thr=0.5))
    assert ys["error_threshold"] == 0.5 and ys["keypoint_name"] == "*IE_2_10" and ys["controls_noise"] == 0.1


def test_summary_yaml_is_byte_identical_to_a_published_file():
    """The VC_2_20, H = 80 walker file of PaperData (walker_run_synchronus_mpc_20240918_0118/summary.yaml), verbatim."""
    from trajoptkp_b200 import sweep
    want = ("optimisation horizon: 80\ncontrols_noise: 0.5\nmodel timestep: 0.005\nkeypoint_name: \"*VC_2_20\"\n"
            "keypoint_min_N: 2\nkeypoint_max_N: 20\nvelocity_change_thresholds:\n  - 2\n  - 2\n  - 2\n  - 10\n  - 10\n"
            "  - 10\n  - 10\n  - 10\n  - 10")
    got = sweep.summary_yaml(80, 0.5, 0.005, "velocity_change", 2, 20, vel_thr=[2, 2, 2, 10, 10, 10, 10, 10, 10])
    assert got == want


def test_save_trajec_information_matches_the_reference_format(tmp_path):
    """FileHandler.cpp:312-383: four CSVs, row-major, trailing commas, default ostream number format, horizon - 1 lines."""
    from trajoptkp_b200 import sweep
    T, dof, m = 3, 1, 1
    A = np.array([[[1.0, 0.004], [-2.5e-7, 0.999999]], [[1.0, 0.004], [0.0, 123456789.0]], [[9.0, 9.0], [9.0, 9.0]]])
    B = np.array([[[0.0], [0.004]], [[1e-5], [-0.0]], [[9.0], [9.0]]])
    X = np.array([[3.14159265, 0.0], [3.1, -0.25], [9.0, 9.0]])
    U = np.array([[0.5], [-1.0], [9.0]])
    root = sweep.save_trajec_information(str(tmp_path), A, B, X, U, "/acrobot")
    assert root == str(tmp_path) + "/savedTrajecInfo/acrobot"
    read = lambda n: open(root + "/" + n).read()
    assert read("A_matrices.csv") == "1,0.004,-2.5e-07,0.999999,\n1,0.004,0,1.23457e+08,\n"
    assert read("B_matrices.csv") == "0,0.004,\n1e-05,-0,\n"
    assert read("states.csv") == "3.14159,0,\n3.1,-0.25,\n"
    assert read("controls.csv") == "0.5,\n-1,\n"
