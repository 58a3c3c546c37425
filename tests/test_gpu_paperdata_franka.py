"""GPU: the device open-loop sweep of the Franka configs (sweep.run_openloop_sweep: 100 TestTasks trials as ONE tokp_traj
batch at the reference's full horizon, the reference's own initial controls - gravity compensation for reaching, the
Jacobian end-effector follower for box_sweep / push) against the per-trial rows the MuJoCo-based reference published
(tests/paperdata_franka.py).  Full tables of all six published methods per task: profiles/r2a_*_openloop_vs_paperdata.json."""
import numpy as np
import pytest

import paperdata_franka as pf
from helpers import make
from trajoptkp_b200 import sweep

pytestmark = pytest.mark.gpu


def _run(task, tag):
    entry = pf.published(task)[tag]
    tc = pf.configure(make(task), tag, entry)
    recs = sweep.run_openloop_sweep(tc, entry["horizon"], num_trials=100)
    return pf.check(task, tag, recs[:, 0], recs[:, 2].astype(int), recs[:, 4])


@pytest.mark.parametrize("tag", ["SI_1000", "AJ_5_100", "VC_1_100", "IE_1_50"])
def test_reaching_sweep_reproduces_published(built, tag):
    r = _run("reaching", tag)
    assert r["iters_equal"] >= 30  # trial by trial the same number of iterations for most scenes (SI_1000: 37, others 72-81)


@pytest.mark.parametrize("tag", ["SI_1000", "AJ_1_1000", "VC_1_100"])
def test_box_sweep_sweep_reproduces_published(built, tag):
    r = _run("box_sweep", tag)
    assert r["cost_reduction"][0] > 0.35  # the arm sweeps the box: iLQR makes the progress the reference made


@pytest.mark.parametrize("tag", ["SI_1000", "AJ_5_100"])
def test_push_ncl_sweep_reproduces_published(built, tag):
    r = _run("push_ncl", tag)
    assert r["cost_reduction"][0] > 0.12
