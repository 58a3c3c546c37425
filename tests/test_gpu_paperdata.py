"""GPU: the device sweep (trajoptkp_b200/sweep.py: 100 acrobot trials as one tokp_traj batch through tokp_optimise)
against the per-trial results the real reference published (see tests/paperdata.py), and against the oracle's sweep."""
import numpy as np
import pytest

import paperdata
from helpers import make
from trajoptkp_b200 import sweep

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("tag", ["SI_1", "SI_5", "SI_1000", "AJ_1_50", "VC_1_200", "IE_1_50"])
def test_device_sweep_reproduces_published_acrobot_sweep(built, tag):
    kp, mn, mx = paperdata.METHODS[tag]
    tc = make("acrobot", kp_method=kp, min_N=mn, max_N=mx)
    recs = sweep.run_openloop_sweep(tc, 500, num_trials=100)
    paperdata.check_against_published(tag, recs[:, 0], recs[:, 2].astype(int), recs[:, 4])


def test_device_sweep_matches_oracle_sweep(built):
    from test_cpu_paperdata import oracle_sweep
    tc = make("acrobot", kp_method=paperdata.METHODS["SI_1000"][0], min_N=1000, max_N=1000)
    recs = sweep.run_openloop_sweep(tc, 500, num_trials=100)
    cr, it, _ = oracle_sweep("SI_1000")
    assert (np.array(it) == recs[:, 2].astype(int)).sum() >= 98
    assert np.median(np.abs(np.array(cr) - recs[:, 0])) < 1e-6
