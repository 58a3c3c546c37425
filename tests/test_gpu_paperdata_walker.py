"""GPU: synchronous walker_run MPC through the device (trajoptkp_b200/mpc.py, 100 trials as one tokp_traj batch, one
tokp_optimise + one tokp_step_worlds per plant step) against the per-trial results the real MuJoCo-based reference
published (tests/golden/paperdata_walker_mpc.json).  The contact-rich config's only available reference outputs.

The runs are noisy (0.5 % control noise with the reference's unseeded RNG) and bimodal (a trial either gets the walker
running or not), so the pin is on distribution statistics; margins are 2-4x the deviations measured in round 1
(profiles/r1n_walker_mpc_vs_paperdata.json)."""
import json
import os

import numpy as np
import pytest

from helpers import make
from trajoptkp_b200 import capi, mpc

pytestmark = pytest.mark.gpu
PUB = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "paperdata_walker_mpc.json")))


def test_adaptive_jerk_h20_statistics(built):
    """AJ_2_20, H = 20: the share of derivative columns depends on the simulated jerk profile under contact."""
    pub = PUB["AJ_2_20:20"]
    tc = make("walker_run", kp_method=capi.KP_ADAPTIVE_JERK, min_N=2, max_N=20)
    res = mpc.run_sync_mpc(tc, 20, list(range(100)), 1500, controls_noise=pub["controls_noise"], fixed_index=1)
    cost, pct = res["summary"][:, 0], res["summary"][:, 4]
    assert abs(pct.mean() - np.mean(pub["pct_derivs"])) < 3.0, (pct.mean(), np.mean(pub["pct_derivs"]))
    assert abs(np.median(cost) / np.median(pub["final_cost"]) - 1) < 0.12, (np.median(cost), np.median(pub["final_cost"]))
    assert abs(cost.mean() / np.mean(pub["final_cost"]) - 1) < 0.15
    # a 20-step horizon is too short for the walker to get going: (almost) no trial reaches a low cost, as published
    assert np.mean(cost < 1000) < 0.15 and np.mean(np.array(pub["final_cost"]) < 1000) < 0.15


def test_set_interval_h80_statistics(built):
    """SI_5, H = 80: with the long horizon most trials end with a low cost (the walker runs); key-point share 21.25 %."""
    pub = PUB["SI_5:80"]
    tc = make("walker_run", kp_method=capi.KP_SET_INTERVAL, min_N=5, max_N=20)
    res = mpc.run_sync_mpc(tc, 80, list(range(48)), 1500, controls_noise=pub["controls_noise"], fixed_index=1)
    cost, pct = res["summary"][:, 0], res["summary"][:, 4]
    assert np.allclose(pct, 21.25) and np.allclose(pub["pct_derivs"], 21.25)
    pc = np.array(pub["final_cost"])
    assert 0.4 < np.median(cost) / np.median(pc) < 2.5, (np.median(cost), np.median(pc))
    assert np.mean(cost < 300) > 0.8 and np.mean(pc < 300) > 0.8
    assert np.percentile(cost, 10) < 30 and np.percentile(pc, 10) < 30


def test_search_variant_applies_only_noise(built):
    """The checked-in index search under strict alternation returns 0 and the first control of an accepted iteration
    is zero: the plant sees noise only, whatever the horizon - documented in DESIGN.md section 2."""
    tc = make("walker_run", kp_method=capi.KP_SET_INTERVAL, min_N=5, max_N=20)
    a = mpc.run_sync_mpc(tc, 20, list(range(4)), 60, seed=1)["summary"][:, 0]
    b = mpc.run_sync_mpc(tc, 40, list(range(4)), 60, seed=1)["summary"][:, 0]
    assert np.allclose(a, b, rtol=1e-12)
