"""CPU: the oracle against OUTPUTS OF THE REAL REFERENCE - the per-trial summary.csv rows the MuJoCo-based reference
published for its acrobot open-loop sweep (100 TestTasks scenes, T = 500, Optimise(max 10, min 3), lambda 0.01;
GenTestingData.cpp:136-259).  This is the only place where reference results exist for a config that runs here."""
import numpy as np
import pytest

import paperdata
from helpers import make, oracle_model, oracle_task
from oracle import oracle


def oracle_sweep(tag, trials=range(100), T=500):
    kp, mn, mx = paperdata.METHODS[tag]
    tc = make("acrobot", kp_method=kp, min_N=mn, max_N=mx)
    om = oracle_model(tc)
    cr, it, pc = [], [], []
    for j in trials:
        start, tg = tc.load_trial(j)
        d = oracle.Data(om)
        d.load(tc.start_snapshot(start_pos=start))
        d.step()                                             # GenTestingData.cpp:189: one plant step first
        d.forward()                                          # CreateInitSetupControls (ModelTranslator.cpp:875-880)
        il = oracle.ILQR(om, oracle_task(tc, target=tg), T, threads=4)
        il.set_lambda(0.01)                                  # GenTestingData.cpp:217
        il.optimise(d.save(), np.zeros((T, tc.m)), 10, 3)
        sc = il.scalars()
        cr.append(sc["cost_reduction"])
        it.append(int(sc["num_iterations"]))
        pc.append(float(np.mean(il.get("pct_derivs"))))
    return cr, it, pc


@pytest.mark.parametrize("tag", ["SI_1000", "SI_5", "VC_1_200"])
def test_oracle_reproduces_published_acrobot_sweep(built, tag):
    cr, it, pc = oracle_sweep(tag)
    paperdata.check_against_published(tag, cr, it, pc)


def test_published_percentages_are_the_closed_form(built):
    """set_interval % derivatives of the published files equal the key-point rule (SURVEY 8(c) pin 4)."""
    assert np.allclose(paperdata.PUBLISHED["SI_5"]["pct_derivs"], 20.2)
    assert np.allclose(paperdata.PUBLISHED["SI_1000"]["pct_derivs"], 0.4)
    assert np.allclose(paperdata.PUBLISHED["SI_1"]["pct_derivs"], 100.0)
