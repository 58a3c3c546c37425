"""SURVEY 8(f) N2 on the device: the Jacobian end-effector path follower (PushBaseClass::JacobianEEControl,
src/ModelTranslator/PushBaseClass.cpp:139-248) with its way-point builders, run in the order of
GenTestingData::GenDataOpenloopOptimisation (src/GenTestingData.cpp:177-220) from the state
ModelTranslator::InitialiseSystemToStartState leaves - device (tokp_init_controls_ee) against the oracle
(oracle/ok_init.h) on TestTasks trials.  Tolerances: controls and start state 1e-9 relative (north_star per-step
states); final optimised cost within 1 %."""
import numpy as np
import pytest

from helpers import make, oracle_model, oracle_task
from oracle import oracle
from trajoptkp_b200 import capi

pytestmark = pytest.mark.gpu

CTRL_RTOL = 1e-9


def _trials(tc, idx):
    snaps, goals, targets = [], [], []
    for i in idx:
        s, g = tc.init_snapshot(i)
        snaps.append(s)
        goals.append(g)
        targets.append(tc.load_trial(i)[1])
    return np.stack(snaps), np.stack(goals), targets


def _max_ncon(om, tc, snap0, U):
    d = oracle.Data(om)
    d.load(snap0)
    worst = 0
    for t in range(len(U)):
        s = d.save()
        s[tc.model.S - tc.model.nu + tc.ctrl_adr] = U[t]
        d.load(s)
        d.step()
        worst = max(worst, int(d.get("ncon")[0]))
    return worst


def test_box_sweep_ee_init_controls_match_oracle(built):
    tc = make("box_sweep")
    T, idx = 120, [0, 1, 7]
    om = oracle_model(tc)
    snaps, goals, targets = _trials(tc, idx)
    tr = capi.Traj(tc, T, len(idx))
    snap0_dev, U_dev, st = tr.init_controls_ee(snaps, goals, pre_steps=1, setup_horizon=0)
    assert not st.any(), st
    for b, i in enumerate(idx):
        ot = oracle_task(tc, target=targets[b])
        snap0, U, _, warn = oracle.init_sequence_ee(om, ot, snaps[b], goals[b], 0, T)
        assert warn == 0
        assert np.abs(snap0_dev[b] - snap0).max() <= CTRL_RTOL * max(1.0, np.abs(snap0).max())
        assert np.abs(U_dev[b] - U).max() <= CTRL_RTOL * max(1.0, np.abs(U).max()), (i, np.abs(U_dev[b] - U).max())
    # the follower is a feedback law that moves the gripper: controls differ from a gravity-compensation hold
    assert np.abs(U_dev[:, 0] - U_dev[:, -1]).max() > 1.0


def test_box_sweep_arm_reaches_box_and_cost_falls(built):
    """with the reference's initial controls the gripper touches the box (ncon > 8 = box on floor 4 +
    display_goal 4) and iLQR makes progress, so "final cost within 1 %" is no longer a comparison of two idle optimisers."""
    tc = make("box_sweep", kp_method=capi.KP_SET_INTERVAL, min_N=5)
    T = 300
    om = oracle_model(tc)
    snaps, goals, targets = _trials(tc, [0])
    ot = oracle_task(tc, target=targets[0])
    tr = capi.Traj(tc, T, 1)
    tr.set_targets(0, targets[0])
    snap0, U, st = tr.init_controls_ee(snaps, goals, pre_steps=1, setup_horizon=0)
    assert not st.any()
    assert _max_ncon(om, tc, snap0[0], U[0]) > 8
    ref = oracle.ILQR(om, ot, T, threads=8)
    ref.set_lambda(0.01)
    ref.optimise(snap0[0], U[0], 4, 2)
    sc = ref.scalars()
    _, stats = tr.optimise(snap0, U, 4, 2, lam=0.01)
    assert abs(stats[0, 0] - sc["initial_cost"]) <= 1e-9 * abs(sc["initial_cost"])
    assert 1.0 - stats[0, 1] / stats[0, 0] > 0.2, stats[0]
    assert abs(stats[0, 1] - sc["new_cost"]) <= 0.01 * abs(sc["new_cost"]), (stats[0], sc)
    assert int(stats[0, 2]) == int(sc["num_iterations"])


@pytest.mark.parametrize("name", ["push_ncl", "push_lcl"])
def test_push_setup_and_push_controls_match_oracle(built, name):
    """TwoDPushing: CreateInitSetupControls moves the gripper behind the cylinder (setup steps), the optimisation
    controls then push it (TwoDPushing.cpp:225-289).  A shortened setup keeps the oracle leg in seconds."""
    tc = make(name)
    T, setup, idx = 60, 150, [0, 3]
    om = oracle_model(tc)
    snaps, goals, targets = _trials(tc, idx)
    tr = capi.Traj(tc, T, len(idx))
    snap0_dev, U_dev, st, Us_dev = tr.init_controls_ee(snaps, goals, pre_steps=1, setup_horizon=setup, want_setup=True)
    assert not st.any(), st
    for b, i in enumerate(idx):
        ot = oracle_task(tc, target=targets[b])
        snap0, U, Us, warn = oracle.init_sequence_ee(om, ot, snaps[b], goals[b], setup, T)
        assert warn == 0
        assert np.abs(Us_dev[b] - Us).max() <= CTRL_RTOL * max(1.0, np.abs(Us).max()), (i, np.abs(Us_dev[b] - Us).max())
        assert np.abs(snap0_dev[b] - snap0).max() <= CTRL_RTOL * max(1.0, np.abs(snap0).max())
        assert np.abs(U_dev[b] - U).max() <= CTRL_RTOL * max(1.0, np.abs(U).max()), (i, np.abs(U_dev[b] - U).max())
