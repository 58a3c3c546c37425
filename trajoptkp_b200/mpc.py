"""Headless synchronous MPC harness: GenTestingData::TestingMPC / SingleMPCRun / AsyncronusMPCWorker
(src/GenTestingData.cpp:328-751) in its synchronous mode, batched over trials.

In the reference every trial is a plant thread and a worker thread in strict alternation (num_steps_replan = 1,
include/GenTestingData.h:115): re-plan with ONE iLQR iteration from the plant state, pick the buffer index whose
nominal state is closest to the plant state, apply that control plus Gaussian noise, step the plant.  All trials run
the same control flow, so B trials advance in lock-step here: one tokp_optimise call re-plans the whole batch, one
tokp_step_worlds call steps all plants.  Trials shard over ranks exactly like the open-loop sweep (sweep.py).

Everything numeric runs behind the C ABI; this module is the harness logic only (control buffer, index search, noise,
bookkeeping, final-cost replay with the task's residuals).
"""
import os
import time

import numpy as np

from . import capi

MPC_SUMMARY_HEADER = ("Final cost,Final dist,Average dofs,Average optimisation time (ms),Average percent derivs,"
                      "Average time derivs (ms),Average time BP (ms),Average time FP (ms),Average surprise")
MPC_TRIAL_HEADER = ("Optimisation time (ms),num dofs,% derivs,derivs time (ms),BP time (ms),FP time (ms),"
                    "surprise,expected,new cost")


def task_residuals(tc, qpos, qvel, ctrl, target):
    """MT::Residuals of the supported tasks on host arrays (Acrobot.cpp:26-55, Walker.cpp:59-90); qpos/qvel/ctrl are
    full model vectors of one state."""
    q, v, u = qpos[tc.qpos_adr], qvel[tc.dof_adr], ctrl[tc.ctrl_adr]
    if tc.kind == capi.TASK_ACROBOT:
        # the double minus of the fourth residual is the reference's (Acrobot.cpp:45)
        return np.array([q[0] - target[0], q[1] - target[1], v[0] - target[2], v[1] - -target[3], u[0] - target[4]])
    if tc.kind == capi.TASK_WALKER:
        return np.concatenate([[q[0] - target[0], q[2] - target[1], v[1] - target[2]], u - target[3:3 + len(u)]])
    raise ValueError("no residuals for task kind %r" % (tc.kind,))


def task_cost(tc, resid, terminal):
    """MT::CostFunction (ModelTranslator.cpp:314-327): sum of w_i r_i^2, terminal weights on the last state."""
    w = tc.weight_terminal if terminal else tc.weight
    return float(np.sum(w * resid * resid))


def best_matching_state_index(X_old, current_state, horizon, default_index=0):
    """AsyncronusMPCWorker's search (GenTestingData.cpp:668-684): first i in [0, horizon-1) with the smallest L1 distance
    between X_old[i] and the plant state, provided it is below 1000; otherwise the (timing-derived) default."""
    best, smallest = min(default_index, horizon - 1), 1000.0
    err = np.abs(X_old[:horizon - 1] - current_state[None, :]).sum(axis=1)
    for i in range(horizon - 1):
        if err[i] < smallest:
            smallest, best = err[i], i
    return best


def shift_controls(U, k):
    """Drop the first k controls and repeat the last one k times (GenTestingData.cpp:638-645)."""
    if k <= 0:
        return U
    return np.concatenate([U[k:], np.repeat(U[-1:], k, axis=0)], axis=0)


def run_sync_mpc(tc, horizon, trials, task_timeout, controls_noise=0.5, seed=0, lam0=0.1, progress=None,
                 worker_lag=False, fixed_index=None):
    """Synchronous MPC for the given trial indices as one device batch.  Returns a dict with the per-trial summary rows
    (final cost, averages) and the per-replan records of each trial.

    worker_lag: in the reference the plant thread raises `reoptimise` BEFORE it draws the noise, sets the control and
    steps vis_data (GenTestingData.cpp:460-510), and the worker spins on that flag and copies vis_data at once (:632):
    from the second re-plan on the worker therefore plans from the state before the pending plant step, while the
    index search after Optimise (:658-684) sees the stepped plant.  True models that ordering; False (default) is the
    strict alternation with the copy after the step.

    fixed_index: None runs the search of the checked-in code.  With strict alternation that search always returns 0
    (X_old[0] IS the plant state) and the first control of an accepted iteration is zero, so the plant
    would only ever see noise.  The reference carries the override `bestMatchingStateIndex = 1` (main.cpp:707; commented
    out at GenTestingData.cpp:686); fixed_index=1 is that variant - apply U[1], drop one control per re-plan - and it is
    the one that reproduces the published PaperData walker statistics."""
    m = tc.model
    B, S, nq, nv, nu, n = len(trials), m.S, m.nq, m.nv, m.nu, tc.n
    rng = np.random.default_rng(seed)
    snaps = np.zeros((B, S))
    targets = np.zeros((B, tc.nr))
    for r, j in enumerate(trials):
        start, tg = tc.load_trial(j % tc.num_trials())
        snaps[r] = tc.start_snapshot(start_pos=start)
        targets[r] = tg
    st = np.zeros(B, dtype=np.int32)

    def plant_step(s):
        out = np.zeros_like(s)
        capi.check(capi.lib().tokp_step_worlds(m.h, capi._d(s), capi._d(out), B, 1, -1, 0.0, capi._i(st)))
        return out

    plant = plant_step(snaps)          # TestingMPC: mj_step on master_reset_data, kept by CreateInitSetupControls
    tr = capi.Traj(tc, horizon, B)
    for r in range(B):
        tr.set_targets(r, targets[r])
    lim = m.real("actuator_ctrlrange").reshape(-1, 2)[tc.ctrl_adr]
    noise_std = (lim[:, 1] - lim[:, 0]) / 100.0 * controls_noise          # SingleMPCRun: percent of the control range
    U = np.zeros((B, horizon, tc.m))   # CreateInitOptimisationControls: zeros for acrobot / walker
    lam = np.full(B, lam0)             # Optimiser::lambda persists from one Optimise call to the next
    best = np.zeros(B, dtype=np.int64)
    states, controls = [], []          # replay buffers (trajectory_states / trajectory_controls)
    per_replan = []                    # [replan][B][opt_ms, pct, t_derivs, t_bp, t_fp]
    qs, vs, cs = slice(1, 1 + nq), slice(1 + nq, 1 + nq + nv), slice(S - nu, S)
    prev_plant = plant
    for task_time in range(task_timeout):
        # ---- worker: re-plan from the plant state with one iteration --------------------------------------------
        for r in range(B):
            U[r] = shift_controls(U[r], int(best[r]))
        plan_from = prev_plant if (worker_lag and task_time > 0) else plant
        t0 = time.perf_counter()
        Uo, stats = tr.optimise(plan_from, U, 1, 1, lam=lam)
        wall_ms = 1e3 * (time.perf_counter() - t0)
        U = Uo.copy()
        lam = stats[:, 3].copy()
        per_replan.append(np.stack([np.full(B, wall_ms / B), stats[:, 4], stats[:, 5], stats[:, 6], stats[:, 7]], axis=1))
        cur = np.concatenate([plant[:, qs][:, tc.qpos_adr], plant[:, vs][:, tc.dof_adr]], axis=1)
        for r in range(B):
            X_old = tr.get(r, "X_old").reshape(horizon + 1, 2 * n)
            best[r] = best_matching_state_index(X_old, cur[r], horizon)
        if fixed_index is not None:    # the reference's commented-out override (GenTestingData.cpp:686, main.cpp:707)
            best[:] = fixed_index
        # ---- plant: apply the selected control with noise, record, step ---------------------------------------------
        # controlBuffer[current_control_index = best]; the next re-plan drops the first `best` controls (the worker's
        # own index, not the incremented buffer index: with best = 0 the sequence is never shifted, as in the reference)
        u = U[np.arange(B), best] + rng.standard_normal((B, tc.m)) * noise_std[None, :]
        full_u = plant[:, cs].copy()
        full_u[:, tc.ctrl_adr] = u
        states.append(plant[:, 1:1 + nq + nv].copy())
        controls.append(full_u.copy())
        plant[:, cs] = full_u
        prev_plant = plant.copy()
        plant = plant_step(plant)
        if progress and (task_time + 1) % progress == 0:
            print("mpc step %d / %d" % (task_time + 1, task_timeout), flush=True)
    # ---- final cost: replay of the recorded states and controls (SingleMPCRun, GenTestingData.cpp:565-585) ----------
    final_cost = np.zeros(B)
    N = len(states)
    for i in range(N):
        for r in range(B):
            res = task_residuals(tc, states[i][r, :nq], states[i][r, nq:], controls[i][r], targets[r])
            final_cost[r] += task_cost(tc, res, i == N - 1)
    rec = np.stack(per_replan)         # [replans][B][5]
    mean = rec.mean(axis=0)
    summary = np.zeros((B, 9))
    summary[:, 0] = final_cost         # Final cost; Final dist and Average dofs are 0 for these tasks, surprise 0
    summary[:, 3], summary[:, 4], summary[:, 5], summary[:, 6], summary[:, 7] = mean.T
    return {"summary": summary, "per_replan": rec, "trials": list(trials)}


def mpc_summary_csv(summary):
    """summary.csv of TestingMPC (GenTestingData.cpp:397-410)."""
    lines = [MPC_SUMMARY_HEADER]
    for r in summary:
        lines.append(",".join("%g" % v for v in r))
    return "\n".join(lines) + "\n"


def mpc_trial_csv(rec_b):
    """<trial>.csv of AsyncronusMPCWorker (GenTestingData.cpp:703-720): one row per re-plan."""
    lines = [MPC_TRIAL_HEADER]
    for opt_ms, pct, td, tb, tf in rec_b:
        lines.append("%g,0,%g,%g,%g,%g,0,0,0" % (opt_ms, pct, td, tb, tf))
    return "\n".join(lines) + "\n"


def write_mpc_results(root, tc, horizon, result, controls_noise=0.5, optimiser_name="iLQR", when=None):
    """Directory + files as TestingMPC writes them (synchronus_mpc naming is the reference's spelling)."""
    from . import sweep
    d = sweep.test_directory(root, optimiser_name, tc.name, "synchronus_mpc", when)
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "summary.csv"), "w") as f:
        f.write(mpc_summary_csv(result["summary"]))
    for r, j in enumerate(result["trials"]):
        with open(os.path.join(d, "%d.csv" % j), "w") as f:
            f.write(mpc_trial_csv(result["per_replan"][:, r, :]))
    with open(os.path.join(d, "summary.yaml"), "w") as f:
        f.write(sweep.summary_yaml(horizon, controls_noise, tc.timestep, sweep.method_name_of(tc), tc.min_N, tc.max_N,
                                   tc.vel_thr, tc.jerk_thr, tc.iter_thr))
    return d
