"""Multi-GPU trial sweep: host-side logic for SURVEY 8(e).

Replaces the serial `for i in 0..99` of GenTestingData::GenDataOpenloopOptimisation
(src/GenTestingData.cpp:136-259) and TestingMPC (:328-420): trials are independent, so trial j goes to rank
j mod G (one process per GPU), each rank runs complete Optimise calls on its own device batch, and ONE
all_gather of fixed-size result records ends the sweep.  There is no data-path collective and a single
trajectory never leaves its GPU.

Record columns follow the reference's summary.csv (GenTestingData.cpp:236-247) plus two device figures:
  cost_reduction, opt_time_ms, iterations, avg_dofs, avg_pct_derivs, t_derivs_ms, t_bp_ms, t_fp_ms,
  jacobians, trial_index
"""
import os

import numpy as np

RECORD_FIELDS = ("cost_reduction", "opt_time_ms", "iterations", "avg_dofs", "avg_pct_derivs", "t_derivs_ms",
                 "t_bp_ms", "t_fp_ms", "jacobians", "trial_index")
RECORD_SIZE = len(RECORD_FIELDS)
SUMMARY_HEADER = ("Cost reduction,Optimisation time (ms),Number iterations,Average num dofs,Average percent derivs,"
                  "Average time derivs (ms),Average time BP (ms),Average time FP (ms)")


def shard_trials(num_trials, rank, world):
    """Trial indices owned by `rank`: j with j mod world == rank (SURVEY 8(e) partitioning)."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world %d" % (rank, world))
    return list(range(rank, num_trials, world))


def pad_count(num_trials, world):
    """Records per rank after padding so that all_gather sees equal shapes."""
    return (num_trials + world - 1) // world


def pack_records(trial_indices, records, num_trials, world):
    """[pad_count, RECORD_SIZE] array; unused rows carry trial_index = -1."""
    out = np.zeros((pad_count(num_trials, world), RECORD_SIZE))
    out[:, -1] = -1.0
    for r, (j, rec) in enumerate(zip(trial_indices, records)):
        rec = np.asarray(rec, dtype=np.float64)
        assert rec.shape == (RECORD_SIZE - 1,), rec.shape
        out[r, :-1] = rec
        out[r, -1] = j
    return out


def gather_records(local, num_trials, dist=None, device=None):
    """The sweep's only collective: all_gather of each rank's packed records, re-ordered by trial index.
    `dist` is torch.distributed (initialised: NCCL over NVLink on the GPU box, gloo in the CPU tests) or None for a
    single process."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        allr = local[None]
    else:
        import torch
        t = torch.as_tensor(local, dtype=torch.float64, device=device)
        bufs = [torch.empty_like(t) for _ in range(dist.get_world_size())]
        dist.all_gather(bufs, t)
        allr = torch.stack(bufs).cpu().numpy()
    flat = allr.reshape(-1, RECORD_SIZE)
    flat = flat[flat[:, -1] >= 0]
    order = np.argsort(flat[:, -1], kind="stable")
    flat = flat[order]
    if len(flat) != num_trials or not np.array_equal(flat[:, -1], np.arange(num_trials)):
        raise RuntimeError("sweep gather: expected %d trial records, got indices %s" % (num_trials, flat[:, -1]))
    return flat


def summary_csv(records):
    """The reference's summary.csv text (GenTestingData.cpp:236-250) from gathered records."""
    lines = [SUMMARY_HEADER]
    for r in records:
        lines.append("%g,%g,%d,%g,%g,%g,%g,%g" % (r[0], r[1], int(r[2]), r[3], r[4], r[5], r[6], r[7]))
    return "\n".join(lines) + "\n"


def keypoint_method_name(method, min_N, max_N, auto_adjust=False):
    """Label written to summary.yaml (GenTestingData::SaveTestSummaryData, GenTestingData.cpp:868-892): SI_<min>,
    *VC_<min>_<max>, AJ_<min>_<max>, *IE_<min>_<max>, AA_<min>_<max> for auto-adjust; other methods get an empty name."""
    if auto_adjust:
        return "AA_%d_%d" % (min_N, max_N)
    if method == "set_interval":
        return "SI_%d" % min_N
    prefix = {"velocity_change": "*VC", "adaptive_jerk": "AJ", "iterative_error": "*IE"}.get(method)
    return "" if prefix is None else "%s_%d_%d" % (prefix, min_N, max_N)


def _yaml_scalar(v):
    if isinstance(v, str):
        # yaml-cpp quotes scalars that start with an indicator character such as '*'
        return '"%s"' % v if (v == "" or v[0] in "*&!|>%@`#,[]{}-?:'\"") else v
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, int):
        return "%d" % v
    # yaml-cpp writes doubles with digits10 + 1 = 16 significant digits (0.005 stays "0.005")
    return "%.16g" % v if float(v) != int(v) or abs(v) >= 1e15 else "%d" % int(v)


def summary_yaml(horizon, controls_noise, timestep, method, min_N, max_N, vel_thr=None, jerk_thr=None, iter_thr=None,
                 auto_adjust=False):
    """The reference's summary.yaml (GenTestingData.cpp:894-958): same keys in the same order, block map with block
    sequences for the threshold lists and no trailing newline, byte for byte what yaml-cpp's emitter writes (checked
    against PaperData/.../walker_run_synchronus_mpc_20240918_0118/summary.yaml)."""
    rows = [("optimisation horizon", int(horizon)), ("controls_noise", float(controls_noise)),
            ("model timestep", float(timestep)),
            ("keypoint_name", keypoint_method_name(method, min_N, max_N, auto_adjust)),
            ("keypoint_min_N", int(min_N)), ("keypoint_max_N", int(max_N))]
    lines = ["%s: %s" % (k, _yaml_scalar(v)) for k, v in rows]
    if method == "velocity_change" and vel_thr is not None:
        lines += ["velocity_change_thresholds:"] + ["  - %s" % _yaml_scalar(float(x)) for x in vel_thr]
    if method == "adaptive_jerk" and jerk_thr is not None:
        lines += ["jerk_thresholds:"] + ["  - %s" % _yaml_scalar(float(x)) for x in jerk_thr]
    if method == "iterative_error" and iter_thr is not None:
        lines.append("error_threshold: %s" % _yaml_scalar(float(iter_thr)))
    return "\n".join(lines)


def test_directory(root, optimiser_name, task_name, testing_method, when=None):
    """GenTestingData::CreateTestName (GenTestingData.cpp:827-860):
    <root>/TestingData/<optimiser>/<task>_<openloop|synchronus_mpc|asynchronus_mpc>_<YYYYMMDD_HHMM>."""
    import os
    import time
    stamp = time.strftime("%Y%m%d_%H%M", time.localtime() if when is None else when)
    return os.path.join(root, "TestingData", optimiser_name, "%s_%s_%s" % (task_name, testing_method, stamp))


def write_openloop_results(root, tc, horizon, records, optimiser_name="iLQR", controls_noise=0.0, when=None):
    """Writes summary.csv + summary.yaml where and how GenDataOpenloopOptimisation does, so that the reference's
    TestingData/scripts/plot_iLQR_keypoint_results.py reads them unchanged.  Returns the directory."""
    import os

    from . import capi
    d = test_directory(root, optimiser_name, tc.name, "openloop", when)
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "summary.csv"), "w") as f:
        f.write(summary_csv(records))
    with open(os.path.join(d, "summary.yaml"), "w") as f:
        f.write(summary_yaml(horizon, controls_noise, tc.timestep, method_name_of(tc), tc.min_N, tc.max_N, tc.vel_thr,
                             tc.jerk_thr, tc.iter_thr))
    return d


def method_name_of(tc):
    """keypoint_method.name of a task config."""
    from . import capi
    return {capi.KP_SET_INTERVAL: "set_interval", capi.KP_ADAPTIVE_JERK: "adaptive_jerk",
            capi.KP_ADAPTIVE_ACCEL: "adaptive_accel", capi.KP_ITERATIVE_ERROR: "iterative_error",
            capi.KP_VELOCITY_CHANGE: "velocity_change"}[tc.kp_method]


def prepare_trials(tc, trial_indices, horizon, tr=None):
    """Everything GenDataOpenloopOptimisation does for a trial before Optimise (GenTestingData.cpp:177-214), for this
    rank's trials at once on the device: InitialiseSystemToStartState, one plant step, CreateInitSetupControls
    (mj_forward; TwoDPushing: 1000 controller steps that park the gripper behind the object),
    CreateInitOptimisationControls of the task (zeros: acrobot / walker, ModelTranslator.cpp:1783-1797; gravity
    compensation: Reaching.cpp:269-296; Jacobian end-effector follower: BoxSweep.cpp:72-99, TwoDPushing.cpp:260-289).
    Returns (snapshot0[B][S], U[B][T][m], targets[B][...])."""
    from . import capi
    B = len(trial_indices)
    snaps = np.zeros((B, tc.model.S))
    goals = np.zeros((B, 2))
    targets = []
    franka = tc.kind in (capi.TASK_REACHING, capi.TASK_BOXSWEEP, capi.TASK_PUSH)
    for r, j in enumerate(trial_indices):
        start, tg = tc.load_trial(j % tc.num_trials())
        targets.append(tg)
        if franka:
            snaps[r], g = tc.init_snapshot(j % tc.num_trials())
            if g is not None:
                goals[r] = g
        else:
            snaps[r] = tc.start_snapshot(start_pos=start)
    targets = np.stack(targets)
    st = np.zeros(B, dtype=np.int32)
    if tr is None:
        tr = capi.Traj(tc, horizon, B)
    if tc.kind in (capi.TASK_BOXSWEEP, capi.TASK_PUSH):
        setup = tc.SETUP_HORIZON if tc.kind == capi.TASK_PUSH else 0
        snap0, U, st = tr.init_controls_ee(snaps, goals, pre_steps=1, setup_horizon=setup)
    else:
        snap0 = np.zeros_like(snaps)
        capi.check(capi.lib().tokp_forward_worlds(tc.model.h, capi._d(snaps), capi._d(snap0), B, 1, -1, 0.0, capi._i(st)))
        if tc.kind == capi.TASK_REACHING:
            U, st2 = tr.init_controls_gravcomp(snap0)
            st = st | st2
        else:
            U = np.zeros((B, horizon, tc.m))
    if st.any():
        raise RuntimeError("prepare_trials: step warnings %s" % sorted(set(int(x) for x in st if x)))
    return snap0, U, targets


def run_openloop_sweep(tc, horizon, num_trials=100, rank=0, world=1, dist=None, device=None, max_iter=10, min_iter=3,
                       lam=0.01, init_controls=None, batch=None):
    """GenDataOpenloopOptimisation on the device: this rank's trials as ONE tokp_traj batch (all trials of the rank
    advance in lock-step through tokp_optimise), then the gather.  Returns the [num_trials, RECORD_SIZE] table."""
    import time

    from . import capi
    mine = shard_trials(num_trials, rank, world)
    B = len(mine)
    recs = []
    if B:
        tr = capi.Traj(tc, horizon, B)
        snaps, U, targets = prepare_trials(tc, mine, horizon, tr)
        if init_controls is not None:
            U = init_controls(tc, mine, horizon)
        for r in range(B):
            tr.set_targets(r, targets[r])
        t0 = time.perf_counter()
        _, stats = tr.optimise(snaps, U, max_iter, min_iter, lam=lam)
        wall_ms = 1e3 * (time.perf_counter() - t0)
        jac = tr.counters()[1] / 1e6
        for r in range(B):
            ini, fin, its = stats[r, 0], stats[r, 1], stats[r, 2]
            recs.append([1.0 - fin / ini if ini != 0 else 0.0, wall_ms / B, its, float(tc.n), stats[r, 4], stats[r, 5],
                         stats[r, 6], stats[r, 7], jac / B])
    local = pack_records(mine, recs, num_trials, world)
    return gather_records(local, num_trials, dist=dist, device=device)


def save_trajec_information(project_parent_path, A, B, states, controls, file_prefix=""):
    """FileHandler::SaveTrajecInformation (src/FileHandler/FileHandler.cpp:312-383): the four CSV files the reference's
    analysis scripts read - `<parent>/savedTrajecInfo<prefix>/{A_matrices,B_matrices,states,controls}.csv`.  One line
    per time step, matrices flattened row-major, every value followed by a comma, numbers in the default ostream format
    (6 significant digits, "%g"), and - as in the reference - only the first `horizon - 1` steps are written.
    A [T][2 dof][2 dof], B [T][2 dof][num_ctrl], states [T][2 dof], controls [T][num_ctrl] (mathematical indexing)."""
    A = np.asarray(A, dtype=np.float64)
    B = np.asarray(B, dtype=np.float64)
    states = np.asarray(states, dtype=np.float64)
    controls = np.asarray(controls, dtype=np.float64)
    horizon = A.shape[0]
    root = project_parent_path + "/savedTrajecInfo" + file_prefix
    os.makedirs(root, exist_ok=True)

    def write(name, rows):
        with open(os.path.join(root, name), "w") as f:
            for r in rows:
                f.write("".join("%g," % v for v in np.ravel(r)) + "\n")

    write("A_matrices.csv", A[:horizon - 1])
    write("B_matrices.csv", B[:horizon - 1])
    write("states.csv", states[:horizon - 1])
    write("controls.csv", controls[:horizon - 1])
    return root


def dump_trajectory(tr, b, project_parent_path, file_prefix=""):
    """save_trajec_information for trajectory b of a device batch: A / B as interpolated by the last
    generate_derivatives, the nominal states and controls."""
    T, nx, m = tr.T, tr.nx, tr.m
    A = tr.get(b, "A").reshape(T, nx, nx).transpose(0, 2, 1)  # device matrices are column-major
    B = tr.get(b, "B").reshape(T, m, nx).transpose(0, 2, 1)
    X = tr.get(b, "X_old").reshape(T + 1, nx)[:T]
    U = tr.get(b, "U_old").reshape(T, m)
    return save_trajec_information(project_parent_path, A, B, X, U, file_prefix)
