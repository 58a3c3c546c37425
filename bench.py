#!/usr/bin/env python3
"""bench.py - FD A/B Jacobians/s and iLQR iteration time for batched MPC replans (BASELINE.json's metric).

A "step" is one MPC replan of every trajectory in the batch = iLQR::Optimise(max_iter = 1): nominal rollout,
key-points, FD dynamics derivatives at the key-points, interpolation, cost derivatives, Riccati backward pass,
6-alpha line-search rollouts and acceptance.  `value` times the step with inputs resident in HBM; `e2e` times
the same call through the C ABI with host buffers (H2D of start snapshots + controls, D2H of controls + stats).

The headline line is walker_run (BASELINE configs[4], velocity_change key-points, H = 80, 592 replans per GPU); the
same line carries `targets.box_sweep` - north_star's second target config (BASELINE configs[2]: iterative_error
key-points, the task's MPC horizon, the reference's Jacobian end-effector initial controls) with its own value / e2e /
roofline / cpu_baseline.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl tokp|reference] [--task T] [--batch B] [--horizon H]
  python bench.py --sweep push_ncl [--gpus N]     # the 100-trial open-loop sweep, strong-scaled over the ranks
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

KP_METHOD = {"velocity_change": 4, "set_interval": 0, "adaptive_jerk": 1, "iterative_error": 3}
# per task: (key-point method BASELINE.json names, horizon, replans per GPU).  The two target configs are sized to
# whole waves of the rollout kernel: 148 SMs x 24 resident walker worlds = 3552 = 6 alphas x 592 replans;
# 148 x 6 resident box_sweep worlds = 888, so 6 x 296 = two full waves.
DEFAULTS = {"walker_run": ("velocity_change", 80, 592), "box_sweep": ("iterative_error", 100, 296),
            "acrobot": ("set_interval", 500, 512), "reaching": ("adaptive_jerk", 50, 512),
            "push_ncl": ("adaptive_jerk", 50, 256), "push_lcl": ("iterative_error", 100, 64),
            "push_mcl": ("set_interval", 50, 16)}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="tokp", choices=["tokp", "reference"])
    ap.add_argument("--task", default="walker_run")
    ap.add_argument("--batch", type=int, default=0, help="replans per GPU (0 = the task's default)")
    ap.add_argument("--horizon", type=int, default=0)
    ap.add_argument("--kp", default="", choices=[""] + sorted(KP_METHOD))
    ap.add_argument("--cpu-replans", type=int, default=0, help="replans per CPU sample step (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-targets", action="store_true", help="skip the targets.box_sweep block")
    ap.add_argument("--sweep", default="", help="run the 100-trial open-loop sweep of this task (strong scaling)")
    ap.add_argument("--sweep-horizon", type=int, default=0)
    return ap.parse_args()


def task_settings(args, task):
    kp, T, B = DEFAULTS[task]
    if task == args.task:
        kp = args.kp or kp
        T = args.horizon or T
        B = args.batch or B
    return kp, T, B


def start_states(task, kp, B, rank):
    """Synthetic inputs: scene i of TestTasks/<task> as InitialiseSystemToStartState leaves it, start
    state perturbed with a seeded RNG (joints +-0.02 rad, body x, y +-0.005 m, state velocities N(0, 0.01))."""
    from trajoptkp_b200 import capi, tasks
    tc = tasks.TaskConfig(task, kp_method=KP_METHOD[kp])
    m = tc.model
    franka = tc.kind in (capi.TASK_REACHING, capi.TASK_BOXSWEEP, capi.TASK_PUSH)
    snaps = np.zeros((B, m.S))
    goals = np.zeros((B, 2))
    targets = np.zeros((B, len(tc.target)))
    for j in range(B):
        gj = rank * B + j                      # global replan index: weak scaling, fixed work per GPU
        scene, p = gj % 100, gj // 100
        start, tg = tc.load_trial(scene)
        targets[j] = tg
        rng = np.random.default_rng(1000 * 7 + 10 * scene + p)
        nj = getattr(tc, "num_robot_dofs", tc.n)
        dq = rng.uniform(-0.02, 0.02, tc.n)
        if tc.n > nj:
            dq[nj:] = 0.0
            for k in range(nj, tc.n, 3):
                dq[k:k + 2] = rng.uniform(-0.005, 0.005, 2)
        v = 0.01 * rng.standard_normal(tc.n)
        if franka:
            s, g = tc.init_snapshot(scene)
            f = m.split(s)
            f["qpos"][tc.qpos_adr] += dq
            f["qvel"][tc.dof_adr] = v
            snaps[j] = m.snapshot(qpos=f["qpos"], qvel=f["qvel"])
            if g is not None:
                goals[j] = g
        else:
            snaps[j] = tc.start_snapshot(start_pos=start + dq, qvel=v)
    return tc, snaps, goals, targets


class ClockSampler(threading.Thread):
    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu, self.samples, self.reasons, self.stop_flag, self.max_mhz = gpu, [], set(), False, None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0]))
                self.max_mhz = float(out[1])
                for nm, v in zip(names, out[2:]):
                    if v.strip().lower().startswith("active"):
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def initial_controls_device(tc, tr, snaps, goals, T):
    """The task's CreateInitOptimisationControls on the device, outside the timed region (the reference also runs it once
    per trial before Optimise): zeros (acrobot, walker), gravity compensation (reaching), Jacobian end-effector follower
    after the setup motion (box_sweep, push)."""
    from trajoptkp_b200 import capi
    if tc.kind in (capi.TASK_BOXSWEEP, capi.TASK_PUSH):
        setup = tc.SETUP_HORIZON if tc.kind == capi.TASK_PUSH else 0
        snap0, U, st = tr.init_controls_ee(snaps, goals, pre_steps=1, setup_horizon=setup)
        return snap0, U, "Jacobian end-effector follower (tokp_init_controls_ee: PushBaseClass.cpp:139-248)"
    if tc.kind == capi.TASK_REACHING:
        U, st = tr.init_controls_gravcomp(snaps)
        return snaps, U, "gravity compensation (tokp_init_controls_gravcomp: Reaching.cpp:269-296)"
    return snaps, np.zeros((len(snaps), T, tc.m)), "zeros (ModelTranslator.cpp:1783-1797)"


def oracle_task_of(tc, target):
    from oracle import oracle
    return oracle.Task(tc.kind, tc.qpos_adr, tc.dof_adr, tc.ctrl_adr, target, tc.weight, tc.weight_terminal,
                       tc.kp_method, tc.min_N, tc.max_N, tc.jerk_thr, tc.vel_thr, tc.iter_thr, aux=tc.aux)


def initial_controls_oracle(tc, om, snaps, goals, targets, T):
    """The same with the oracle, for the CPU arm (untimed input preparation)."""
    from oracle import oracle
    B = len(snaps)
    if tc.kind in (3, 4):
        setup = tc.SETUP_HORIZON if tc.kind == 4 else 0
        s0, U = np.zeros_like(snaps), np.zeros((B, T, tc.m))
        for j in range(B):
            s0[j], U[j], _, _ = oracle.init_sequence_ee(om, oracle_task_of(tc, targets[j]), snaps[j], goals[j], setup, T)
        return s0, U
    U = np.zeros((B, T, tc.m))
    if tc.kind == 2:
        S, nu = tc.model.S, tc.model.nu
        d = oracle.Data(om)
        for j in range(B):
            d.load(snaps[j])
            for t in range(T):
                d.forward()
                bias = d.get("qfrc_bias")
                s = d.save()
                s[S - nu + tc.ctrl_adr] = bias[tc.dof_adr]
                d.load(s)
                d.step()
                U[j, t] = bias[tc.dof_adr]
    return snaps, U


def cpu_arm(tc, T, snaps, goals, targets, steps, warmup, cpu_replans=0, want_flops=True, budget_s=3.0):
    """The reference's CPU implementation of the path, restated (oracle/): multithreaded FD workers + 6-thread line
    search, all host cores, one replan after the other - the reference's way of using a CPU (a CPU that ran one replan
    per core instead would scale better).  Each step = a bounded sample of the workload."""
    from oracle import oracle
    cores = os.cpu_count() or 1
    om = oracle.Model(tc.model.blob())
    nprep = min(len(snaps), 32 if tc.kind in (3, 4) else len(snaps))   # the oracle's N2 generator costs ~0.1 s per trial
    snap0, U = initial_controls_oracle(tc, om, snaps[:nprep], goals[:nprep], targets[:nprep], T)
    ilqr = None
    times, jac = [], []

    def one_replan(j):
        nonlocal ilqr
        ilqr = oracle.ILQR(om, oracle_task_of(tc, targets[j]), T, threads=max(1, cores - 1))
        ilqr.set_lambda(0.1)
        ilqr.optimise(snap0[j], U[j], 1, 1)
        rows = ilqr.keypoints()
        return sum(len(set(r)) for r in rows) / tc.n

    n_rep = cpu_replans
    if n_rep <= 0:
        t0 = time.perf_counter()
        one_replan(0)
        dt = time.perf_counter() - t0
        n_rep = int(max(1, min(nprep, round(budget_s / max(dt, 1e-4)))))   # ~budget_s of CPU work per step
    for s in range(warmup + steps):
        t0 = time.perf_counter()
        nj = 0.0
        for j in range(n_rep):
            nj += one_replan((s * n_rep + j) % nprep)
        dt = time.perf_counter() - t0
        if s >= warmup:
            times.append(dt)
            jac.append(nj)
    flops_per_jac = None
    if want_flops:
        # ALGORITHMIC flops per Jacobian: instrumented oracle run of the reference schedule (stage skipping included)
        saved = ilqr.get("saved").reshape(T + 1, -1)
        ot = oracle_task_of(tc, targets[0])
        fl = [oracle.flops_dynamics_derivs(om, ot, saved[t], list(range(tc.n))) for t in range(1, T, max(1, T // 8))]
        flops_per_jac = float(np.mean(fl))
    return dict(value=float(np.sum(jac) / np.sum(times)), ms_per_step=1e3 * float(np.mean(times)), replans_per_step=n_rep,
                cores=cores, flops_per_jacobian=flops_per_jac, jacobians_per_step=float(np.mean(jac)))


def source_hash():
    """Hash of the kernel sources: an ncu traffic capture only describes the build it was taken on."""
    h = hashlib.sha1()
    d = os.path.join(ROOT, "trajoptkp_b200", "csrc")
    for fn in sorted(os.listdir(d)):
        if fn.endswith((".cu", ".cuh", ".h", ".cpp")):
            h.update(open(os.path.join(d, fn), "rb").read())
    return h.hexdigest()[:12]


def workload_config(task, kp, T, B):
    return {"workload": "%s MPC replan batch (iLQR::Optimise max_iter=1), %s key-points" % (task, kp),
            "task": task, "horizon": T, "replans_per_gpu": B, "keypoint_method": kp,
            "l2_policy": "per-step working set (FD world outputs + snapshots + A/B) exceeds L2; no explicit flush"}


def reference_line(args, task):
    kp, T, B = task_settings(args, task)
    tc, snaps, goals, targets = start_states(task, kp, min(B, 64), 0)
    r = cpu_arm(tc, T, snaps, goals, targets, args.steps, args.warmup, args.cpu_replans, want_flops=False)
    sample = "%d replans (H=%d) per step on %d host threads; CPU restatement of the MuJoCo-based reference" % (
        r["replans_per_step"], T, r["cores"])
    return {"metric": "fd_jacobians_per_sec", "value": r["value"], "unit": "jacobians/s", "ms_per_step": r["ms_per_step"],
            "cpu_baseline": {"value": r["value"], "unit": "jacobians/s", "cores": r["cores"], "kind": "port", "sample": sample},
            "e2e": {"value": r["value"], "unit": "jacobians/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "ilqr_iter_ms": r["ms_per_step"] / r["replans_per_step"],
            "config": workload_config(task, kp, T, B)}


def run_task(args, task, rank, world, local, dist, torch, headline):
    """One bench line (dict) for `task`; collectives only after the timed regions."""
    from trajoptkp_b200 import capi
    kp, T, B = task_settings(args, task)
    tc, snaps, goals, targets = start_states(task, kp, B, rank)
    tr = capi.Traj(tc, T, B)
    for b in range(B):
        tr.set_targets(b, targets[b])
    snap0, U, how = initial_controls_device(tc, tr, snaps, goals, T)
    # e2e inputs live in pinned host memory (the C ABI takes plain host pointers; pinning is the caller's choice)
    pin_s = torch.empty(snap0.shape, dtype=torch.float64, pin_memory=True)
    pin_u = torch.empty(U.shape, dtype=torch.float64, pin_memory=True)
    pin_s.numpy()[...] = snap0
    pin_u.numpy()[...] = U
    snap0, U = pin_s.numpy(), pin_u.numpy()
    config = workload_config(task, kp, T, B)
    config["initial_controls"] = how

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    tr.optimise(snap0, U, 1, 1, lam=0.1)         # first call uploads the inputs (also the first warm-up step)
    for _ in range(max(0, args.warmup - 1)):
        tr.optimise_resident(1, 1, lam=0.1, want_stats=False)
    sampler = ClockSampler(local)
    sampler.start()
    # ---- timed region 1: device-resident steps ---------------------------------------------------------------------
    fd0 = tr.fd_kernel_ms()
    l0 = capi.lib().tokp_launch_count()
    barrier()
    t0 = time.perf_counter()
    jac = 0.0
    phase = np.zeros(3)
    step_ms = []
    for _ in range(args.steps):
        ts = time.perf_counter()
        stats = tr.optimise_resident(1, 1, lam=0.1)
        jac += tr.counters()[1] / 1e6
        phase += stats[:, 5:8].max(axis=0)
        step_ms.append(1e3 * (time.perf_counter() - ts))
    barrier()
    dt = time.perf_counter() - t0
    if os.environ.get("TOKP_BENCH_VERBOSE"):
        print("rank %d %s resident steps ms: %s  derivs/bp/fp of last: %s  reruns %d" % (
            rank, task, ["%.1f" % x for x in step_ms], stats[:, 5:8].max(axis=0).tolist(), tr.counters()[3]), file=sys.stderr)
    fd1 = tr.fd_kernel_ms()
    l1 = capi.lib().tokp_launch_count()
    # ---- timed region 2: end to end through the C ABI with host buffers -------------------------------------------------
    barrier()
    t0 = time.perf_counter()
    jac_e = 0.0
    for _ in range(args.steps):
        Uo, st = tr.optimise(snap0, U, 1, 1, lam=0.1)
        jac_e += tr.counters()[1] / 1e6
    barrier()
    dt_e = time.perf_counter() - t0
    # ---- single-trajectory replan latency (B = 1, host buffers in, controls out), rank 0 only ---------------------------
    lat_ms = None
    if rank == 0:
        tr1 = capi.Traj(tc, T, 1)
        tr1.set_targets(0, targets[0])
        for _ in range(3):
            tr1.optimise(snap0[:1], U[:1], 1, 1, lam=0.1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(10):
            tr1.optimise(snap0[:1], U[:1], 1, 1, lam=0.1)
        torch.cuda.synchronize()
        lat_ms = 1e3 * (time.perf_counter() - t0) / 10
        del tr1
    sampler.stop_flag = True
    sampler.join(timeout=2)
    h2d = snap0.nbytes + U.nbytes + 8 * B
    d2h = Uo.nbytes + st.nbytes
    tt = torch.tensor([dt, dt_e, jac, jac_e], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = tt.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = tt.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dt, dt_e = float(mx[0]), float(mx[1])
        jac, jac_e = float(sm[2]), float(sm[3])
        # the data path has no collective; this gather of per-rank result records is the only exchange
        rec = torch.tensor(stats[:, :3].mean(axis=0), dtype=torch.float64, device="cuda")
        gathered = [torch.zeros_like(rec) for _ in range(world)]
        dist.all_gather(gathered, rec)
    if rank != 0:
        return None
    fd_launches = fd1[2] - fd0[2]
    fd_ms = (fd1[1] - fd0[1]) / max(fd_launches, 1)
    line = {"metric": "fd_jacobians_per_sec", "value": jac / dt, "unit": "jacobians/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
            "ilqr_iter_ms": 1e3 * dt / args.steps, "ilqr_iter_ms_per_replan": 1e3 * dt / args.steps / (B * world),
            "phase_ms": {"derivs": phase[0] / args.steps, "backward": phase[1] / args.steps, "forward": phase[2] / args.steps},
            "jacobians_per_step": jac / args.steps,
            "replan_latency_ms_b1": lat_ms,   # one trajectory alone on the GPU (serial rollouts / Riccati: latency-bound)
            "e2e": {"value": jac_e / dt_e, "unit": "jacobians/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": 1e3 * dt_e / args.steps},
            "gpu_launches": int(l1 - l0), "clocks": sampler.summary(),
            "fd_kernel": {"name": "FD world kernels (k_fd_worlds + k_fd_dep_lanes), CUDA events around them on the launching stream",
                          "avg_ms": fd_ms, "launches": int(fd_launches),
                          "worlds_last": int(tr.counters()[0]), "worlds_rerun_full_capacity_last": int(tr.counters()[3]),
                          "share_of_step": fd_ms * fd_launches / (1e3 * dt) if dt > 0 else None}}
    if headline and task == "walker_run":
        # context only (never a ratio input): the reference's own published per-replan times for walker_run synchronous
        # MPC at H = 80 (PaperData, hardware not stated)
        line["published_reference_ms_per_replan_h80"] = {"SI_1": 8.474, "SI_5": 5.364, "AJ_2_20": 5.708, "VC_2_20": 5.674,
                                                         "IE_2_20": 16.570}
    cpu = None
    if not args.no_cpu_baseline:
        cpu = cpu_arm(tc, T, snaps, goals, targets, steps=2, warmup=1, budget_s=3.0 if headline else 2.0)
        line["cpu_baseline"] = {"value": cpu["value"], "unit": "jacobians/s", "cores": cpu["cores"], "kind": "port",
                                "sample": "%d replans (H=%d) per step, 2 steps, %d host threads, one replan after the other as the "
                                          "reference does (a CPU running one replan per core would scale better); CPU restatement "
                                          "of the MuJoCo-based reference (MuJoCo itself is not installable here)" % (
                                              cpu["replans_per_step"], T, cpu["cores"]),
                                "ilqr_iter_ms_per_replan": cpu["ms_per_step"] / cpu["replans_per_step"]}
    # roofline of the FD world kernels: algorithmic fp64 flops (oracle-instrumented, stage skipping included) / time
    fpj = cpu["flops_per_jacobian"] if cpu else None
    jac_per_launch = (jac / world) / max(fd_launches, 1)
    SD = 1 + tr.model.nq + 3 * tr.model.nv + tr.model.nu
    bytes_per_jac = 8 * SD + 8 * tr.nx * (tr.nx + tr.m)
    if fpj and fd_ms > 0:
        peaks = capi.measure_peaks()
        ach = fpj * jac_per_launch / (fd_ms * 1e-3) / 1e12
        traffic = None
        try:  # dram bytes per FD launch from an ncu --set full capture: only valid for the build it was taken on
            e = json.load(open(os.path.join(ROOT, "profiles", "fd_traffic.json"))).get(task)
            if e and e.get("batch") == B and e.get("horizon") == T and e.get("kp") == kp and e.get("source_hash") == source_hash():
                traffic = e["dram_bytes_per_launch"]
        except Exception:
            pass
        hbm_peak, hbm_src = peaks["copy_gbs"], "measured live (copy kernel)"
        try:  # the driver-written figure wins when present
            hbm_peak, hbm_src = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json"
        except Exception:
            pass
        line["roofline"] = {"bound": "fp64", "achieved": ach, "peak": peaks["fp64_tflops"], "unit": "TFLOP/s",
                            "frac": ach / peaks["fp64_tflops"], "traffic": traffic,
                            "peak_source": "measured live (DFMA loop, tokp_measure_peaks); MEASURED_PEAKS.json has no FP64 figure",
                            "flops_per_jacobian": fpj,
                            "hbm": {"achieved_gbs": bytes_per_jac * jac_per_launch / (fd_ms * 1e-3) / 1e9,
                                    "peak_gbs": hbm_peak, "peak_source": hbm_src, "live_copy_gbs": peaks["copy_gbs"],
                                    "bytes_per_jacobian": bytes_per_jac}}
    return line


def run_sweep(args, rank, world, local, dist, torch):
    """GenTestingData::GenDataOpenloopOptimisation (src/GenTestingData.cpp:136-259): the 100 TestTasks trials of a task,
    split over the ranks (trial j -> rank j mod world: strong scaling), initial controls + Optimise(max 10, min 3) per
    trial, ONE all_gather of the result records - inside the timed region."""
    from trajoptkp_b200 import sweep, tasks
    tc = tasks.TaskConfig(args.sweep)
    T = args.sweep_horizon or tc.openloop_horizon
    dev = torch.device("cuda", local)
    d = dist if world > 1 else None

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    sweep.run_openloop_sweep(tc, T, num_trials=100, rank=rank, world=world, dist=d, device=dev)  # warm-up
    barrier()
    t0 = time.perf_counter()
    recs = sweep.run_openloop_sweep(tc, T, num_trials=100, rank=rank, world=world, dist=d, device=dev)
    barrier()
    dt = time.perf_counter() - t0
    tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    if rank != 0:
        return None
    dt = float(tt[0])
    return {"metric": "openloop_sweep_trials_per_sec", "value": 100.0 / dt, "unit": "trials/s", "n_gpus": world, "steps": 1,
            "warmup": 1, "ms_per_step": 1e3 * dt, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "TestTasks fixtures (100 trials)",
            "config": {"workload": "%s open-loop sweep: 100 trials, Optimise(max 10, min 3), T = %d" % (args.sweep, T),
                       "task": args.sweep, "horizon": T, "trials_per_rank": (100 + world - 1) // world},
            "sweep_seconds": dt, "jacobians_per_sec": float(recs[:, 8].sum() / dt),
            "mean_cost_reduction": float(recs[:, 0].mean()), "mean_iterations": float(recs[:, 2].mean())}


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank != 0:
            return
        import __graft_entry__ as g
        g.build()
        line = reference_line(args, args.task)
        line.update({"impl": "reference", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                     "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                     "gpu_launches": 0})
        if args.task == "walker_run" and not args.no_targets:
            line["targets"] = {"box_sweep": reference_line(args, "box_sweep")}
        print(json.dumps(line))
        return

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import __graft_entry__ as g
    if rank == 0:
        g.build()
    if world > 1:
        dist.barrier()
    if args.sweep:
        line = run_sweep(args, rank, world, local, dist, torch)
    else:
        line = run_task(args, args.task, rank, world, local, dist, torch, headline=True)
        if args.task == "walker_run" and not args.no_targets:
            tb = run_task(args, "box_sweep", rank, world, local, dist, torch, headline=False)
            if line is not None:
                line["targets"] = {"box_sweep": tb}
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
