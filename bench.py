#!/usr/bin/env python3
"""bench.py - FD A/B Jacobians/s and iLQR iteration time for batched walker_run MPC replans.

A "step" is one MPC replan of every trajectory in the batch = iLQR::Optimise(max_iter = 1): nominal rollout,
key-points, FD dynamics derivatives at the key-points, interpolation, cost derivatives, Riccati backward pass,
6-alpha line-search rollouts and acceptance.  `value` times the step with inputs resident in HBM; `e2e` times
the same call through the C ABI with host buffers (H2D of start snapshots + controls, D2H of controls + stats).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl tokp|reference] [--batch B] [--horizon H]
"""
import argparse
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="tokp", choices=["tokp", "reference"])
    ap.add_argument("--task", default="walker_run")
    ap.add_argument("--batch", type=int, default=512, help="replans per GPU")
    ap.add_argument("--horizon", type=int, default=80)
    ap.add_argument("--kp", default="velocity_change", choices=sorted(KP_METHOD))
    ap.add_argument("--cpu-replans", type=int, default=0, help="replans per CPU sample step (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def workload(args, rank, world):
    """Synthetic inputs: scene i of TestTasks/<task>, start state perturbed with a seeded RNG."""
    from trajoptkp_b200 import tasks
    tc = tasks.TaskConfig(args.task, kp_method=KP_METHOD[args.kp])
    B, T = args.batch, args.horizon
    snaps = np.zeros((B, tc.model.S))
    targets = np.zeros((B, len(tc.target)))
    for j in range(B):
        gj = rank * B + j                      # global replan index: weak scaling, fixed work per GPU
        scene, p = gj % 100, gj // 100
        start, tg = tc.load_trial(scene)
        rng = np.random.default_rng(1000 * 7 + 10 * scene + p)
        # joints +-0.02 rad, body x, y +-0.005 m (z untouched), velocities N(0, 0.01)
        nj = getattr(tc, "num_robot_dofs", tc.n)
        dq = rng.uniform(-0.02, 0.02, tc.n)
        if tc.n > nj:
            dq[nj:] = 0.0
            for k in range(nj, tc.n, 3):
                dq[k:k + 2] = rng.uniform(-0.005, 0.005, 2)
        q = start + dq
        v = 0.01 * rng.standard_normal(tc.n)
        snaps[j] = tc.start_snapshot(start_pos=q, qvel=v)
        targets[j] = tg
    U = np.zeros((B, T, tc.m))                 # walker initial controls are zeros (ModelTranslator.cpp:1783-1797)
    return tc, snaps, targets, U


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


def cpu_arm(args, tc, snaps, targets, U, steps, warmup, want_flops=True):
    """The reference's CPU implementation of the path, restated (oracle/): multithreaded FD workers + 6-thread line
    search, all host cores.  Each step = a bounded sample of the workload (cpu_replans replans)."""
    from oracle import oracle
    cores = os.cpu_count() or 1
    om = oracle.Model(tc.model.blob())
    T = args.horizon
    n_rep = args.cpu_replans
    if tc.kind in (2, 3) and not np.any(U):
        # reaching: gravity-compensation initial controls (Reaching.cpp:269-296), restated with the oracle, untimed
        U = np.array(U, copy=True)
        S, nu = tc.model.S, tc.model.nu
        d = oracle.Data(om)
        for j in range(len(snaps)):
            d.load(snaps[j])
            for t in range(T):
                d.forward()
                bias = d.get("qfrc_bias")
                s = d.save()
                s[S - nu + tc.ctrl_adr] = bias[tc.dof_adr]
                d.load(s)
                d.step()
                U[j, t] = bias[tc.dof_adr]
    ilqr = None
    times, jac = [], []
    flops_per_jac = None
    total = warmup + steps

    def one_replan(j):
        nonlocal ilqr
        ot = oracle.Task(tc.kind, tc.qpos_adr, tc.dof_adr, tc.ctrl_adr, targets[j], tc.weight, tc.weight_terminal,
                         tc.kp_method, tc.min_N, tc.max_N, tc.jerk_thr, tc.vel_thr, tc.iter_thr, aux=tc.aux)
        ilqr = oracle.ILQR(om, ot, T, threads=max(1, cores - 1))
        ilqr.set_lambda(0.1)
        ilqr.optimise(snaps[j], U[j], 1, 1)
        rows = ilqr.keypoints()
        return sum(len(set(r)) for r in rows) / tc.n

    if n_rep <= 0:
        t0 = time.perf_counter()
        one_replan(0)
        dt = time.perf_counter() - t0
        n_rep = int(max(1, min(len(snaps), round(3.0 / max(dt, 1e-4)))))   # ~3 s of CPU work per step
    for s in range(total):
        t0 = time.perf_counter()
        nj = 0.0
        for j in range(n_rep):
            nj += one_replan((s * n_rep + j) % len(snaps))
        dt = time.perf_counter() - t0
        if s >= warmup:
            times.append(dt)
            jac.append(nj)
    if want_flops:
        # ALGORITHMIC flops per Jacobian: instrumented oracle run of the reference schedule (stage skipping included)
        saved = ilqr.get("saved").reshape(T + 1, -1)
        ot = oracle.Task(tc.kind, tc.qpos_adr, tc.dof_adr, tc.ctrl_adr, targets[0], tc.weight, tc.weight_terminal,
                         tc.kp_method, tc.min_N, tc.max_N, tc.jerk_thr, tc.vel_thr, tc.iter_thr, aux=tc.aux)
        fl = [oracle.flops_dynamics_derivs(om, ot, saved[t], list(range(tc.n))) for t in range(1, T, max(1, T // 8))]
        flops_per_jac = float(np.mean(fl))
    ms = 1e3 * float(np.mean(times))
    value = float(np.sum(jac) / np.sum(times))
    return dict(value=value, ms_per_step=ms, replans_per_step=n_rep, cores=cores, flops_per_jacobian=flops_per_jac,
                jacobians_per_step=float(np.mean(jac)))


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    config = {"workload": "%s MPC replan batch (iLQR::Optimise max_iter=1), %s key-points" % (args.task, args.kp),
              "task": args.task, "horizon": args.horizon, "replans_per_gpu": args.batch, "keypoint_method": args.kp,
              "l2_policy": "per-step working set (FD world outputs + snapshots + A/B) exceeds L2; no explicit flush"}

    if args.impl == "reference":
        if rank != 0:
            return
        import __graft_entry__ as g
        g.build()
        tc, snaps, targets, U = workload(args, 0, 1)
        r = cpu_arm(args, tc, snaps, targets, U, args.steps, args.warmup, want_flops=False)
        sample = "%d replans (H=%d) per step on %d host threads; CPU restatement of the MuJoCo-based reference" % (
            r["replans_per_step"], args.horizon, r["cores"])
        line = {"impl": "reference", "metric": "fd_jacobians_per_sec", "value": r["value"], "unit": "jacobians/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": config,
                "cpu_baseline": {"value": r["value"], "unit": "jacobians/s", "cores": r["cores"], "kind": "port", "sample": sample},
                "e2e": {"value": r["value"], "unit": "jacobians/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "ilqr_iter_ms": r["ms_per_step"] / r["replans_per_step"], "gpu_launches": 0}
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
    from trajoptkp_b200 import capi
    tc, snaps, targets, U = workload(args, rank, world)
    B, T = args.batch, args.horizon
    # e2e inputs live in pinned host memory (the C ABI takes plain host pointers; pinning is the caller's choice)
    pin_s = torch.empty(snaps.shape, dtype=torch.float64, pin_memory=True)
    pin_u = torch.empty(U.shape, dtype=torch.float64, pin_memory=True)
    pin_s.numpy()[...] = snaps
    pin_u.numpy()[...] = U
    snaps, U = pin_s.numpy(), pin_u.numpy()
    tr = capi.Traj(tc, T, B)
    for b in range(B):
        tr.set_targets(b, targets[b])
    if tc.kind in (capi.TASK_REACHING, capi.TASK_BOXSWEEP):
        # pandaReaching::CreateInitOptimisationControls (Reaching.cpp:269-296): gravity-compensation rollout, on the
        # device, outside the timed region (the reference also generates it once per trial before Optimise)
        U[...] = tr.init_controls_gravcomp(snaps)[0]
        config["initial_controls"] = "gravity compensation (tokp_init_controls_gravcomp)" + (
            "; the reference's Jacobian end-effector path follower (PushBaseClass.cpp:8-248, SURVEY N2) is not built yet"
            if tc.kind == capi.TASK_BOXSWEEP else "")
    launches0 = capi.lib().tokp_launch_count()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # first call uploads the inputs (also the first warm-up step)
    tr.optimise(snaps, U, 1, 1, lam=0.1)
    for _ in range(max(0, args.warmup - 1)):
        tr.optimise_resident(1, 1, lam=0.1, want_stats=False)
    sampler = ClockSampler(local)
    sampler.start()
    # ---- timed region 1: device-resident steps --------------------------------------------------------------
    fd0 = tr.fd_kernel_ms()
    l0 = capi.lib().tokp_launch_count()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    jac = 0.0
    phase = np.zeros(3)
    for _ in range(args.steps):
        stats = tr.optimise_resident(1, 1, lam=0.1)
        jac += tr.counters()[1] / 1e6
        phase += stats[:, 5:8].max(axis=0)
    barrier()
    dt = time.perf_counter() - t0
    fd1 = tr.fd_kernel_ms()
    l1 = capi.lib().tokp_launch_count()
    # ---- timed region 2: end to end through the C ABI with host buffers ----------------------------------------
    barrier()
    t0 = time.perf_counter()
    jac_e = 0.0
    for _ in range(args.steps):
        Uo, st = tr.optimise(snaps, U, 1, 1, lam=0.1)
        jac_e += tr.counters()[1] / 1e6
    barrier()
    dt_e = time.perf_counter() - t0
    # ---- single-trajectory replan latency (B = 1, host buffers in, controls out), rank 0 only ----------------------
    lat_ms = None
    if rank == 0:
        tr1 = capi.Traj(tc, T, 1)
        tr1.set_targets(0, targets[0])
        for _ in range(3):
            tr1.optimise(snaps[:1], U[:1], 1, 1, lam=0.1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(10):
            tr1.optimise(snaps[:1], U[:1], 1, 1, lam=0.1)
        torch.cuda.synchronize()
        lat_ms = 1e3 * (time.perf_counter() - t0) / 10
        del tr1
    sampler.stop_flag = True
    sampler.join(timeout=2)
    h2d = snaps.nbytes + U.nbytes + 8 * B
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
        if world > 1:
            dist.destroy_process_group()
        return

    value = jac / dt
    e2e_value = jac_e / dt_e
    fd_launches = fd1[2] - fd0[2]
    fd_ms = (fd1[1] - fd0[1]) / max(fd_launches, 1)
    peaks = capi.measure_peaks()
    line = {"metric": "fd_jacobians_per_sec", "value": value, "unit": "jacobians/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
            "ilqr_iter_ms": 1e3 * dt / args.steps, "ilqr_iter_ms_per_replan": 1e3 * dt / args.steps / (B * world),
            "phase_ms": {"derivs": phase[0] / args.steps, "backward": phase[1] / args.steps, "forward": phase[2] / args.steps},
            "jacobians_per_step": jac / args.steps,
            "replan_latency_ms_b1": lat_ms,   # one trajectory alone on the GPU (serial rollouts / Riccati: latency-bound)
            # context only (never a ratio input): the reference's own published per-replan times for walker_run
            # synchronous MPC at H = 80 (PaperData, hardware not stated)
            "published_reference_ms_per_replan_h80": {"SI_1": 8.474, "SI_5": 5.364, "AJ_2_20": 5.708, "VC_2_20": 5.674,
                                                       "IE_2_20": 16.570} if args.task == "walker_run" else None,
            "e2e": {"value": e2e_value, "unit": "jacobians/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": 1e3 * dt_e / args.steps},
            "gpu_launches": int(l1 - l0), "clocks": sampler.summary(),
            "fd_kernel": {"name": "k_fd_worlds", "avg_ms": fd_ms, "launches": int(fd_launches),
                          "worlds_last": int(tr.counters()[0]), "worlds_rerun_full_capacity_last": int(tr.counters()[3]),
                          "share_of_step": fd_ms * fd_launches / (1e3 * dt) if dt > 0 else None}}
    cpu = None
    if not args.no_cpu_baseline:
        cpu = cpu_arm(args, tc, snaps, targets, U, steps=2, warmup=1)
        line["cpu_baseline"] = {"value": cpu["value"], "unit": "jacobians/s", "cores": cpu["cores"], "kind": "port",
                                "sample": "%d replans (H=%d) per step, 2 steps, %d host threads; CPU restatement of the "
                                          "MuJoCo-based reference (MuJoCo itself is not installable here)" % (
                                              cpu["replans_per_step"], T, cpu["cores"]),
                                "ilqr_iter_ms_per_replan": cpu["ms_per_step"] / cpu["replans_per_step"]}
    # roofline of the dominant kernel: algorithmic fp64 flops (oracle-instrumented, stage skipping included) / time
    fpj = cpu["flops_per_jacobian"] if cpu else None
    jac_per_launch = (jac / world) / max(fd_launches, 1)
    S, SD = tr.model.S, 1 + tr.model.nq + 3 * tr.model.nv + tr.model.nu
    bytes_per_jac = 8 * SD + 8 * tr.nx * (tr.nx + tr.m)
    if fpj and fd_ms > 0:
        ach = fpj * jac_per_launch / (fd_ms * 1e-3) / 1e12
        traffic = None
        try:  # dram bytes per launch of k_fd_worlds from the committed ncu --set full capture of this workload
            tj = json.load(open(os.path.join(ROOT, "profiles", "fd_traffic.json")))
            if tj.get("task") == args.task and tj.get("batch") == B and tj.get("horizon") == T and tj.get("kp") == args.kp:
                traffic = tj["dram_bytes_per_launch"]
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
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
