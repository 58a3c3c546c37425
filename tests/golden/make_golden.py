#!/usr/bin/env python3
"""Generates tests/golden/*.npz: seeded inputs and the CPU oracle's outputs for every stage of the path.

The reference itself cannot run here (MuJoCo, Eigen, yaml-cpp are not installable), and it ships no
numeric golden vectors.  These fixtures therefore pin the ORACLE (oracle/, a CPU restatement of the reference's
algorithm) at the commit that generated them - a regression pin for both the oracle and the CUDA path - and are
NOT outputs of the real MuJoCo-based reference.  Re-generate with:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import __graft_entry__ as g  # noqa: E402

g.build()
from helpers import make, oracle_model, oracle_task, perturbed_start, random_controls  # noqa: E402
from oracle import oracle  # noqa: E402
from trajoptkp_b200 import capi  # noqa: E402

CASES = {
    # name: (horizon, key-point method, min_N, max_N)
    "acrobot": (60, capi.KP_SET_INTERVAL, 5, 50),
    "walker_run": (40, capi.KP_VELOCITY_CHANGE, 1, 20),
}


def main():
    for name, (T, method, min_N, max_N) in CASES.items():
        tc = make(name, kp_method=method, min_N=min_N, max_N=max_N)
        om, ot = oracle_model(tc), oracle_task(tc)
        m = tc.model
        snap0 = perturbed_start(tc, 7)
        U = random_controls(tc, T, 3)
        il = oracle.ILQR(om, ot, T, threads=1)
        cost = il.rollout(snap0, U)
        il.generate_keypoints()
        rows = il.keypoints()
        il.generate_derivatives()
        counts, idx = oracle.pack_rows(rows)
        out = dict(T=T, method=method, min_N=min_N, max_N=max_N, snap0=snap0, U=U, rollout_cost=cost,
                   saved=il.get("saved").reshape(T + 1, -1), X_old=il.get("X_old"), resid=il.get("resid"),
                   kp_counts=counts, kp_indices=idx[:counts.sum()], pct=il.get("last_percentages"),
                   A=il.get("A"), B=il.get("B"), l_x=il.get("l_x"), l_xx=il.get("l_xx"), l_u=il.get("l_u"),
                   l_uu=il.get("l_uu"))
        il.set_lambda(0.1)
        ok = il.backward()
        out.update(bp_ok=int(ok), K=il.get("K"), k=il.get("k"))
        # a full optimise from the same inputs
        il2 = oracle.ILQR(om, ot, T, threads=1)
        il2.set_lambda(0.1)
        Uopt = il2.optimise(snap0, U, 3, 1)
        sc = il2.scalars()
        out.update(U_opt=Uopt, initial_cost=sc["initial_cost"], final_cost=sc["new_cost"], iterations=sc["num_iterations"])
        # single-step pins: a few worlds, one full step with model solver options
        d = oracle.Data(om)
        steps_in = np.stack([perturbed_start(tc, 100 + i, dq=0.2, dv=0.5) for i in range(6)])
        steps_in[:, m.S - m.nu:] = np.random.default_rng(1).uniform(-1, 1, (6, m.nu))
        steps_out = np.zeros_like(steps_in)
        for i in range(6):
            d.load(steps_in[i])
            for _ in range(10):
                d.step()
            steps_out[i] = d.save()
        out.update(steps_in=steps_in, steps_out=steps_out, nsteps=10)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, "cost", cost, "final", sc["new_cost"], "kp rows", int((counts > 0).sum()))


if __name__ == "__main__":
    main()
