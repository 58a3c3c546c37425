#!/usr/bin/env python3
"""Open-loop iLQR sweep over the 100 TestTasks scenes (GenDataOpenloopOptimisation: Optimise(max 10, min 3, T),
lambda 0.01) on the device, summarised like BASELINE.md section 1 so that the means can be put beside the reference's
published PaperData rows (statistical pin of SURVEY 8(c)-6: iteration count, % derivatives, cost reduction).
usage: openloop_vs_paperdata.py [task] [T] [out.json]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from trajoptkp_b200 import capi, sweep, tasks  # noqa: E402

task = sys.argv[1] if len(sys.argv) > 1 else "acrobot"
T = int(sys.argv[2]) if len(sys.argv) > 2 else 500
out = sys.argv[3] if len(sys.argv) > 3 else None
# published rows (PaperData/iLQR_C++_Data/iLQR/acrobot_openloop_*): iters, % derivs, cost reduction
PUBLISHED = {"acrobot": {"SI_1": (7.49, 100.00, 0.768), "SI_5": (7.06, 20.20, 0.764), "SI_1000": (4.15, 0.40, 0.359),
                         "AJ_1_50": (6.08, 28.16, 0.731), "VC_1_200": (7.62, 36.25, 0.770)}}
METHODS = [("SI_1", capi.KP_SET_INTERVAL, 1, 1), ("SI_5", capi.KP_SET_INTERVAL, 5, 5),
           ("SI_1000", capi.KP_SET_INTERVAL, 1000, 1000), ("AJ_1_50", capi.KP_ADAPTIVE_JERK, 1, 50),
           ("VC_1_200", capi.KP_VELOCITY_CHANGE, 1, 200)]
rows = []
for tag, kp, mn, mx in METHODS:
    tc = tasks.TaskConfig(task, kp_method=kp, min_N=mn, max_N=mx)
    recs = sweep.run_openloop_sweep(tc, T, num_trials=100)
    pub = PUBLISHED.get(task, {}).get(tag)
    row = {"task": task, "method": tag, "T": T, "iters": float(recs[:, 2].mean()), "pct_derivs": float(recs[:, 4].mean()),
           "cost_reduction": float(recs[:, 0].mean()), "opt_ms_per_trial_batched": float(recs[:, 1].mean()),
           "published_iters_pct_costred": pub,
           "per_trial": {"cost_reduction": [float(x) for x in recs[:, 0]], "iterations": [int(x) for x in recs[:, 2]],
                         "pct_derivs": [float(x) for x in recs[:, 4]]}}
    rows.append(row)
    print(json.dumps({k: v for k, v in row.items() if k != "per_trial"}))
if out:
    json.dump(rows, open(out, "w"), indent=1)
