#!/usr/bin/env python3
"""Synchronous walker_run MPC (GenDataMPCHorizons: 100 trials, ~1500 plant steps, one iLQR iteration per step) on the
device, summarised like BASELINE.md section 2 next to the reference's published PaperData rows.
usage: mpc_vs_paperdata.py [out.json] [timeout] [trials] [method:H ...]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from trajoptkp_b200 import capi, mpc, tasks  # noqa: E402

out = sys.argv[1] if len(sys.argv) > 1 else None
timeout = int(sys.argv[2]) if len(sys.argv) > 2 else 1500
ntr = int(sys.argv[3]) if len(sys.argv) > 3 else 100
runs = sys.argv[4:] or ["SI_1:20", "SI_5:40", "SI_20:80", "AJ_2_20:40", "VC_2_20:40"]
METHODS = {"SI_1": (capi.KP_SET_INTERVAL, 1, 20), "SI_5": (capi.KP_SET_INTERVAL, 5, 20), "SI_20": (capi.KP_SET_INTERVAL, 20, 20),
           "AJ_2_20": (capi.KP_ADAPTIVE_JERK, 2, 20), "VC_2_20": (capi.KP_VELOCITY_CHANGE, 2, 20),
           "IE_2_20": (capi.KP_ITERATIVE_ERROR, 2, 20)}
PUB = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                                  "paperdata_walker_mpc.json")))
rows = []
for spec in runs:
    tag, H = spec.split(":")
    H = int(H)
    kp, mn, mx = METHODS[tag]
    tc = tasks.TaskConfig("walker_run", kp_method=kp, min_N=mn, max_N=mx)
    t0 = time.time()
    res = mpc.run_sync_mpc(tc, H, list(range(ntr)), timeout, progress=500, worker_lag=os.environ.get("MPC_LAG", "0") == "1",
                           fixed_index=int(os.environ.get("MPC_BEST", "1")) if os.environ.get("MPC_BEST", "1") != "search" else None)
    s = res["summary"]
    row = {"method": tag, "H": H, "trials": ntr, "timeout": timeout, "final_cost_mean": float(s[:, 0].mean()),
           "final_cost_median": float(np.median(s[:, 0])), "pct_derivs": float(s[:, 4].mean()),
           "final_cost_p10_p90": [float(np.percentile(s[:, 0], 10)), float(np.percentile(s[:, 0], 90))],
           "wall_s": time.time() - t0, "final_costs": [float(v) for v in s[:, 0]], "pct_per_trial": [float(v) for v in s[:, 4]]}
    pub = PUB.get(spec)
    if pub:
        c = np.array(pub["final_cost"])
        row["published"] = {"final_cost_mean": float(c.mean()), "final_cost_median": float(np.median(c)),
                            "final_cost_p10_p90": [float(np.percentile(c, 10)), float(np.percentile(c, 90))],
                            "pct_derivs": float(np.mean(pub["pct_derivs"]))}
    rows.append(row)
    print(json.dumps({k: v for k, v in row.items() if k not in ("final_costs", "pct_per_trial")}), flush=True)
    if out:
        json.dump(rows, open(out, "w"))
