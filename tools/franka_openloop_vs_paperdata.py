#!/usr/bin/env python3
"""Open-loop sweeps of the Franka configs on the device against the reference's published per-trial rows
(tests/golden/paperdata_<task>_openloop.json).  One tokp_traj batch of 100 trials per (task, method), the reference's
own initial controls (sweep.prepare_trials).  Writes one JSON line per run and, with an output path, the full table.
usage: franka_openloop_vs_paperdata.py task[,task...] [method,method|all] [out.json] [horizon]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

import paperdata_franka as pf  # noqa: E402
from trajoptkp_b200 import sweep, tasks  # noqa: E402

names = sys.argv[1].split(",")
methods = sys.argv[2] if len(sys.argv) > 2 else "all"
out = sys.argv[3] if len(sys.argv) > 3 and sys.argv[3] != "-" else None
rows = []
for task in names:
    pub = pf.published(task)
    for tag, entry in pub.items():
        if methods != "all" and tag not in methods.split(","):
            continue
        T = int(sys.argv[4]) if len(sys.argv) > 4 else entry["horizon"]
        tc = pf.configure(tasks.TaskConfig(task), tag, entry)
        t0 = time.perf_counter()
        recs = sweep.run_openloop_sweep(tc, T, num_trials=100)
        r = pf.compare(task, tag, recs[:, 0], recs[:, 2].astype(int), recs[:, 4])
        r["T"] = T
        r["sweep_seconds"] = time.perf_counter() - t0
        r["published_source"] = entry["source"]
        r["per_trial"] = {"cost_reduction": [float(x) for x in recs[:, 0]], "iterations": [int(x) for x in recs[:, 2]],
                          "pct_derivs": [float(x) for x in recs[:, 4]]}
        rows.append(r)
        print(json.dumps({k: v for k, v in r.items() if k != "per_trial"}), flush=True)
        if out:
            json.dump(rows, open(out, "w"), indent=1)
