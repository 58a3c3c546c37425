#!/usr/bin/env python3
"""Extracts the reference's own published per-trial results of the Franka open-loop sweeps
(PaperData/iLQR_C++_Data/iLQR/<task>_openloop_*/summary.{csv,yaml}; produced by the MuJoCo-based reference with
GenTestingData::GenDataOpenloopOptimisation, src/GenTestingData.cpp:136-259: 100 TestTasks scenes,
Optimise(max 10, min 3), lambda 0.01) into one small fixture per task: tests/golden/paperdata_<task>_openloop.json.
Each entry carries the key-point parameters of that run as summary.yaml recorded them (they differ from the task YAML).
Outputs of the real reference: they pin the whole path (model compiler, simulator step, initial-control generators, cost,
key-points, iLQR) statistically.
Run in the build container (reads /root/reference): python tests/golden/make_paperdata_openloop_fixture.py"""
import json
import os

import pandas as pd
import yaml

ROOT = "/root/reference/PaperData/iLQR_C++_Data/iLQR"
TASKS = ["reaching", "box_sweep", "push_ncl", "push_lcl", "push_mcl"]
here = os.path.dirname(os.path.abspath(__file__))
for task in TASKS:
    out = {}
    for d in sorted(os.listdir(ROOT)):
        if not d.startswith(task + "_openloop_"):
            continue
        y = yaml.safe_load(open(os.path.join(ROOT, d, "summary.yaml")))
        df = pd.read_csv(os.path.join(ROOT, d, "summary.csv"))
        e = {"source": "PaperData/iLQR_C++_Data/iLQR/%s" % d, "horizon": int(y["optimisation horizon"]),
             "timestep": float(y["model timestep"]), "min_N": int(y["keypoint_min_N"]), "max_N": int(y["keypoint_max_N"]),
             "cost_reduction": [float(v) for v in df["Cost reduction"]],
             "iterations": [int(v) for v in df["Number iterations"]],
             "pct_derivs": [float(v) for v in df["Average percent derivs"]]}
        for k in ("jerk_thresholds", "velocity_change_thresholds", "error_threshold"):
            if k in y:
                e[k] = y[k]
        out[y["keypoint_name"].lstrip("*")] = e
    path = os.path.join(here, "paperdata_%s_openloop.json" % task)
    json.dump(out, open(path, "w"))
    print(path, {k: (len(v["iterations"]), round(sum(v["cost_reduction"]) / 100, 3)) for k, v in out.items()})
