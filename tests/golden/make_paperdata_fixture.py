#!/usr/bin/env python3
"""Extracts the reference's own published per-trial results for the acrobot open-loop sweep
(PaperData/iLQR_C++_Data/iLQR/acrobot_openloop_*/summary.{csv,yaml}, produced by the MuJoCo-based reference with
GenDataOpenloopOptimisation: 100 TestTasks scenes, T = 500, Optimise(max 10, min 3)) into a small fixture.  These are
outputs of the real reference, so they pin the whole path (model compiler, simulator step, cost, key-points, iLQR)
statistically - trial by trial where the optimisation does not branch, in the mean otherwise.
Run in the build container (reads /root/reference): python tests/golden/make_paperdata_fixture.py"""
import json
import os

import pandas as pd
import yaml

ROOT = "/root/reference/PaperData/iLQR_C++_Data/iLQR"
out = {}
for d in sorted(os.listdir(ROOT)):
    if not d.startswith("acrobot_openloop_"):
        continue
    y = yaml.safe_load(open(os.path.join(ROOT, d, "summary.yaml")))
    df = pd.read_csv(os.path.join(ROOT, d, "summary.csv"))
    out[y["keypoint_name"].lstrip("*")] = {
        "source": "PaperData/iLQR_C++_Data/iLQR/%s" % d, "horizon": int(y["optimisation horizon"]),
        "timestep": float(y["model timestep"]), "min_N": int(y["keypoint_min_N"]), "max_N": int(y["keypoint_max_N"]),
        "cost_reduction": [float(v) for v in df["Cost reduction"]],
        "iterations": [int(v) for v in df["Number iterations"]],
        "pct_derivs": [float(v) for v in df["Average percent derivs"]]}
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "paperdata_acrobot_openloop.json")
json.dump(out, open(path, "w"))
print(path, {k: len(v["iterations"]) for k, v in out.items()})
