#!/usr/bin/env python3
"""Extracts the reference's published per-trial results of the walker_run synchronous-MPC sweep
(PaperData/iLQR_C++_Data/iLQR/walker/walker_run_synchronus_mpc_*/summary.{csv,yaml}; GenDataMPCHorizons: 100 trials,
1500 plant steps, one iLQR iteration per step, control noise 0.5 %) into a small fixture: final cost and average
% derivatives per trial for every (key-point method, horizon).  Outputs of the real MuJoCo-based reference on the
contact-rich config.  Run in the build container: python tests/golden/make_paperdata_walker_fixture.py"""
import json
import os

import pandas as pd
import yaml

ROOT = "/root/reference/PaperData/iLQR_C++_Data/iLQR/walker"
out = {}
for d in sorted(os.listdir(ROOT)):
    y = yaml.safe_load(open(os.path.join(ROOT, d, "summary.yaml")))
    df = pd.read_csv(os.path.join(ROOT, d, "summary.csv"))
    n_replans = sum(1 for _ in open(os.path.join(ROOT, d, "0.csv"))) - 1
    key = "%s:%d" % (y["keypoint_name"].lstrip("*"), int(y["optimisation horizon"]))
    out[key] = {"source": "PaperData/iLQR_C++_Data/iLQR/walker/%s" % d, "controls_noise": float(y["controls_noise"]),
                "timestep": float(y["model timestep"]), "min_N": int(y["keypoint_min_N"]), "max_N": int(y["keypoint_max_N"]),
                "replans_trial0": n_replans,
                "final_cost": [round(float(v), 3) for v in df["Final cost"]],
                "pct_derivs": [round(float(v), 3) for v in df["Average percent derivs"]]}
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "paperdata_walker_mpc.json")
json.dump(out, open(path, "w"))
print(path, len(out), sorted(out)[:6], os.path.getsize(path))
