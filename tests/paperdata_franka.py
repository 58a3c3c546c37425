"""Shared checks of the Franka open-loop sweeps against the per-trial rows the MuJoCo-based reference published
(tests/golden/paperdata_<task>_openloop.json, made by tests/golden/make_paperdata_openloop_fixture.py from
PaperData/iLQR_C++_Data/iLQR/<task>_openloop_*; GenTestingData::GenDataOpenloopOptimisation, 100 TestTasks scenes,
Optimise(max 10, min 3), lambda 0.01, initial controls of the task).

The MuJoCo version behind the files is unknown and contact-rich optimisations branch, so the pin is statistical like
the acrobot one (tests/paperdata.py): means over the 100 trials of iteration count, % derivatives and cost reduction,
and the median per-trial deviation of the cost reduction.  The margins are per task (MARGINS) and sit at 2-4x the
deviations measured in round 2 (DESIGN.md section 2 tabulates the measured values)."""
import json
import os

import numpy as np

from trajoptkp_b200 import capi

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
KP = {"SI": capi.KP_SET_INTERVAL, "AJ": capi.KP_ADAPTIVE_JERK, "VC": capi.KP_VELOCITY_CHANGE, "IE": capi.KP_ITERATIVE_ERROR}
# task -> (|d mean iterations|, |d mean % derivs|, |d mean cost reduction|, median per-trial |d cost reduction|)
# measured in round 2 (profiles/r2a_*_openloop_vs_paperdata.json), worst over the six published methods of a task:
#   reaching  0.42 / 0.09 / 0.005 / 0.0015      box_sweep 0.26 / 3.3 / 0.110 / 0.104      push_ncl 0.17 / 0.8 / 0.055 / 0.068
MARGINS = {"reaching": (0.6, 0.3, 0.01, 0.005), "box_sweep": (0.6, 5.0, 0.15, 0.15), "push_ncl": (0.5, 2.0, 0.08, 0.1),
           "push_lcl": (0.6, 5.0, 0.12, 0.15), "push_mcl": (0.6, 5.0, 0.15, 0.2)}


def published(task):
    return json.load(open(os.path.join(_DIR, "paperdata_%s_openloop.json" % task)))


def configure(tc, tag, entry):
    """Key-point method of one published run onto a TaskConfig: method, min_N / max_N and the thresholds summary.yaml
    recorded for that run (they differ from the task YAML's defaults)."""
    tc.kp_method = KP[tag.split("_")[0]]
    tc.min_N, tc.max_N = entry["min_N"], entry["max_N"]
    if "jerk_thresholds" in entry:
        tc.jerk_thr = np.array(entry["jerk_thresholds"], dtype=np.float64)
    if "velocity_change_thresholds" in entry:
        tc.vel_thr = np.array(entry["velocity_change_thresholds"], dtype=np.float64)
    if "error_threshold" in entry:
        tc.iter_thr = float(entry["error_threshold"])
    return tc


def compare(task, tag, cost_reduction, iterations, pct_derivs, trials=None):
    pub = published(task)[tag]
    sel = slice(None) if trials is None else list(trials)
    cr, it, pc = np.asarray(cost_reduction), np.asarray(iterations), np.asarray(pct_derivs)
    pcr, pit, ppc = np.array(pub["cost_reduction"])[sel], np.array(pub["iterations"])[sel], np.array(pub["pct_derivs"])[sel]
    return {"task": task, "method": tag, "trials": len(cr),
            "iters": (float(it.mean()), float(pit.mean())), "pct": (float(pc.mean()), float(ppc.mean())),
            "cost_reduction": (float(cr.mean()), float(pcr.mean())),
            "median_abs_dcr": float(np.median(np.abs(cr - pcr))), "iters_equal": int((it == pit).sum())}


def check(task, tag, cost_reduction, iterations, pct_derivs, trials=None):
    r = compare(task, tag, cost_reduction, iterations, pct_derivs, trials)
    mi, mp, mc, mm = MARGINS[task]
    assert abs(r["iters"][0] - r["iters"][1]) < mi, r
    assert abs(r["pct"][0] - r["pct"][1]) < mp, r
    assert abs(r["cost_reduction"][0] - r["cost_reduction"][1]) < mc, r
    assert r["median_abs_dcr"] < mm, r
    return r
