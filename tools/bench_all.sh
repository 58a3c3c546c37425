#!/bin/bash
# bench lines of every BASELINE config (profiles/<tag>_bench_<task>.json)
tag=${1:-r2}
mkdir -p gpurun_out
for t in reaching box_sweep push_ncl push_lcl push_mcl acrobot; do
  timeout 600 python bench.py --task $t --no-targets --steps 3 > gpurun_out/${tag}_bench_$t.json 2> gpurun_out/${tag}_bench_$t.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${tag}_bench_$t.json").read().strip().splitlines()[-1])
    print("$t", "B", d["config"]["replans_per_gpu"], "value %.0f e2e %.0f ms %.1f fd %.2f x%d frac %.4f cpu %.0f lat %.1f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["fd_kernel"]["avg_ms"], d["fd_kernel"]["launches"], d["roofline"]["frac"], d["cpu_baseline"]["value"], d["replan_latency_ms_b1"]), d["phase_ms"])
except Exception as e:
    print("$t failed", e, open("gpurun_out/${tag}_bench_$t.err").read()[-400:])
PY
done
