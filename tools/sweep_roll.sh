# tuning knobs of the nominal rollout's spread layout at the default bench batch (walker)
for w in 1 2 3 4; do
  echo "SPREAD_WARPS=$w"; TOKP_SPREAD_WARPS=$w TOKP_BENCH_VERBOSE=1 python bench.py --no-cpu-baseline --no-targets --steps 3 2>/dev/null | python tools/show_bench.py
done
echo "SPREAD_KEEP_LANES"; TOKP_SPREAD_KEEP_LANES=1 python bench.py --no-cpu-baseline --no-targets --steps 3 2>/dev/null | python tools/show_bench.py
