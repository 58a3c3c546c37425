# block-synchronous level of the spread (one world per warp) rollouts at the default bench batch (walker)
for bs in 0 1 2 3; do
  echo "SPREAD_BLOCKSYNC=$bs"; TOKP_SPREAD_BLOCKSYNC=$bs python bench.py --no-cpu-baseline --no-targets --steps 5 2>/dev/null | python tools/show_bench.py
done
