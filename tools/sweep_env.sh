for bs in 0 1 2 3; do
  echo "BLOCKSYNC=$bs"; TOKP_FD_BLOCKSYNC=$bs python bench.py --no-cpu-baseline --steps 3 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['fd_kernel']['avg_ms'], d['phase_ms'])"
done
for w in 3 4 8 12; do
  echo "WARPS=$w"; TOKP_WARPS=$w python bench.py --no-cpu-baseline --steps 3 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['fd_kernel']['avg_ms'], d['phase_ms'])"
done
for l in 8 32; do
  echo "LANES=$l"; TOKP_LANES=$l python bench.py --no-cpu-baseline --steps 3 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['fd_kernel']['avg_ms'], d['phase_ms'])"
done
