#!/bin/bash
# ncu evidence for profiles/ (B200_PROFILING.md recipe): launch list of the default bench command, then one --set full
# capture of each FD world kernel of the timed step (walker_run, the default batch, H = 80).  Run under gpurun; read back locally
# with tools/ncu_summary.py and tools/make_fd_traffic.py.
tag=${1:-r2}
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-targets > gpurun_out/${tag}_bench_plain.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-targets > gpurun_out/${tag}_ncu_launches.log 2>&1
# launches 0-2 of k_fd_worlds belong to warm-up step 1 (phase 0 + its empty redo pass); skip to a later step
ncu --set full --clock-control none --import-source on -k regex:k_fd_worlds -s 6 -c 1 -f -o gpurun_out/${tag}_fd_worlds \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-targets > gpurun_out/${tag}_ncu_fd_worlds.log 2>&1
ncu --set full --clock-control none -k regex:k_fd_dep_lanes -s 6 -c 1 -f -o gpurun_out/${tag}_fd_dep_lanes \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-targets > gpurun_out/${tag}_ncu_fd_dep_lanes.log 2>&1
ls -la gpurun_out | tail -8
