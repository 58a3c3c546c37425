#!/bin/bash
# ncu evidence for profiles/ (B200_PROFILING.md recipe) of the default bench command (walker_run, default batch, H = 80):
# the launch list, then --set full captures of (a) the five world kernels of ONE FD launch (k_fd_worlds position stage +
# k_fd_dep_lanes nominal light / heavy + the rest light / heavy; their DRAM traffic is roofline.traffic), (b) the dominant
# lanes launch with source, (c) the nominal and the line-search rollout kernels with source.  The reports stay on the box
# (25-40 MB each); what comes back are the raw page of (a) and the summaries.  Run under gpurun; then
#   python tools/make_fd_traffic.py walker_run <batch> 80 velocity_change gpurun_out/<tag>_fd_launch_raw.csv
tag=${1:-r2}
mkdir -p gpurun_out
B="python bench.py --warmup 3 --no-cpu-baseline --no-targets"
$B --steps 2 > gpurun_out/${tag}_bench_plain.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
  $B --steps 2 > gpurun_out/${tag}_ncu_launches.log 2>&1
summarise() {  # name
  python tools/ncu_summary.py /tmp/${tag}_$1.ncu-rep 40 > gpurun_out/${tag}_$1_ncu_summary.txt 2>&1
  python tools/ncu_by_function.py /tmp/${tag}_$1.ncu-rep > gpurun_out/${tag}_$1_by_function.txt 2>&1
}
# five world kernels per FD launch; the first launch belongs to warm-up step 1
ncu --set full --clock-control none -k 'regex:k_fd_worlds|k_fd_dep_lanes' -s 5 -c 5 -f -o /tmp/${tag}_fd_launch \
  $B --steps 1 > gpurun_out/${tag}_ncu_fd_launch.log 2>&1
ncu -i /tmp/${tag}_fd_launch.ncu-rep --page raw --csv > gpurun_out/${tag}_fd_launch_raw.csv 2>/dev/null
python tools/ncu_summary.py /tmp/${tag}_fd_launch.ncu-rep 0 > gpurun_out/${tag}_fd_launch_ncu_summary.txt 2>&1
# the largest of them (all worlds but the nominal ones, <= 24 rows) with source
ncu --set full --clock-control none --import-source on -k regex:k_fd_dep_lanes -s 6 -c 1 -f -o /tmp/${tag}_fd_dep_lanes \
  $B --steps 1 > gpurun_out/${tag}_ncu_fd_dep_lanes.log 2>&1
summarise fd_dep_lanes
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:k_rollout<\(int\)32' -s 4 -c 1 -f \
  -o /tmp/${tag}_rollout_nom $B --steps 1 > gpurun_out/${tag}_ncu_rollout_nom.log 2>&1
summarise rollout_nom
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:k_rollout<\(int\)16' -s 2 -c 1 -f \
  -o /tmp/${tag}_rollout_fwd $B --steps 1 > gpurun_out/${tag}_ncu_rollout_fwd.log 2>&1
summarise rollout_fwd
ls -la gpurun_out | grep ${tag}_
