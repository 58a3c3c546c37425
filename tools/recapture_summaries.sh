tag=r2s
B="python bench.py --warmup 3 --no-cpu-baseline --no-targets"
summarise() { python tools/ncu_summary.py /tmp/${tag}_$1.ncu-rep 40 > gpurun_out/${tag}_$1_ncu_summary.txt 2>&1; python tools/ncu_by_function.py /tmp/${tag}_$1.ncu-rep > gpurun_out/${tag}_$1_by_function.txt 2>&1; }
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:k_rollout<\(int\)32' -s 4 -c 1 -f -o /tmp/${tag}_rollout_nom $B --steps 1 > /dev/null 2>&1
summarise rollout_nom
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:k_rollout<\(int\)16' -s 2 -c 1 -f -o /tmp/${tag}_rollout_fwd $B --steps 1 > /dev/null 2>&1
summarise rollout_fwd
ncu --set full --clock-control none --import-source on -k regex:k_fd_dep_lanes -s 6 -c 1 -f -o /tmp/${tag}_fd_dep_lanes $B --steps 1 > /dev/null 2>&1
summarise fd_dep_lanes
