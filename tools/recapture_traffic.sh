# the five world kernels of one FD launch of the default bench command, --set full; only the raw page leaves the box
# (then: python tools/make_fd_traffic.py walker_run <batch> 80 velocity_change gpurun_out/<tag>_fd_launch_raw.csv)
tag=${1:-r2}
B="python bench.py --warmup 3 --no-cpu-baseline --no-targets"
ncu --set full --clock-control none -k 'regex:k_fd_worlds|k_fd_dep_lanes' -s 5 -c 5 -f -o /tmp/${tag}_fd_launch $B --steps 1 > gpurun_out/${tag}_ncu_fd_launch.log 2>&1
ncu -i /tmp/${tag}_fd_launch.ncu-rep --page raw --csv > gpurun_out/${tag}_fd_launch_raw.csv 2>/dev/null
python tools/ncu_summary.py /tmp/${tag}_fd_launch.ncu-rep 0 > gpurun_out/${tag}_fd_launch_ncu_summary.txt 2>&1
