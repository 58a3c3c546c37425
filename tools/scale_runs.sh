#!/bin/bash
# multi-GPU records for profiles/: bench.py (weak scaling, walker + targets.box_sweep) and the strong-scaled 100-trial
# open-loop sweeps at N GPUs.  usage (under gpurun --gpus N): scale_runs.sh N tag [bench]
N=$1; tag=$2; mkdir -p gpurun_out
run() { if [ "$N" = 1 ]; then python bench.py "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; fi; }
for t in push_ncl push_lcl; do
  run --sweep $t 2> gpurun_out/${tag}_sweep_${t}_${N}gpu.err | tail -1 > gpurun_out/${tag}_sweep_${t}_${N}gpu.json
  python -c "import json;d=json.load(open('gpurun_out/${tag}_sweep_${t}_${N}gpu.json'));print('$t N=$N sweep %.2f s  %.1f trials/s  cost red %.3f'%(d['sweep_seconds'],d['value'],d['mean_cost_reduction']))" || tail -3 gpurun_out/${tag}_sweep_${t}_${N}gpu.err
done
if [ "$3" = bench ]; then
  run --no-cpu-baseline 2> gpurun_out/${tag}_bench_${N}gpu.err | tail -1 > gpurun_out/${tag}_bench_${N}gpu.json
  python -c "import json;d=json.load(open('gpurun_out/${tag}_bench_${N}gpu.json'));print('bench N=$N walker %.0f Jac/s %.2f ms | box_sweep %.0f Jac/s %.2f ms'%(d['value'],d['ms_per_step'],d['targets']['box_sweep']['value'],d['targets']['box_sweep']['ms_per_step']))" || tail -3 gpurun_out/${tag}_bench_${N}gpu.err
fi
