#!/bin/bash
# Round-2 ncu captures (run on the GPU box under gpurun): the launch list of the replayed C2 step, and `--set full` of the
# transposed-convolution kernels at C2 size and of the other top kernels of the step.  Outputs under gpurun_out/.
set -x
cd "$(dirname "$0")/.."
if [ "$1" != "step" ]; then
python bench.py --steps 2 --warmup 1 --no-eval --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches_graph.csv \
    python bench.py --steps 2 --warmup 1 --no-eval --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
for L in l3 l2; do
  python tools/ts_one.py $L > gpurun_out/plain_$L.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:convt_ts -s 4 -c 4 -o gpurun_out/r2_ts_$L -f python tools/ts_one.py $L > gpurun_out/ncu_$L.log 2>&1
done
fi
if [ "$1" = "step" ]; then
python bench.py --steps 1 --warmup 1 --no-graph --no-eval --no-cpu-baseline > gpurun_out/plain_bench2.log 2>&1 &&
ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section LaunchStats --section Occupancy --clock-control none \
    -k regex:'gemm_tc_kernel|loglik|convt_l4|convt_small_wgrad|gemm_skinny|moe_logw' -s 300 -c 40 -o gpurun_out/r2_step_top -f \
    python bench.py --steps 1 --warmup 1 --no-graph --no-eval --no-cpu-baseline > gpurun_out/ncu_step.log 2>&1
fi
ls -la gpurun_out/*.ncu-rep
