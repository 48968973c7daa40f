#!/bin/bash
# Round-2 final ncu captures (GPU box, under gpurun): launch list of the replayed C2 step; `--set full` of the six
# transposed-convolution kernels at C2 size and of the tcgen05 GEMM launches of one eager step.  Only text summaries are
# kept (the reports exceed the 64 MiB merge limit of gpurun_out/).
set -x
cd "$(dirname "$0")/.."
python bench.py --steps 2 --warmup 1 --no-eval --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches_graph.csv \
    python bench.py --steps 2 --warmup 1 --no-eval --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
REPS=""
for L in l3 l2; do
  python tools/ts_one.py $L > gpurun_out/plain_$L.log 2>&1 &&
  ncu --set full --clock-control none -k regex:'convt_ts_kernel|convt_ts_wgrad_kernel' -s 3 -c 3 -o gpurun_out/r2f_ts_$L -f python tools/ts_one.py $L > gpurun_out/ncu_$L.log 2>&1
  REPS="$REPS gpurun_out/r2f_ts_$L.ncu-rep"
done
python bench.py --steps 1 --warmup 1 --no-graph --no-eval --no-cpu-baseline > gpurun_out/plain_bench3.log 2>&1 &&
ncu --set full --clock-control none -k regex:gemm_tc_kernel -s 20 -c 20 -o gpurun_out/r2f_gemm_tc -f \
    python bench.py --steps 1 --warmup 1 --no-graph --no-eval --no-cpu-baseline > gpurun_out/ncu_gemmtc.log 2>&1
ls -la gpurun_out/*.ncu-rep
python tools/ncu_summary2.py $REPS gpurun_out/r2f_gemm_tc.ncu-rep --json gpurun_out/traffic_final.json > gpurun_out/r2f_summary.txt 2>&1
for f in r2f_ts_l3 r2f_ts_l2; do ncu -i gpurun_out/$f.ncu-rep --page details > gpurun_out/$f.details.txt 2>&1; done
rm -f gpurun_out/*.ncu-rep
du -sh gpurun_out
