#!/bin/bash
# Round-2 ncu captures, second part: `--set full` of the transposed-convolution weight-gradient kernel at C2 size (both layers)
# and of every tcgen05 GEMM launch of one eager C2 step.  Outputs under gpurun_out/.
set -x
cd "$(dirname "$0")/.."
for L in l3 l2; do
  python tools/ts_one.py $L > gpurun_out/plain_$L.log 2>&1 &&
  ncu --set full --clock-control none -k regex:convt_ts_wgrad_kernel -s 1 -c 1 -o gpurun_out/r2_tsw_$L -f python tools/ts_one.py $L > gpurun_out/ncu_w$L.log 2>&1
done
python bench.py --steps 1 --warmup 1 --no-graph --no-eval --no-cpu-baseline > gpurun_out/plain_bench3.log 2>&1 &&
ncu --set full --clock-control none -k regex:gemm_tc_kernel -s 37 -c 37 -o gpurun_out/r2_gemm_tc -f \
    python bench.py --steps 1 --warmup 1 --no-graph --no-eval --no-cpu-baseline > gpurun_out/ncu_gemmtc.log 2>&1
ls -la gpurun_out/*.ncu-rep
python tools/ncu_summary2.py gpurun_out/r2_tsw_l3.ncu-rep gpurun_out/r2_tsw_l2.ncu-rep gpurun_out/r2_gemm_tc.ncu-rep --json gpurun_out/traffic_b.json > gpurun_out/r2_wgrad_gemm_summary.txt 2>&1
for f in r2_tsw_l3 r2_tsw_l2; do ncu -i gpurun_out/$f.ncu-rep --page details > gpurun_out/$f.details.txt 2>&1; done
ncu -i gpurun_out/r2_gemm_tc.ncu-rep --page details > gpurun_out/r2_gemm_tc.details.txt 2>&1
rm -f gpurun_out/*.ncu-rep            # the text summaries travel back; the reports exceed the 64 MiB merge limit
du -sh gpurun_out
