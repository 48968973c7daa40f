#!/bin/bash
# Round-2 closing captures (GPU box, under gpurun): launch list of the replayed C2 step of the final build and
# `--set full` of the three last-layer kernels (forward, dgrad, convt_l4w_ts_kernel) at C2 size.  Text summaries only.
set -x
cd "$(dirname "$0")/.."
python bench.py --steps 2 --warmup 1 --no-eval --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches_graph_final.csv \
    python bench.py --steps 2 --warmup 1 --no-eval --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python tools/l4_prof.py 2 > gpurun_out/plain_l4.log 2>&1 &&
ncu --set full --clock-control none -k regex:'convt_l4' -s 3 -c 3 -o gpurun_out/r2f_l4 -f python tools/l4_prof.py 2 > gpurun_out/ncu_l4.log 2>&1
python tools/ncu_summary2.py gpurun_out/r2f_l4.ncu-rep --json gpurun_out/traffic_l4.json > gpurun_out/r2f_l4_summary.txt 2>&1
ncu -i gpurun_out/r2f_l4.ncu-rep --page details > gpurun_out/r2f_l4.details.txt 2>&1
rm -f gpurun_out/*.ncu-rep
du -sh gpurun_out
