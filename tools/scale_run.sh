#!/bin/bash
# tools/scale_run.sh N: weak and strong bench lines on N GPUs of the box (run under `gpurun --gpus N`)
N=$1
cd "$(dirname "$0")/.."
P=$((29600 + RANDOM % 300))
for sc in weak strong; do
  P=$((P + 7))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N \
      --steps 30 --warmup 5 --scaling $sc --no-cpu-baseline > gpurun_out/r2_bench_n${N}_$sc.json 2> gpurun_out/r2_bench_n${N}_$sc.err
  python tools/bench_summary.py gpurun_out/r2_bench_n${N}_$sc.json 2>/dev/null | head -1
done
