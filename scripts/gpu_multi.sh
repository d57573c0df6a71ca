#!/bin/bash
# usage: gpu_multi.sh N   (run under gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 benchmarks/collectives_sweep.py --out gpurun_out/collectives_sweep_${N}gpu.jsonl 2>&1 | grep -v "^{" | tail -20 | tee gpurun_out/collectives_sweep_${N}gpu.md
for impl in ours torch; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --steps 20 --warmup 8 --impl $impl 2>&1 | grep -E '^\{|Error|error' | tail -2 | tee gpurun_out/bench_${impl}_${N}gpu.log
done
