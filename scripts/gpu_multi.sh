#!/bin/bash
# usage: gpu_multi.sh N [max_bytes]  (run under gpurun --gpus N)
N=${1:-2}
MAXB=${2:-1073741824}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_collectives_gpu.py -m gpu -q --timeout 280 -x -k "real_peer" -s 2>&1 | grep -E "^\[|passed|failed" | tee gpurun_out/real_${N}gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 benchmarks/collectives_sweep.py --max-bytes $MAXB --out gpurun_out/collectives_sweep_${N}gpu.jsonl 2>&1 | grep -v "^{" | tail -14 | tee gpurun_out/collectives_sweep_${N}gpu.md
for cfg in resnet50_ddp_amp resnet50_oss_sddp vitb16_fsdp_bf16; do
for impl in ours torch; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --steps 10 --warmup 6 --impl $impl --config $cfg --no-e2e 2>&1 | grep -E '^\{|Error|error' | tail -2 | tee gpurun_out/bench_${cfg}_${impl}_${N}gpu.log
done
done
