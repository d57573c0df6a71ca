#!/bin/bash
# usage: gpu_multi.sh N [quick]   (run under gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
for cfg in resnet50_ddp_amp resnet50_oss_sddp vitb16_fsdp_bf16; do
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --steps 15 --warmup 6 --config $cfg --no-e2e 2>&1 | grep -E '^\{|Error|error|b200\]' | tail -3 | tee gpurun_out/bench_${cfg}_ours_${N}gpu.log
done
timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 benchmarks/gather_gemm.py 2>&1 | grep -E "^\{|rror" | tail -5
if [ "$2" != "quick" ]; then
timeout 200 python -m pytest tests/test_collectives_gpu.py -m gpu -q --timeout 180 -x -k "real" -s 2>&1 | grep -E "^\[|passed|failed|Error" | tee gpurun_out/real_${N}gpu.log
fi
