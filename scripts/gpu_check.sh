#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q --timeout 300 -s 2>&1 | tail -30 | tee gpurun_out/pytest_kernels.log
timeout 300 python scripts/profile_step.py --impl ours --out gpurun_out/prof_ours.txt > gpurun_out/prof_ours.log 2>&1; tail -3 gpurun_out/prof_ours.log
timeout 300 python scripts/profile_step.py --impl torch --out gpurun_out/prof_torch.txt > gpurun_out/prof_torch.log 2>&1; tail -3 gpurun_out/prof_torch.log
timeout 600 python bench.py --steps 20 --warmup 8 2>&1 | tail -1 | tee gpurun_out/bench_ours.log
