#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/profile_step.py --impl ours --config vitb16_fsdp_bf16 --out gpurun_out/prof_vit_ours.txt > gpurun_out/prof_vit_ours.log 2>&1; tail -2 gpurun_out/prof_vit_ours.log
B200DP_DISABLE=gemm timeout 300 python scripts/profile_step.py --impl ours --config vitb16_fsdp_bf16 --out gpurun_out/prof_vit_nogemm.txt > gpurun_out/prof_vit_nogemm.log 2>&1; tail -2 gpurun_out/prof_vit_nogemm.log
