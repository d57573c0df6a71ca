#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_collectives_gpu.py -m gpu -q --timeout 600 -x -s 2>&1 | tail -40 | tee gpurun_out/pytest_collectives.log
