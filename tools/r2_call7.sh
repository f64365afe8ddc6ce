#!/bin/bash
# round 2, GPU call 7: seed-count-on-device tests
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_heads.py -q -m gpu > gpurun_out/r2g_heads.log 2>&1; echo "heads rc=$?"; tail -5 gpurun_out/r2g_heads.log
