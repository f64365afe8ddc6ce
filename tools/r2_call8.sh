#!/bin/bash
# round 2, GPU call 8: ROIMerge / ROILabel / CMIL layer tests
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cmil.py -q -m gpu > gpurun_out/r2h_cmil.log 2>&1; echo "cmil rc=$?"; tail -30 gpurun_out/r2h_cmil.log
