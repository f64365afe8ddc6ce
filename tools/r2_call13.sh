#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_roi_heads_variants.py -q -m gpu > gpurun_out/r2m_heads.log 2>&1; echo "rc=$?"; tail -40 gpurun_out/r2m_heads.log
