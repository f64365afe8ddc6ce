#!/bin/bash
# round 2, GPU call 11: PCL targets tests, then the whole GPU suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pcl_targets.py -q -m gpu > gpurun_out/r2k_pcl.log 2>&1; echo "pcl rc=$?"; tail -15 gpurun_out/r2k_pcl.log
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2k_pytest.log
