#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_hf_bf16_gpu.py tests/test_stage2_gpu.py -q > gpurun_out/t28_tests.log 2>&1
echo "exit $?" >> gpurun_out/t28_tests.log
