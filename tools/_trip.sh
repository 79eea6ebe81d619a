#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_infer_gpu.py tests/test_stage2_gpu.py tests/test_step_gpu.py -q > gpurun_out/t31_tests.log 2>&1
echo "exit $?" >> gpurun_out/t31_tests.log
timeout 600 python tools/bench_aux.py > gpurun_out/t31_aux.json 2> gpurun_out/t31_aux.err
