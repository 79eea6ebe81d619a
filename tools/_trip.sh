#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_image_gpu.py tests/test_stage2_gpu.py -q > gpurun_out/t22_tests.log 2>&1
echo "exit $?" >> gpurun_out/t22_tests.log
timeout 300 python tools/bench_image.py > gpurun_out/t22_image_bench.json 2> gpurun_out/t22_image_bench.err
