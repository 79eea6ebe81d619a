#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 600 python tools/bench_hf_gpu.py > gpurun_out/t27_hf_gpu.json 2> gpurun_out/t27_hf_gpu.err
echo "exit $?" >> gpurun_out/t27_hf_gpu.err
if ! grep -q images_per_s gpurun_out/t27_hf_gpu.json; then
  timeout 600 python tools/bench_hf_gpu.py --batch 16 > gpurun_out/t27_hf_gpu_b16.json 2> gpurun_out/t27_hf_gpu_b16.err
fi
timeout 900 python -m pytest tests/test_attn_gpu.py tests/test_stage2_gpu.py -q > gpurun_out/t27_tests.log 2>&1
echo "exit $?" >> gpurun_out/t27_tests.log
