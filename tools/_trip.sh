#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_attn_gpu.py tests/test_step_gpu.py tests/test_hf_bf16_gpu.py -q > gpurun_out/t29_tests.log 2>&1
echo "exit $?" >> gpurun_out/t29_tests.log
timeout 600 python tools/bench_qwen3.py > gpurun_out/t29_qwen3.json 2> gpurun_out/t29_qwen3.err
timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/t29_bench.json 2> gpurun_out/t29_bench.err
