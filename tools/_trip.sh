#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/t25_all_tests.log 2>&1
echo "exit $?" >> gpurun_out/t25_all_tests.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/t25_bench.json 2> gpurun_out/t25_bench.err
timeout 600 python tools/bench_qwen3.py > gpurun_out/t25_qwen3.json 2> gpurun_out/t25_qwen3.err
