#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_step_gpu.py tests/test_gemm_gpu.py tests/test_stage2_gpu.py -q > gpurun_out/t33_tests.log 2>&1
echo "exit $?" >> gpurun_out/t33_tests.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/t33_bench.json 2> gpurun_out/t33_bench.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ce_ -c 6 --csv --log-file gpurun_out/t33_ce_ncu.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --graph 0 > gpurun_out/t33_ncu.log 2>&1
