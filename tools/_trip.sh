#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_step_gpu.py -q > gpurun_out/t36_tests.log 2>&1
echo "exit $?" >> gpurun_out/t36_tests.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/t36_bench.json 2> gpurun_out/t36_bench.err
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k 'regex:int.256, .int.5>' -c 30 --csv --log-file gpurun_out/t36_ds_list.csv python bench.py --profile --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/t36_ncu.log 2>&1
