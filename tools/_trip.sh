#!/bin/bash
# trip 44: after dropping the backward row kernel — attention / step tests, bench, launch list
timeout 900 python -m pytest tests/test_attn_gpu.py tests/test_step_gpu.py -m gpu -x -q > gpurun_out/t44_tests.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/t44_bench.json 2> gpurun_out/t44_bench.err
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/t44_launches.csv python bench.py --profile --steps 1 > gpurun_out/t44_ncu.log 2>&1
tail -n 3 gpurun_out/t44_tests.log
