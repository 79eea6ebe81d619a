#!/bin/bash
# trip 63: split-K LM-head input gradient: tests + bench A/B
timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_step_gpu.py tests/test_stage2_gpu.py -m gpu -x -q 2>&1 | tail -n 3
PT_STAGE1_LM_HEAD_SPLITK=1 timeout 240 python bench.py --no-cpu-baseline > gpurun_out/s29_bench_sk1.json 2> gpurun_out/s29_bench_sk1.err
timeout 240 python bench.py --no-cpu-baseline > gpurun_out/s29_bench_sk8.json 2> gpurun_out/s29_bench_sk8.err
PT_STAGE1_LM_HEAD_SPLITK=1 timeout 240 python bench.py --no-cpu-baseline > gpurun_out/s29_bench_sk1b.json 2>> gpurun_out/s29_bench_sk1.err
timeout 240 python bench.py --no-cpu-baseline > gpurun_out/s29_bench_sk8b.json 2>> gpurun_out/s29_bench_sk8.err
for f in sk1 sk8 sk1b sk8b; do echo $f; grep -o '"ms_per_step": [0-9.]*' gpurun_out/s29_bench_$f.json | head -n 1; grep -o '"clocks": {[^}]*}' gpurun_out/s29_bench_$f.json; grep -o '"last_loss": [0-9.]*' gpurun_out/s29_bench_$f.json; done
tail -n 3 gpurun_out/s29_bench_sk8.err
