#!/bin/bash
# trip 47: NF4 de-quantise with four loads in flight; arena byte budget test
timeout 600 python -m pytest tests/test_nf4_gpu.py tests/test_stage2_gpu.py -m gpu -x -q > gpurun_out/t47_tests.log 2>&1
timeout 600 python tools/bench_nf4.py > gpurun_out/t47_nf4_bench.json 2> gpurun_out/t47_nf4_bench.err
tail -n 3 gpurun_out/t47_tests.log; cat gpurun_out/t47_nf4_bench.json; tail -n 5 gpurun_out/t47_nf4_bench.err
