#!/bin/bash
# trip 50: NF4 quantise with the table look-up code search — parity + bandwidth
timeout 600 python -m pytest tests/test_nf4_gpu.py -m gpu -x -q > gpurun_out/t50_tests.log 2>&1
timeout 600 python tools/bench_nf4.py > gpurun_out/t50_nf4_bench.json 2> gpurun_out/t50_nf4_bench.err
tail -n 3 gpurun_out/t50_tests.log; cat gpurun_out/t50_nf4_bench.json; tail -n 5 gpurun_out/t50_nf4_bench.err
