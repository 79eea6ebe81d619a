#!/bin/bash
# trip 52: NF4 tests against the flush-to-zero oracle
timeout 300 python -m pytest tests/test_nf4_gpu.py -m gpu -x -q > gpurun_out/t52_tests.log 2>&1; tail -n 3 gpurun_out/t52_tests.log
