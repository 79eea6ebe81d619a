#!/bin/bash
# trip 40: full suite after the capture-mode / fork-gc / qknorm row-kernel changes + fork-gc probe
timeout 120 python tools/_probe_fork_gc.py > gpurun_out/t40_probe.log 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/t40_tests.log 2>&1
tail -n 3 gpurun_out/t40_tests.log
tail -n 3 gpurun_out/t40_probe.log
