#!/bin/bash
# trip 48: Qwen3-8B-shaped step — B=8 and B=32 timings, ncu launch list at B=8
timeout 600 python tools/bench_qwen3.py --batch 8 > gpurun_out/t48_qwen3_b8.json 2> gpurun_out/t48_qwen3_b8.err
timeout 600 python tools/bench_qwen3.py --batch 32 > gpurun_out/t48_qwen3_b32.json 2> gpurun_out/t48_qwen3_b32.err
PT_STAGE1_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/t48_qwen3_launches.csv python tools/bench_qwen3.py --batch 8 --iters 1 > gpurun_out/t48_ncu.log 2>&1
cat gpurun_out/t48_qwen3_b8.json gpurun_out/t48_qwen3_b32.json; tail -n 3 gpurun_out/t48_qwen3_b32.err
