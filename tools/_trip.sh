#!/bin/bash
# trip 61: A/B of the GeGLU-forward epilogue with packed-f32 math (old = HEAD)
for i in 1 2; do
  PT_STAGE1_LIB=$PWD/projectiontrainer_b200/libpt_stage1_old.so timeout 120 python tools/geglu_one.py 2>&1 | head -n 1 | sed "s/^/old: /"
  timeout 120 python tools/geglu_one.py 2>&1 | head -n 2 | sed "s/^/new: /"
done
timeout 300 python -m pytest tests/test_gemm_gpu.py tests/test_step_gpu.py -m gpu -x -q 2>&1 | tail -n 2
