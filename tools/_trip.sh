#!/bin/bash
# trip 42: NF4 frozen weights
timeout 600 python -m pytest tests/test_nf4_gpu.py -m gpu -x -q > gpurun_out/t42_nf4_tests.log 2>&1
timeout 600 python -m projectiontrainer_b200.train_projection_stage1 --synthetic 128 --batch_size 16 --num_epochs 1 \
  --enable_qlora --disable_wandb --output_dir /tmp/s1_nf4 > gpurun_out/t42_cli_nf4.log 2>&1
echo "exit $?" >> gpurun_out/t42_cli_nf4.log
timeout 600 python -m projectiontrainer_b200.train_projection_stage1 --synthetic 32 --batch_size 8 --num_epochs 1 \
  --synthetic_llm qwen3_8b --enable_qlora --disable_wandb --output_dir /tmp/s1_nf4q > gpurun_out/t42_cli_nf4_qwen.log 2>&1
echo "exit $?" >> gpurun_out/t42_cli_nf4_qwen.log
tail -n 5 gpurun_out/t42_nf4_tests.log
