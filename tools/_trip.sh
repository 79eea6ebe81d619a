#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_stage2_gpu.py -q > gpurun_out/t21_stage2_tests.log 2>&1
echo "exit $?" >> gpurun_out/t21_stage2_tests.log
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
  -m projectiontrainer_b200.train_vqa_stage2 --synthetic 96 --batch_size 8 --gradient_accumulation_steps 2 --num_epochs 1 \
  --unfreeze_projection_layer --disable_wandb --output_dir /tmp/vqa_out2 > gpurun_out/t21_stage2_cli_2gpu.log 2>&1
echo "exit $?" >> gpurun_out/t21_stage2_cli_2gpu.log
ls -R /tmp/vqa_out2 >> gpurun_out/t21_stage2_cli_2gpu.log 2>&1
