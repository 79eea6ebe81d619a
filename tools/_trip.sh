#!/bin/bash
# trip 41: 2-GPU soaks (Stage 2 ragged, Stage 1 multi-epoch with evaluation)
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 \
  -m projectiontrainer_b200.train_vqa_stage2 --synthetic 648 --batch_size 8 --gradient_accumulation_steps 4 --num_epochs 2 \
  --unfreeze_projection_layer --disable_wandb --output_dir /tmp/vqa_stress2 > gpurun_out/t41_stage2_2gpu.log 2>&1
echo "exit $?" >> gpurun_out/t41_stage2_2gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 \
  -m projectiontrainer_b200.train_projection_stage1 --synthetic 400 --batch_size 16 --gradient_accumulation_steps 2 --num_epochs 3 \
  --disable_wandb --output_dir /tmp/s1_stress2 > gpurun_out/t41_stage1_2gpu.log 2>&1
echo "exit $?" >> gpurun_out/t41_stage1_2gpu.log
grep -c . gpurun_out/t41_stage2_2gpu.log
