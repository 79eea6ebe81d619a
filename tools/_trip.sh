#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_stage2_gpu.py -x -q > gpurun_out/t20_stage2_tests.log 2>&1
echo "exit $?" >> gpurun_out/t20_stage2_tests.log
timeout 1200 python -m pytest tests -q -m gpu --deselect tests/test_stage2_gpu.py > gpurun_out/t20_all_tests.log 2>&1
echo "exit $?" >> gpurun_out/t20_all_tests.log
timeout 300 python tools/bench_stage2.py > gpurun_out/t20_stage2_bench.json 2> gpurun_out/t20_stage2_bench.err
timeout 600 python -m projectiontrainer_b200.train_vqa_stage2 --synthetic 48 --batch_size 8 --gradient_accumulation_steps 2 --num_epochs 1 --unfreeze_projection_layer --disable_wandb --output_dir /tmp/vqa_out > gpurun_out/t20_stage2_cli.log 2>&1
echo "exit $?" >> gpurun_out/t20_stage2_cli.log
ls -R /tmp/vqa_out >> gpurun_out/t20_stage2_cli.log 2>&1
