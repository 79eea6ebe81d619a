#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29577 bench.py --gpus 8 --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/t26_bench_8gpu.json 2> gpurun_out/t26_bench_8gpu.err
echo "exit $?" >> gpurun_out/t26_bench_8gpu.err
tail -5 gpurun_out/t26_bench_8gpu.err
