#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/t32_all_tests.log 2>&1
echo "exit $?" >> gpurun_out/t32_all_tests.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/t32_smoke.log 2>&1
echo "exit $?" >> gpurun_out/t32_smoke.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus 4 --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/t32_bench_4gpu.json 2> gpurun_out/t32_bench_4gpu.err
echo "exit $?" >> gpurun_out/t32_bench_4gpu.err
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/t32_bench_full.json 2> gpurun_out/t32_bench_full.err
echo "exit $?" >> gpurun_out/t32_bench_full.err
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/t32_bench_ref.json 2> gpurun_out/t32_bench_ref.err
echo "exit $?" >> gpurun_out/t32_bench_ref.err
