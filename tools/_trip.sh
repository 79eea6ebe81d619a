set -x
timeout 200 python -m pytest tests/test_attn_gpu.py -x -q -k gemma 2>&1 | tail -25 > gpurun_out/t16_attn_tests.log
timeout 100 python tools/attn_gemma_probe.py > gpurun_out/t16_probe.log 2>&1
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/t16_tests.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/t16_bench.json 2> gpurun_out/t16_bench.err
