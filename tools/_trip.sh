set -x
timeout 300 python -m pytest tests/test_attn_gpu.py -x -q -k gemma 2>&1 | tail -25 > gpurun_out/t11_attn_tests.log
timeout 120 python tools/attn_gemma_probe.py > gpurun_out/t11_probe.log 2>&1
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/t11_tests.log
timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/t11_bench.json 2> gpurun_out/t11_bench.err
