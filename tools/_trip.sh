set -x
timeout 300 python -m pytest tests/test_attn_gpu.py -x -q 2>&1 | tail -5 > gpurun_out/t8_attn_tests.log
timeout 120 python tools/attn_probe.py > gpurun_out/t8_attn_probe.log 2>&1
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/t8_tests.log
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/t8_bench.json 2> gpurun_out/t8_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/t8_launches.csv python bench.py --profile --steps 1 > gpurun_out/t8_ncu.log 2>&1
