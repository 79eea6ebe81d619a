set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/t18_tests.log
timeout 400 python tools/bench_aux.py > gpurun_out/t18_aux.json 2> gpurun_out/t18_aux.err
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/t18_bench.json 2> gpurun_out/t18_bench.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/t18_launches.csv python bench.py --profile --steps 1 > gpurun_out/t18_ncu.log 2>&1
