set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/t3_tests.log
PT_STAGE1_FUSE_GEGLU=0 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/t3_bench_unfused.json 2> gpurun_out/t3_bench_unfused.err
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/t3_bench_fused.json 2> gpurun_out/t3_bench_fused.err
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/t3_launches.csv python bench.py --profile --steps 1 > gpurun_out/t3_ncu.log 2>&1
