set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/t10_tests.log
timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/t10_bench.json 2> gpurun_out/t10_bench.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/t10_launches.csv python bench.py --profile --steps 1 > gpurun_out/t10_ncu.log 2>&1
timeout 120 python tools/geglu_one.py > gpurun_out/t10_geglu_one.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm2_bf16_kernel -s 5 -c 1 -o gpurun_out/t10_gemm2_geglu -f python tools/geglu_one.py > gpurun_out/t10_ncu_gemm.log 2>&1
timeout 300 python tools/gemm_breakdown.py --cublas 0 > gpurun_out/t10_gemm_breakdown.txt 2>&1
ls -la gpurun_out
