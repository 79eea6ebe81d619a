set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/t1_tests.log
python bench.py --steps 8 --warmup 3 --graph 0 --no-cpu-baseline > gpurun_out/t1_bench_eager.json 2> gpurun_out/t1_bench_eager.err
python bench.py --steps 8 --warmup 3 --graph 1 > gpurun_out/t1_bench_graph.json 2> gpurun_out/t1_bench_graph.err
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/t1_launches.csv python bench.py --profile --steps 1 > gpurun_out/t1_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attn_siglip -s 30 -c 1 -o gpurun_out/t1_attn_siglip -f python bench.py --profile --steps 1 > gpurun_out/t1_ncu_attn.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm2_bf16_kernel -s 700 -c 6 -o gpurun_out/t1_gemm2 -f python bench.py --profile --steps 1 > gpurun_out/t1_ncu_gemm.log 2>&1
ls -la gpurun_out
