set -x
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/t19_smoke.log 2>&1
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/t19_bench_a.json 2> gpurun_out/t19_bench_a.err
PT_STAGE1_BN_DKV=128 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/t19_bench_b.json 2> gpurun_out/t19_bench_b.err
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/t19_bench_c.json 2> gpurun_out/t19_bench_c.err
PT_STAGE1_BN_DKV=128 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/t19_bench_d.json 2> gpurun_out/t19_bench_d.err
