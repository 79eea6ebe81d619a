#!/bin/bash
# trip 66: half-width last n tile: full GPU suite + step A/B
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 2)
for r in a b; do
PT_GEMM_HALF_TILE=0 timeout 240 python bench.py --no-cpu-baseline > gpurun_out/s29_bench_ht0$r.json 2> gpurun_out/s29_bench_ht0.err
timeout 240 python bench.py --no-cpu-baseline > gpurun_out/s29_bench_ht1$r.json 2> gpurun_out/s29_bench_ht1.err
done
for f in ht0a ht1a ht0b ht1b; do echo $f; grep -o '"ms_per_step": [0-9.]*' gpurun_out/s29_bench_$f.json | head -n 1; grep -o '"clocks": {[^}]*}' gpurun_out/s29_bench_$f.json; grep -o '"last_loss": [0-9.]*' gpurun_out/s29_bench_$f.json; done
