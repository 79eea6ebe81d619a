set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/t13_tests.log
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/t13_bench.json 2> gpurun_out/t13_bench.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/t13_launches.csv python bench.py --profile --steps 1 > gpurun_out/t13_ncu.log 2>&1
timeout 200 ncu --set full --clock-control none --import-source on -k regex:attn_gemma_fwd -s 2 -c 1 -o gpurun_out/t13_attn_gemma -f python tools/attn_gemma_probe.py > gpurun_out/t13_ncu_attn.log 2>&1
