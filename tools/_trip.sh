#!/bin/bash
# trip 67: final tree: smoke, bench line (with CPU baseline), launch list of one step
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" > gpurun_out/s29g_smoke.log 2>&1
timeout 300 python bench.py > gpurun_out/s29g_bench.json 2> gpurun_out/s29g_bench.err
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/s29g_bench20.json 2> gpurun_out/s29g_bench20.err
timeout 200 python bench.py --profile --steps 1 > gpurun_out/s29g_profile_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s29g_launches.csv \
   python bench.py --profile --steps 1 > gpurun_out/s29g_profile_ncu.log 2>&1
tail -n 2 gpurun_out/s29g_smoke.log; cut -c1-230 gpurun_out/s29g_bench.json; cut -c1-230 gpurun_out/s29g_bench20.json; tail -n 1 gpurun_out/s29g_profile_plain.log
