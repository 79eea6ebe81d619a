#!/bin/bash
# trip 60: final tree of the session: GPU tests, smoke, bench line (with CPU baseline), reference arm, ncu launch list of one step
(time timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 4) > gpurun_out/s29f_tests.log 2>&1
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" > gpurun_out/s29f_smoke.log 2>&1
timeout 300 python bench.py > gpurun_out/s29f_bench.json 2> gpurun_out/s29f_bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/s29f_bench_ref.json 2> gpurun_out/s29f_bench_ref.err
timeout 200 python bench.py --profile --steps 1 > gpurun_out/s29f_profile_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s29f_launches.csv \
   python bench.py --profile --steps 1 > gpurun_out/s29f_profile_ncu.log 2>&1
tail -n 4 gpurun_out/s29f_tests.log; tail -n 2 gpurun_out/s29f_smoke.log; cut -c1-230 gpurun_out/s29f_bench.json; cut -c1-200 gpurun_out/s29f_bench_ref.json
cat gpurun_out/s29f_profile_plain.log | tail -n 1; wc -l gpurun_out/s29f_launches.csv
