#!/bin/bash
# trip 59: attention-backward GEMMs on three streams: step tests + bench A/B
timeout 600 python -m pytest tests/test_step_gpu.py tests/test_stage2_gpu.py -m gpu -x -q 2>&1 | tail -n 3
PT_STAGE1_ATTN_BWD_STREAMS=0 timeout 240 python bench.py --no-cpu-baseline > gpurun_out/s29_bench_nostreams.json 2> gpurun_out/s29_bench_nostreams.err
timeout 240 python bench.py --no-cpu-baseline > gpurun_out/s29_bench_streams.json 2> gpurun_out/s29_bench_streams.err
PT_STAGE1_ATTN_BWD_STREAMS=0 timeout 240 python bench.py --no-cpu-baseline > gpurun_out/s29_bench_nostreams2.json 2>> gpurun_out/s29_bench_nostreams.err
timeout 240 python bench.py --no-cpu-baseline > gpurun_out/s29_bench_streams2.json 2>> gpurun_out/s29_bench_streams.err
for f in nostreams streams nostreams2 streams2; do echo $f; cut -c1-200 gpurun_out/s29_bench_$f.json; grep -o '"clocks": {[^}]*}' gpurun_out/s29_bench_$f.json; done
tail -n 3 gpurun_out/s29_bench_streams.err
