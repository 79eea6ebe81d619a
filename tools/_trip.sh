#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_image_gpu.py -q > gpurun_out/t24_tests.log 2>&1
echo "exit $?" >> gpurun_out/t24_tests.log
timeout 300 python tools/bench_image.py > gpurun_out/t24_image_bench.json 2> gpurun_out/t24_image_bench.err
timeout 120 python tools/image_one.py && timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:resample --csv --log-file gpurun_out/t24_image_ncu.csv python tools/image_one.py > gpurun_out/t24_ncu.log 2>&1
