set -e
python tools/bench_jpeg.py 8 > gpurun_out/jpeg_b8.json 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_jpeg_launches.csv python tools/bench_jpeg.py 8 > gpurun_out/ncu_jpeg.log 2>&1 || true
tail -2 gpurun_out/jpeg_b8.json | cut -c1-600
