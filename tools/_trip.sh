#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 120 python tools/image_one.py && timeout 400 ncu --set full --clock-control none --import-source on -k regex:resample_horizontal -s 2 -c 1 -o gpurun_out/t30_image_h python tools/image_one.py > gpurun_out/t30_ncu.log 2>&1
ncu -i gpurun_out/t30_image_h.ncu-rep --page raw --csv > gpurun_out/t30_image_h_raw.csv 2>/dev/null
ls -la gpurun_out/t30* >> gpurun_out/t30_ncu.log
rm -f gpurun_out/t30_image_h.ncu-rep
