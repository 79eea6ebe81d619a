#!/bin/bash
# --set full capture of pt_decode_attn at B = $1, nsplit = $2 (run under gpurun; the probe has exited 0 without ncu before)
set -e
mkdir -p gpurun_out
PROBE_NSPLIT=$2 timeout -k 5 600 ncu --set full --import-source on --clock-control none -k regex:decode_attn -c 1 \
  -o gpurun_out/decode_attn_B$1_ns$2 -f python tools/decode_attn_probe.py $1 > gpurun_out/ncu_da_$1_$2.log 2>&1
ls -la gpurun_out/decode_attn_B$1_ns$2.ncu-rep
