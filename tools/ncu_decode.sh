#!/bin/bash
# launch list of one decode step (run under gpurun; the program has to exit 0 without ncu first)
set -e
mkdir -p gpurun_out
timeout -k 5 300 python tools/decode_ncu.py ${1:-64} > gpurun_out/decode_plain.log 2>&1
timeout -k 5 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
  --log-file gpurun_out/decode_launches_b${1:-64}.csv python tools/decode_ncu.py ${1:-64} > gpurun_out/decode_ncu.log 2>&1
tail -2 gpurun_out/decode_ncu.log
