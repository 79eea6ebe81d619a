# ncu --set full captures added late in round 2 (run under gpurun, one GPU): LM-head GEMM with the CE-statistics
# epilogue, and the parallel Huffman kernels of the JPEG decode.
python tools/lmhead_one.py > gpurun_out/lmhead_one.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm2 -s 12 -c 1 -o gpurun_out/r02_lmhead_ce_stats -f python tools/lmhead_one.py > gpurun_out/ncu_lmhead.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:jpeg_sync\|jpeg_write\|jpeg_stitch_par\|jpeg_dc -c 4 -o gpurun_out/r02_jpeg_huffman -f python tools/bench_jpeg.py 8 > gpurun_out/ncu_jpeg_full.log 2>&1
cat gpurun_out/lmhead_one.txt
ls -la gpurun_out/*.ncu-rep
