set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/t2_tests.log
python tools/bench_aux.py > gpurun_out/t2_aux.json 2> gpurun_out/t2_aux.err
