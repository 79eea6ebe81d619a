#!/bin/bash
# default trip: GPU tests, smoke, one bench line
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -n 2
timeout 300 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; cut -c1-240 gpurun_out/bench.json
