#!/bin/bash
# scratch: the command list of the current gpurun call
mkdir -p gpurun_out
timeout 300 python bench.py 2> gpurun_out/AT_bench1.err | grep "^{" > gpurun_out/AT_bench1.json; python -c "
import json; d=json.load(open('gpurun_out/AT_bench1.json')); print(d['ms_per_step'], d['clocks'], d['verified'], d['e2e']['vs_h2d_floor'])"
tail -3 gpurun_out/AT_bench1.err
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/AT_gpu_suite.txt 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/AT_gpu_suite.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -1
