#!/bin/bash
# scratch: the command list of the current gpurun call (edit, then `gpurun [--gpus N] -- 'bash bench/gpu_call.sh'`).
# As left: the one-GPU regression -- GPU test suite, smoke(), bench.py.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/gpu_suite.txt 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/gpu_suite.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -1
timeout 300 python bench.py 2> gpurun_out/bench1.err | grep "^{" > gpurun_out/bench1.json; tail -c 600 gpurun_out/bench1.json
