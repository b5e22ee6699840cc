#!/bin/bash
# scratch: the command list of the current gpurun call
mkdir -p gpurun_out
for nl in 1 0 1; do
AAR_NATIVE_LOOP=$nl timeout 120 python bench.py --steps 30 2>/dev/null | grep "^{" > gpurun_out/AW_bench1_nl$nl.json; python -c "
import json; d=json.load(open('gpurun_out/AW_bench1_nl$nl.json')); e=d['e2e']; print('native_loop=$nl', d['ms_per_step'], d['verified'], round(e['ms_per_step'],4), round(e['h2d_floor_ms_per_step'],4), round(e['vs_h2d_floor'],4), e.get('error'))"
done
timeout 100 python -m pytest tests/test_gpu_api.py -x -q -m gpu 2>&1 | tail -1
