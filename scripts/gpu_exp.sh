#!/bin/bash
python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; tail -2 gpurun_out/pytest.log
for g in 4 8; do
HBT_BENCH_TRACE=1 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-groups $g > gpurun_out/last.log 2> gpurun_out/last.err
grep trace gpurun_out/last.err | tail -1 | cut -c1-200
python -c "
import json,sys
d=json.loads(open('gpurun_out/last.log').read().strip().splitlines()[-1]); print('groups $g', d['ms_per_step'], d['gpu_launches'], d['e2e']['ms_per_step'])" || tail -5 gpurun_out/last.err
done
