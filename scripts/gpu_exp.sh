#!/bin/bash
timeout 900 python -m pytest tests/test_switch_events.py -m gpu -q -x > gpurun_out/pytest_sw.log 2>&1; tail -3 gpurun_out/pytest_sw.log
python bench.py --steps 5 --warmup 3 --workload c4_switch > gpurun_out/last.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/last.log').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['config']['merge_ms'])" || tail -5 gpurun_out/last.log
