#!/bin/bash
nvidia-smi topo -m 2>/dev/null | head -14
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/last.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/last.log').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e'])" || tail -5 gpurun_out/last.log
