#!/bin/bash
for w in c4_mixed c4_switch; do
timeout 600 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --workload $w --cpus 32 --ring-mib 1024 > gpurun_out/last.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/last.log').read().strip().splitlines()[-1]); r=d['roofline']; print('1GiB x32 $w', round(d['ms_per_step'],3), round(r['kernel_ms'],3), round(r['frac'],4), d['config'])" || tail -5 gpurun_out/last.log
done
nvidia-smi --query-gpu=memory.used --format=csv
