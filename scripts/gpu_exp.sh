#!/bin/bash
# Scratch driver for one gpurun call while iterating: GPU tests, then the bench lines that
# matter (rewritten per experiment; scripts/gpu_round_measure.sh is the round-end recipe).
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; tail -3 gpurun_out/pytest.log
for w in c2_ip_hist "c3_deltas --events 8" c4_mixed c4_switch; do
python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload $w > gpurun_out/last.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/last.log').read().strip().splitlines()[-1]); r=d['roofline']; print('$w', round(d['ms_per_step'],3), r.get('decode_kernel_ms'), round(r['kernel_ms'],3), round(r['frac'],4), d['config']['slow_tiles'])" || tail -3 gpurun_out/last.log
done
