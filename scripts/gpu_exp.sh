#!/bin/bash
# Scratch driver for one gpurun call while iterating: GPU tests, then the bench lines that
# matter (rewritten per experiment; scripts/gpu_round_measure.sh is the round-end recipe).
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; tail -3 gpurun_out/pytest.log
show() { python -c "
import json,sys
d=json.loads(open('gpurun_out/last.log').read().strip().splitlines()[-1]); r=d['roofline']; c=d['config']
print('$1', 'ms/step', round(d['ms_per_step'],3), 'decode', r.get('decode_kernel_ms'), 'span', round(r['kernel_ms'],3), 'frac', round(r['frac'],4), 'slow', c.get('slow_tiles'), 'reruns', c.get('reruns'), c.get('rerun_segments'))" || tail -3 gpurun_out/last.log; }
for w in "c2_ip_hist" "c2_ip_hist --agg none" "c3_deltas --events 8" "c3_deltas --events 32" "c3_hostile --events 8" c4_mixed c4_switch; do
python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-diagnostics --workload $w > gpurun_out/last.log 2>&1; show "$w"
done
