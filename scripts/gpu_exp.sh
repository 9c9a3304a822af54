#!/bin/bash
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; tail -3 gpurun_out/pytest.log
for w in c4_switch c4_mixed; do
python bench.py --steps 5 --warmup 3 --workload $w > gpurun_out/last.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/last.log').read().strip().splitlines()[-1]); r=d['roofline']; print('$w', d['ms_per_step'], r.get('decode_kernel_ms'), r['kernel_ms'], r['frac'], d['config']['slow_tiles'])"
done
