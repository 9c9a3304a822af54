#!/bin/bash
# Tuning aid (not part of the product): parity tests, then bench.py under a few settings.
python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; tail -3 gpurun_out/pytest.log
show() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1', 'ms/step %.3f kernel_ms %.3f frac %.3f slow %s reruns %s' % (d['ms_per_step'], r['kernel_ms'], r['frac'], d['config']['slow_tiles'], d['config']['reruns']))" || tail -5 gpurun_out/last.log; }
run() { name=$1; shift; "$@" > gpurun_out/last.log 2>&1; cat gpurun_out/last.log | show $name; }
run none python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --agg none
for s in "$@"; do run hist_segs$s python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --segs-per-cpu $s; done
