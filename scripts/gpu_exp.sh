#!/bin/bash
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; tail -3 gpurun_out/pytest.log
show() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1', 'ms/step %.3f kernel_ms %.3f (decode %.3f fold %.3f) frac %.3f slow %s reruns %s' % (d['ms_per_step'], r['kernel_ms'], r.get('decode_kernel_ms',0), r.get('fold_kernel_ms',0), r['frac'], d['config']['slow_tiles'], d['config']['reruns']))" || tail -5 gpurun_out/last.log; }
run() { name=$1; shift; "$@" > gpurun_out/last.log 2>&1; cat gpurun_out/last.log | show $name; }
B="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline"
run hist_a $B
run hist_b $B
