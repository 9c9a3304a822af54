#!/bin/bash
python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; tail -2 gpurun_out/pytest.log
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline"
for w in "--workload c5_counts --cpus 64 --events 8 --intervals 400000" "--workload c5_counts --cpus 512 --events 1 --intervals 400000" "--workload c5_counts --cpus 256 --events 32 --intervals 50000" "--workload c5_counts --cpus 64 --events 2 --intervals 1000000" "--workload c5_counts --cpus 256 --events 8 --intervals 100000"; do
  echo "== $w"; $B $w 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('value %.3e %s ms/step %.3f kernel_ms %.3f achieved %.0f GB/s frac %.3f' % (d['value'], d['unit'], d['ms_per_step'], r['kernel_ms'], r['achieved'], r['frac']), d['config'].get('slow_tiles'), d['config'].get('reruns'))"
done
