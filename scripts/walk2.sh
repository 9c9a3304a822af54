timeout 900 python -m pytest tests/test_switch_events.py -x -q -m gpu 2>&1 | tail -5
B="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-diagnostics"
timeout 300 $B --workload c4_switch > gpurun_out/w2_switch.json 2> gpurun_out/w2_switch.err
python - <<'PY'
import json
j=json.loads(open('gpurun_out/w2_switch.json').read().strip().splitlines()[-1])
print('ms/step',j['ms_per_step'],'frac',j['roofline']['frac'],'decode',j['roofline'].get('decode_kernel_ms'),'merge_ms',j['config'].get('merge_ms'))
PY
tail -3 gpurun_out/w2_switch.err
