set -x
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_switch_events.py -x -q -m gpu 2>&1 | tail -15
B="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-diagnostics"
for k in tiles walk; do
  timeout 300 $B --workload c4_switch --kernel $k > gpurun_out/w_switch_$k.json 2> gpurun_out/w_switch_$k.err
  timeout 300 $B --workload c4_mixed --kernel $k > gpurun_out/w_mixed_$k.json 2> gpurun_out/w_mixed_$k.err
  timeout 300 $B --workload c3_deltas --kernel $k > gpurun_out/w_deltas_$k.json 2> gpurun_out/w_deltas_$k.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/w_*.json')):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1])
        r=j['roofline']
        print(f, 'ms/step',round(j['ms_per_step'],3),'frac',round(r['frac'],3),'decode_ms',r.get('decode_kernel_ms'), 'reruns', j['config'].get('reruns'))
    except Exception as e:
        print(f,'ERR',e, open(f.replace('.json','.err')).read()[-400:])
PY
