timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
B="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-diagnostics"
for w in c4_mixed c3_deltas c4_switch; do
timeout 300 $B --workload $w > gpurun_out/w4_$w.json 2> gpurun_out/w4_$w.err
python - $w <<'PY'
import json,sys
j=json.loads(open('gpurun_out/w4_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1],'ms/step',round(j['ms_per_step'],3),'frac',round(j['roofline']['frac'],3),'kernel_ms',j['roofline'].get('kernel_ms'))
PY
done
