for w in c4_mixed c4_switch; do
timeout 900 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-diagnostics --workload $w --cpus 32 --ring-mib 1024 > gpurun_out/c4full_$w.json 2> gpurun_out/c4full_$w.err
python - $w <<'PY'
import json,sys
try:
    j=json.loads(open('gpurun_out/c4full_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1],'ms/step',round(j['ms_per_step'],3),'frac',round(j['roofline']['frac'],3),'value',j['value'],'merge',j['config'].get('merge_ms'),'levels',j['config'].get('merge_levels'),'reruns',j['config'].get('reruns'))
except Exception as e:
    print('ERR', open('gpurun_out/c4full_%s.err'%sys.argv[1]).read()[-500:])
PY
done
