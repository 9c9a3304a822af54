B="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-diagnostics"
for w in c4_mixed c4_switch; do for s in 12 24 36 48 56 80 128; do
timeout 300 $B --workload $w --segs-per-cpu $s > gpurun_out/sw_$w.json 2> gpurun_out/sw_$w.err
python - $w $s <<'PY'
import json,sys
j=json.loads(open('gpurun_out/sw_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1],'segs',sys.argv[2],'ms/step',round(j['ms_per_step'],3),'frac',round(j['roofline']['frac'],3),'decode',j['roofline'].get('decode_kernel_ms'))
PY
done; done
