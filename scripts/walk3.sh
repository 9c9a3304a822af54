timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_switch_events.py tests/test_golden.py -x -q -m gpu 2>&1 | tail -5
B="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-diagnostics"
for w in c4_switch c4_mixed; do
timeout 300 $B --workload $w > gpurun_out/w3_$w.json 2> gpurun_out/w3_$w.err
python - $w <<'PY'
import json,sys
j=json.loads(open('gpurun_out/w3_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1],'ms/step',j['ms_per_step'],'frac',j['roofline']['frac'],'decode',j['roofline'].get('decode_kernel_ms'),'slow',j['config'].get('slow_tiles'),'reruns',j['config'].get('reruns'))
PY
done
B2="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-diagnostics"
timeout 900 ncu --set full --import-source on --clock-control none -k regex:'k_decode|k_hist_aggregate' -s 12 -c 4 -o gpurun_out/r2_c2_hist -f $B2 > /dev/null 2>&1
ls -la gpurun_out/r2_c2_hist.ncu-rep
