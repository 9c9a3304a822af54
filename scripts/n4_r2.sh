N=${1:-4}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
timeout 900 $T bench.py --gpus $N --steps 10 --warmup 3 --no-diagnostics > gpurun_out/bench_r2_n$N.json 2> gpurun_out/bench_r2_n$N.err; tail -c 300 gpurun_out/bench_r2_n$N.err
timeout 600 $T bench.py --gpus $N --steps 10 --warmup 3 --no-diagnostics --no-e2e --no-cpu-baseline --workload c5_counts --shard > gpurun_out/bench_r2_c5_sharded_n$N.json 2> gpurun_out/bench_r2_c5_n$N.err; tail -c 300 gpurun_out/bench_r2_c5_n$N.err
for f in gpurun_out/bench_r2_n$N.json gpurun_out/bench_r2_c5_sharded_n$N.json; do python - "$f" <<'PY'
import json,sys
try:
    j=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], {k:j.get(k) for k in ('value','unit','n_gpus','ms_per_step','scaling')}, j.get('roofline',{}).get('frac'), j.get('collective'), (j.get('e2e') or {}).get('value'))
except Exception as e: print('ERR',e)
PY
done
