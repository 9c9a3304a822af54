N=${1:-8}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
timeout 600 $T bench.py --gpus $N --steps 10 --warmup 3 --no-diagnostics --no-e2e --no-cpu-baseline --workload c5_counts --cpus 512 --events 1 --intervals 1000000 --shard > gpurun_out/bench_r2_c5_full_sharded_n$N.json 2> gpurun_out/c5f.err; tail -c 300 gpurun_out/c5f.err
timeout 600 $T bench.py --gpus $N --steps 10 --warmup 3 --no-diagnostics --no-e2e --no-cpu-baseline --workload c4_mixed --cpus 256 --shard > gpurun_out/bench_r2_c4_mixed_sharded_n$N.json 2> gpurun_out/c4s.err; tail -c 300 gpurun_out/c4s.err
for f in gpurun_out/bench_r2_c5_full_sharded_n$N.json gpurun_out/bench_r2_c4_mixed_sharded_n$N.json; do python - "$f" <<'PY'
import json,sys
try:
    j=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], {k:j.get(k) for k in ('value','unit','n_gpus','ms_per_step','scaling')}, j.get('roofline',{}).get('frac'), j.get('collective'), j['config'].get('cpus_per_gpu'))
except Exception as e: print('ERR',e)
PY
done
