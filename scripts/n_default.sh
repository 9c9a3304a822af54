N=${1:-4}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
timeout 900 $T bench.py --gpus $N --steps 10 --warmup 3 --no-diagnostics > gpurun_out/bench_r2_n$N.json 2> gpurun_out/bench_r2_n$N.err; tail -c 300 gpurun_out/bench_r2_n$N.err
python - gpurun_out/bench_r2_n$N.json <<'PY'
import json,sys
j=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); e=j['e2e']
print({k:j.get(k) for k in ('value','n_gpus','ms_per_step')}, 'e2e', e['value'], round(e['ms_per_step'],1), 'upload-only', round(e['upload_only_ms_per_step'],1))
PY
