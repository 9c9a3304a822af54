#!/bin/bash
show() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1', 'ms/step %.3f kernel_ms %.3f frac %.3f slow %s reruns %s' % (d['ms_per_step'], r['kernel_ms'], r['frac'], d['config']['slow_tiles'], d['config']['reruns']))" || tail -5 gpurun_out/last.log; }
run() { name=$1; shift; "$@" > gpurun_out/last.log 2>&1; cat gpurun_out/last.log | show $name; }
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --hist-slots-per-cpu 524288"
export HBT_MIN_TILES_PER_SEG=2
FL="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --fmad=false -Xcompiler -fPIC -shared"
nvcc $FL -DHBT_KPENDING=0 -DHBT_HIST_CTAS_PER_SM=4 -o /tmp/lib0.so dynolog-1_b200/csrc/hbt_b200.cu 2>&1 | grep -E "error"
export HBT_B200_LIB=/tmp/lib0.so
for s in 64 128 256; do
  HBT_NO_ZERO_FIRST=1 HBT_SEGS_PER_CPU=$s run sync_memset_segs$s $B
done
HBT_NO_ZERO_FIRST=1 HBT_SEGS_PER_CPU=256 run sync_memset_segs256_256k python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --hist-slots-per-cpu 262144
HBT_NO_ZERO_FIRST=1 HBT_SEGS_PER_CPU=128 run sync_memset_segs128_256k python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --hist-slots-per-cpu 262144
