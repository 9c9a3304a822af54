#!/bin/bash
# Round-end measurement recipe (B200_PROFILING.md): bench (both arms), then the ncu
# launch list and one full capture of the top kernels for the same command.
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
tail -c 300 gpurun_out/bench.err
rm -f gpurun_out/bench_diag.jsonl
for w in "c2_ip_hist --agg none" "c3_deltas --events 8" "c3_deltas --events 32" "c4_mixed --events 8" "c4_switch" "c5_counts --cpus 64 --events 8 --intervals 400000" "c5_counts --cpus 512 --events 1 --intervals 400000"; do
  python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload $w >> gpurun_out/bench_diag.jsonl 2>> gpurun_out/bench.err
done
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_decode|k_finalize|k_hist|k_seed|k_locate|k_counts|k_scaled|k_prepare|k_sw" -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
# one decode step = k_prepare, k_decode, k_finalize, 3 fold passes; steps 0-2 warm up, capture the 4th
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_decode|k_hist_aggregate" -s 12 -c 4 -o gpurun_out/prof_r1_top $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
