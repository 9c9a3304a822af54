B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-diagnostics --workload c3_deltas"
timeout 600 $B > gpurun_out/c3_plain.json 2> gpurun_out/c3_plain.err || { tail -5 gpurun_out/c3_plain.err; exit 1; }
timeout 900 ncu --set full --import-source on --clock-control none -k regex:'k_decode' -s 3 -c 1 -o gpurun_out/r2_c3_deltas -f $B > /dev/null 2>&1
ls -la gpurun_out/r2_c3*
