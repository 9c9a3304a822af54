B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-diagnostics --workload c5_counts --cpus 512 --events 1 --intervals 1000000"
timeout 600 $B > gpurun_out/c5_plain.json 2> gpurun_out/c5_plain.err || { tail -5 gpurun_out/c5_plain.err; exit 1; }
timeout 900 ncu --set full --import-source on --clock-control none -k regex:'k_counts' -s 4 -c 2 -o gpurun_out/r2_c5_counts -f $B > /dev/null 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_' -c 100 --csv --log-file gpurun_out/r2_launches_c5.csv $B > /dev/null 2>&1
ls -la gpurun_out/r2_c5*; tail -c 900 gpurun_out/c5_plain.json
