# Round-2 ncu evidence: launch lists (our kernels only) and --set full of the dominant kernels.
# Each ncu command runs only after the same command has exited 0 without ncu.
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-diagnostics"
timeout 900 $B > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_' -c 400 --csv --log-file gpurun_out/r2_launches_c2_ip_hist.csv $B > /dev/null 2> gpurun_out/prof_launch.err
timeout 900 $B --workload c4_mixed > gpurun_out/prof_plain_mixed.json 2>> gpurun_out/prof_plain.err || exit 1
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_' -c 200 --csv --log-file gpurun_out/r2_launches_c4_mixed.csv $B --workload c4_mixed > /dev/null 2>&1
timeout 900 $B --workload c4_switch > gpurun_out/prof_plain_switch.json 2>> gpurun_out/prof_plain.err || exit 1
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_' -c 200 --csv --log-file gpurun_out/r2_launches_c4_switch.csv $B --workload c4_switch > /dev/null 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_walk -s 3 -c 1 -o gpurun_out/r2_walk_mixed -f $B --workload c4_mixed > /dev/null 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:'k_walk|k_sw_scan|k_sw_emit' -s 9 -c 3 -o gpurun_out/r2_walk_switch -f $B --workload c4_switch > /dev/null 2>&1
ls -la gpurun_out/r2_*
