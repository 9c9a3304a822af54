B="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-diagnostics"
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_sw_merge -s 4 -c 1 -o gpurun_out/sw_merge -f $B --workload c4_switch > /dev/null 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_walk -s 2 -c 1 -o gpurun_out/walk_switch -f $B --workload c4_switch > /dev/null 2>&1
ls -la gpurun_out/*.ncu-rep
