#!/bin/bash
python bench.py --steps 5 --warmup 3 --workload c4_switch > gpurun_out/last.log 2>&1; tail -2 gpurun_out/last.log | cut -c1-900
