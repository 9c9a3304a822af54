#!/bin/bash
timeout 900 python -m pytest tests/test_switch_events.py -m gpu -q -x > gpurun_out/pytest_sw.log 2>&1; tail -15 gpurun_out/pytest_sw.log
