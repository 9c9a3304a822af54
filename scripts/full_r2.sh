set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -6
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_switch_events.py "tests/test_gpu_parity.py::test_mixed_stream_deltas" "tests/test_gpu_parity.py::test_payload_that_looks_like_headers" "tests/test_gpu_parity.py::test_unknown_types_and_giant_records" "tests/test_gpu_parity.py::test_timestamp_regression_and_bad_size" -x -q -m gpu -k walk 2>&1 | tail -8
timeout 900 python bench.py > gpurun_out/bench_r2_default.json 2> gpurun_out/bench_r2_default.err; tail -c 600 gpurun_out/bench_r2_default.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
