#!/usr/bin/env python
"""Generate tests/golden/metric_frame.npz: a metric-frame case evaluated by the REFERENCE's own
MetricFrameVector / MetricFrameTsUnitFixInterval / MetricSeriesSlice (oracle/_ref).  Run in the build
container: python tests/golden/make_golden_frame.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle as orc  # noqa: E402
from test_metric_frame import random_case  # noqa: E402


def main():
    assert orc.have_ref()
    n_rows, S, frame_len, interval_us = 400, 5, 128, 500_000
    rows, isd, t_last, queries = random_case(99, n_rows, S, frame_len, interval_us, 50)
    period_ns, pct = 10**9, 0.37
    length, ranges, o = orc.ref_frame_query(rows, isd, frame_len, interval_us, t_last, queries, period_ns, pct)
    ok = np.array([r is not None for r in ranges], np.uint8)
    ra = np.array([r if r is not None else (0, 0, 0, 0) for r in ranges], dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "metric_frame.npz"), rows=rows, is_double=isd, frame_len=frame_len,
                        interval_us=interval_us, t_last_ns=t_last, queries=np.array(queries, dtype=np.int64),
                        period_ns=period_ns, percentage=np.float32(pct), length=length, ok=ok, ranges=ra, **o)
    print("metric_frame.npz:", length, "samples held,", int(ok.sum()), "of", len(queries), "queries match")


if __name__ == "__main__":
    main()
