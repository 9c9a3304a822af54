#!/usr/bin/env python
"""Generate tests/golden/builtin_metrics.json from the REFERENCE's own metric registry
(makeAvailableMetrics, hbt/src/perf_event/BuiltinMetrics.cpp:470-1175): oracle/dump_builtin_metrics.cpp
is compiled against the reference sources where they lie under /root/reference (outputs only into
oracle/_ref/) and its output committed.  Run in the build container:

    python tests/golden/make_builtin_metrics.py

Only the two translation units that define the registry are compiled; everything they reference but
the dump never calls (PMU discovery, the generated Intel event tables) is left unresolved.
"""
import json
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
OUT = os.path.join(ROOT, "oracle", "_ref")


def main():
    os.makedirs(OUT, exist_ok=True)
    inc = [f"-I{REF}"] + [f"-I{REF}/third_party/{d}/include" for d in ("fmt", "pfs", "range-v3")]
    exe = os.path.join(OUT, "dump_builtin_metrics")
    subprocess.check_call(
        ["g++", "-std=c++20", "-O1", "-DNDEBUG", "-DFMT_HEADER_ONLY", "-w", *inc, "-o", exe,
         os.path.join(ROOT, "oracle", "dump_builtin_metrics.cpp"),
         f"{REF}/hbt/src/perf_event/BuiltinMetrics.cpp", f"{REF}/hbt/src/perf_event/Metrics.cpp",
         "-Wl,--unresolved-symbols=ignore-all"])
    table = json.loads(subprocess.check_output([exe]))
    with open(os.path.join(HERE, "builtin_metrics.json"), "w") as f:
        json.dump(table, f, indent=1, sort_keys=True)
        f.write("\n")
    print(f"{len(table)} metric descriptors")


if __name__ == "__main__":
    main()
