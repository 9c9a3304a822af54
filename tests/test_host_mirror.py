"""The C++ mirror of the reference's host interface (dynolog-1_b200/host):
GroupReadValues / PerCpuCountReader / Monitor / PerfMonitor / Logger shaped classes
over the C-ABI.  The binary is built by __graft_entry__.build()."""
import os
import subprocess

import pytest

import dynolog_1_b200 as hbt

BIN = os.path.join(hbt.REPO_ROOT, "dynolog-1_b200", "host", "tests", "host_test")


def run(mode):
    r = subprocess.run([BIN, mode], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stdout[-3000:]
    assert f"{mode}: ok" in r.stdout


def test_host_mirror_cpu(built):
    """Argument checks, exceptions, JsonLogger formatting, loud failure without a GPU."""
    run("cpu")


@pytest.mark.gpu
def test_host_mirror_gpu(built):
    """PerfMonitor::step/log over a replayed trace (mips, mega_cycles_per_second, logUint),
    failed read -> nullopt, Monitor state machine + mux queue, ring decode through
    onCpuDataBufferRead-shaped slices."""
    run("gpu")
