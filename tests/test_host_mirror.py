"""The C++ mirror of the reference's host interface (dynolog-1_b200/host):
GroupReadValues / PerCpuCountReader / Monitor / PerfMonitor / Logger shaped classes
over the C-ABI.  The binary is built by __graft_entry__.build()."""
import os
import subprocess

import pytest

import dynolog_1_b200 as hbt

BIN = os.path.join(hbt.REPO_ROOT, "dynolog-1_b200", "host", "tests", "host_test")


REF_LIB = os.path.join(hbt.REPO_ROOT, "oracle", "_ref", "libhbt_ref.so")


def run(mode):
    env = dict(os.environ)
    if os.path.exists(REF_LIB):
        # lets the row-sink test compare with the reference's own ring buffer
        env["HBT_REF_LIB"] = REF_LIB
    r = subprocess.run([BIN, mode], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-3000:]
    assert f"{mode}: ok" in r.stdout
    return r.stdout


def test_host_mirror_cpu(built):
    """Argument checks, exceptions, JsonLogger formatting, loud failure without a GPU;
    count-sample row sink (capacity, -ENOSPC, drop-if-full) against the reference's
    ringbuffer::Producer/Consumer, CountData, applyToCountSamplesAndConsume / releaseCountData."""
    out = run("cpu")
    if os.path.exists(REF_LIB):
        assert "row sink vs reference ring buffer: 44 cases" in out


@pytest.mark.gpu
def test_host_mirror_gpu(built):
    """PerfMonitor::step/log over a replayed trace (mips, mega_cycles_per_second, logUint),
    failed read -> nullopt, Monitor state machine + mux queue, ring decode through
    onCpuDataBufferRead-shaped slices, decoded rows through the TraceMonitor-shaped sink
    into the applyToCountSamplesAndConsume callback."""
    run("gpu")
