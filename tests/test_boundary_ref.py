"""The drop-in boundary driven through the REFERENCE's own classes (oracle/boundary_ref.cpp, built
against the reference headers into oracle/_ref/boundary_ref where the reference tree is present;
the binary travels to the GPU box): the real CpuCountSampleGenerator's onCpuDataBufferRead feeds
hbt_submit_ring, and the GPU results must equal the same object's own consume(), counters, output
ring and members; GPU counter-group sums land in a real GroupReadValues<mode::Counting> and go
through the reference's unmodified MetricSeries."""
import os
import subprocess

import pytest

import dynolog_1_b200 as hbt

EXE = os.path.join(hbt.REPO_ROOT, "oracle", "_ref", "boundary_ref")


def test_boundary_binary_refuses_to_run_without_a_gpu(built):
    """No CPU fallback behind the C-ABI: without a device hbt_create fails and the test binary
    says so (exit code 2) instead of computing anything."""
    if not os.path.exists(EXE):
        pytest.skip("oracle/_ref/boundary_ref not built (reference tree absent)")
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: covered by the gpu test below")
    r = subprocess.run([EXE], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r.returncode == 2 and "hbt_create failed" in r.stdout


@pytest.mark.gpu
def test_boundary_with_reference_types(built):
    assert os.path.exists(EXE), "oracle/_ref/boundary_ref is missing: it is built where /root/reference exists"
    r = subprocess.run([EXE], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    print(r.stdout)
    assert r.returncode == 0, r.stdout[-4000:]
    assert "boundary ok" in r.stdout
