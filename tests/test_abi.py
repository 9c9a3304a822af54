"""The C-ABI library loads and exports every symbol include/hbt_b200.h declares;
without a GPU nothing computes: hbt_create fails loudly with -ENODEV."""
import ctypes as C
import errno
import os
import re

import pytest

import dynolog_1_b200 as hbt

HEADER = os.path.join(hbt.REPO_ROOT, "include", "hbt_b200.h")


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"^\s*(?:const\s+)?[A-Za-z_][A-Za-z0-9_]*\s*\*?\s*(hbt_[a-z0-9_]+)\s*\(", text,
                       flags=re.M)
    return sorted(set(names))


def test_header_and_binding_agree(built):
    declared = declared_functions()
    bound = sorted(n for n, _, _ in hbt.EXPORTS)
    assert declared == bound, (set(declared) ^ set(bound))


def test_library_exports_every_declared_symbol(built):
    lib = C.CDLL(hbt.LIB_PATH)
    for name in declared_functions():
        assert hasattr(lib, name), name
    assert lib.hbt_abi_version() == hbt.ABI_VERSION


def test_struct_layouts_match_header():
    # sizes the C side static-asserts through its own use; the binding must agree
    assert C.sizeof(hbt.DecodeDesc) == 56
    assert hbt.SWITCH_EVENT_DTYPE.itemsize == 16  # sizeof(tagstack::Event), Event.h:222
    assert C.sizeof(hbt.CpuResult) == 8 * (4 + hbt.NUM_COUNTERS + 2)
    assert C.sizeof(hbt.HistEntry) == 24
    assert C.sizeof(hbt.TscParams) == 16


def test_no_cpu_fallback(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the -m gpu tests")
    lib = hbt.load_library()
    h = C.c_void_p()
    rc = lib.hbt_create(C.byref(h), 0, 8, 8)
    assert rc == -errno.ENODEV and not h
    with pytest.raises(hbt.HbtError):
        hbt.Context(0)


def test_create_argument_checks(built):
    lib = hbt.load_library()
    h = C.c_void_p()
    assert lib.hbt_create(C.byref(h), 0, hbt.MAX_CPUS + 1, 8) == -errno.EINVAL
    assert lib.hbt_create(C.byref(h), 0, 8, hbt.MAX_EVENTS + 1) == -errno.EINVAL
    assert lib.hbt_create(None, 0, 8, 8) == -errno.EINVAL


def test_oracle_is_not_linked_into_the_product(built):
    """The product library must not depend on anything under oracle/."""
    import subprocess
    out = subprocess.run(["ldd", hbt.LIB_PATH], stdout=subprocess.PIPE, text=True).stdout
    assert "oracle" not in out and "hbt_ref" not in out
    syms = subprocess.run(["nm", "-D", "--defined-only", hbt.LIB_PATH], stdout=subprocess.PIPE,
                          text=True).stdout
    assert "hbt_oracle_" not in syms and "hbtref_" not in syms
