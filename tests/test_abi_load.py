"""CPU-side checks of the drop-in boundary: the shared library loads, exports
every symbol include/kgx.h declares, and refuses to work without a GPU (no CPU
fallback)."""
import ctypes as C
import os
import re

import pytest
import torch

import kgx_b200.abi as abi
from kgx_b200 import kgx, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return kgx.lib()


def test_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(ROOT, "include", "kgx.h")).read()
    declared = set(re.findall(r"\b(kgx_[a-z_0-9]+)\s*\(", hdr))
    assert declared, "no declarations found"
    for name in sorted(declared):
        assert hasattr(lib, name), "libkgx.so does not export %s" % name
    assert set(kgx.EXPORTS) <= declared


def test_abi_version_and_strerror(lib):
    assert lib.kgx_abi_version() == abi.KGX_ABI_VERSION
    assert lib.kgx_strerror(0) == b"ok"
    assert b"CUDA" in lib.kgx_strerror(abi.KGX_E_CUDA)
    assert lib.kgx_strerror(-99) == b"unknown error"


def test_struct_sizes_match_header(tmp_path):
    # ask the C compiler for the layout of include/kgx.h and compare with the ctypes mirror
    import subprocess
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "kgx.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(kgx_opts),sizeof(kgx_table),sizeof(kgx_cols),sizeof(kgx_stats),sizeof(kgx_profile),'
                   '__builtin_offsetof(kgx_table,gpu_uuid_off),__builtin_offsetof(kgx_cols,gpu_total));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    want = [C.sizeof(abi.kgx_opts), C.sizeof(abi.kgx_table), C.sizeof(abi.kgx_cols), C.sizeof(abi.kgx_stats),
            C.sizeof(abi.kgx_profile), abi.kgx_table.gpu_uuid_off.offset, abi.kgx_cols.gpu_total.offset]
    assert got == want


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback(lib):
    with pytest.raises(kgx.KgxError) as e:
        kgx.Kgx(abi.MODE_R_INTENDED)
    assert e.value.code == abi.KGX_E_CUDA


def test_bad_opts_rejected(lib):
    h = C.c_void_p()
    o = abi.kgx_opts(999, 0, None, 0, 0)
    assert lib.kgx_create(0, C.byref(o), C.byref(h)) == abi.KGX_E_INVAL
    o = abi.kgx_opts(abi.KGX_ABI_VERSION, 7, None, 0, 0)
    assert lib.kgx_create(0, C.byref(o), C.byref(h)) == abi.KGX_E_INVAL
