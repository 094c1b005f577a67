"""Builds csrc/libkgx.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(CSRC, "libkgx.so")
SOURCES = ["kgx_api.cu", "kgx_table.cpp"]
DEPS = SOURCES + ["kgx_kernels.cuh", "kgx_writer.cuh", "kgx_writer5.cuh", "kgx_build.cuh", "kgx_build_host.cuh", "kgx_async.cuh", "kgx_ingest2.cuh", "kgx_fmt.cuh", "kgx_table.hpp", "pow10_table.inc", os.path.join("..", "..", "include", "kgx.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-Wall,-Wextra", "-shared", "-Xptxas", "-v"]


def needs_build():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in DEPS)


def build_variant(out_so, extra_flags):
    """Experiment builds (tools/): the same sources with extra -D flags, into another file."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + list(extra_flags) + ["-o", out_so] + [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        print(res.stdout)
        raise RuntimeError("nvcc failed building " + out_so)
    return out_so


def build(force=False, verbose=False):
    if not force and not needs_build():
        return SO
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + ["-o", SO] + [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed building libkgx.so")
    with open(os.path.join(CSRC, "ptxas.log"), "w") as f:
        f.write(res.stdout)
    return SO
