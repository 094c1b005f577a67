"""kgx_b200 — B200-native join -> reduce -> render hot path of a per-pod GPU
metrics exporter (reference: dorkamotorka/kubernetes-gpu-exporter, main.go).

Loaded under the module name `kgx_b200` by kgx_loader.load() because the
directory name is not a Python identifier.  The compute lives in
csrc/libkgx.so (hand-written sm_100a CUDA behind the C ABI of include/kgx.h);
this package is the thin ctypes host side.  There is no CPU fallback."""
from . import abi  # noqa: F401
from .abi import Cols, Table, MODE_R_INTENDED, MODE_R_LITERAL, MODE_X  # noqa: F401
