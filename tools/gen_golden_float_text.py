#!/usr/bin/env python3
"""Writes tests/golden/float_text.npz: doubles and the exact text expfmt would
print for them, produced from CPython's repr() (independent of both the oracle
and the CUDA printer) through tests/goformat_model.py."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from goformat_model import go_format  # noqa: E402

rng = np.random.default_rng(20261106)
bits = rng.integers(0, 2**64, size=20000, dtype=np.uint64)
used = rng.integers(1, 65537, size=5000).astype(np.uint64) * np.uint64(2 << 20)
vals = np.concatenate([
    bits.view(np.float64), used.astype(np.float64), used.astype(np.float64) / np.float64(192 << 30) * 100,
    rng.integers(0, 5000, 5000) / rng.integers(1, 50, 5000), np.ldexp(1.0, np.arange(-1074, 1024)),
    np.array([0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, np.nan, 5e-324, 1.7976931348623157e308, 1e21, 1e-5, 1e-4, 999999.0, 1e6]),
])
text = np.zeros((len(vals), 32), dtype=np.uint8)
for i, v in enumerate(vals):
    s = go_format(float(v)).encode()
    text[i, :len(s)] = np.frombuffer(s, dtype=np.uint8)
    text[i, 31] = len(s)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "float_text.npz"), bits=vals.view(np.uint64), text=text)
print("wrote", len(vals), "vectors")
