#!/usr/bin/env python3
"""Time to first byte of a fleet-sized scrape: kgx_render (whole body to host, then serve) against kgx_render_device +
kgx_body_read chunk by chunk (chunked transfer).  BASELINE config 4, X mode, pinned-free host buffers (what a Go
handler has).  Prints milliseconds; run on the GPU box."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kgx_loader  # noqa: E402

kgx_loader.load()
import numpy as np  # noqa: E402
from kgx_b200 import kgx  # noqa: E402
from kgx_b200.abi import MODE_X  # noqa: E402
from kgx_b200.fleet import Fleet  # noqa: E402

nodes = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
f = Fleet.config(4, n_nodes=nodes)
g = kgx.Kgx(MODE_X)
g.set_table(f.table(MODE_X))
g.ingest(f.scrape(0))
n = g.render_device()
whole = np.empty(n + n // 8, dtype=np.uint8)
for chunk_bytes in (1 << 20, 8 << 20, 64 << 20):
    chunk = np.empty(chunk_bytes, dtype=np.uint8)
    first, total = [], []
    for _ in range(3):
        g.sync()
        t0 = time.perf_counter()
        g.render_device()
        got = g.body_read(0, chunk)
        t1 = time.perf_counter()
        at = got
        while got:
            got = g.body_read(at, chunk)
            at += got
        t2 = time.perf_counter()
        assert at == n
        first.append((t1 - t0) * 1e3); total.append((t2 - t0) * 1e3)
    print("chunks of %3d MiB: first chunk after %.2f ms, whole body (%.2f GB) after %.1f ms" % (chunk_bytes >> 20, min(first), n / 1e9, min(total)))
t = []
for _ in range(3):
    g.sync()
    t0 = time.perf_counter()
    ln = kgx.C.c_size_t()
    rc = kgx.lib().kgx_render(g._h, whole.ctypes.data, whole.nbytes, kgx.C.byref(ln))
    assert rc == 0 and ln.value == n
    t.append((time.perf_counter() - t0) * 1e3)
print("kgx_render into pageable host memory: first byte = whole body after %.1f ms" % min(t))
