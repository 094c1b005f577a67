#!/usr/bin/env python3
"""Can the join of scrape k+1 run under the writer of scrape k?  Two contexts with the same table on one GPU, each on
its own stream and thread: one renders, the other resets + ingests.  Prints the per-call time of each alone and of
both together (wall clock around blocking calls; ctypes drops the GIL).  Experiment only, no library change needed."""
import ctypes as C
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kgx_loader  # noqa: E402

kgx_loader.load()
import numpy as np  # noqa: E402
import torch  # noqa: E402
from kgx_b200 import abi, kgx  # noqa: E402
from kgx_b200.abi import MODE_X  # noqa: E402
from kgx_b200.fleet import Fleet  # noqa: E402

nodes = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
K = 20
f = Fleet.config(4, n_nodes=nodes)
table, cols = f.table(MODE_X), f.scrape(0)
dev = torch.device("cuda", 0)
ctxs = []
for _ in range(2):
    st = torch.cuda.Stream(device=dev)
    g = kgx.Kgx(MODE_X, device=0, stream=st.cuda_stream)
    g.set_table(table)
    ctxs.append((g, st))


def dev_t(a):
    return None if a is None else torch.from_numpy(a.view(np.int64) if a.dtype == np.uint64 else a.view(np.int32)).to(dev)


dcols = {k: dev_t(getattr(cols, k)) for k in ("pid", "gpu", "used", "node", "sm_util", "mem_util", "ts_us", "node_off", "gpu_total")}
ds = abi.kgx_cols()
ds.n = cols.n
for k, t in dcols.items():
    if t is not None:
        setattr(ds, k, C.cast(t.data_ptr(), abi.u64p if t.dtype == torch.int64 else abi.u32p))
for g, _ in ctxs:
    for _ in range(3):
        g.reset_window()
        g.ingest_device(ds)
        g.render_device()
torch.cuda.synchronize()


def renders(g):
    for _ in range(K):
        g.render_device()


def ingests(g):
    for _ in range(K):
        g.reset_window()
        g.ingest_device(ds)


def timed(fns):
    torch.cuda.synchronize()
    th = [threading.Thread(target=fn, args=(g,)) for fn, g in fns]
    t0 = time.perf_counter()
    for t in th:
        t.start()
    for t in th:
        t.join()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / K * 1e3


for rep in range(2):
    r = timed([(renders, ctxs[0][0])])
    i = timed([(ingests, ctxs[1][0])])
    both = timed([(renders, ctxs[0][0]), (ingests, ctxs[1][0])])
    print("render alone %.3f ms  ingest alone %.3f ms  sum %.3f  both together %.3f ms per pair" % (r, i, r + i, both))
