#!/usr/bin/env python3
"""Where a high-churn cycle spends its time at configs[1] (1 node, 16k processes):
set_table (host plan + carry-over + upload), ingest (index rebuild), render."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kgx_loader; kgx_loader.load()
from kgx_b200 import kgx
from kgx_b200.abi import MODE_R_INTENDED, Table
from kgx_b200.fleet import Fleet
f = Fleet.config(2)
base = f.table(MODE_R_INTENDED)
g = kgx.Kgx(MODE_R_INTENDED)
rng = np.random.default_rng(2)
n_pods = len(base.podkey_node)
for cyc in range(6):
    keep = rng.random(n_pods) > 0.1 * (cyc > 0)
    names = [("pod-0-%d" % i) if keep[i] else ("pod-0-%d-gen%d" % (i, cyc)) for i in range(n_pods)]
    t = Table(1, f.G, base.podkey_node, names, base.entry_podkey, base.entry_pid)
    s = f.scrape(cyc, x_cols=False)
    t0 = time.perf_counter(); g.set_table(t); t1 = time.perf_counter(); g.ingest(s); t2 = time.perf_counter(); b = g.render(); t3 = time.perf_counter()
    print("cycle %d: set_table %.2f ms, ingest %.2f ms, render %.2f ms (%d bytes, %d series)" % (cyc, (t1-t0)*1e3, (t2-t1)*1e3, (t3-t2)*1e3, len(b), g.stats()["n_series"]))
