"""Where an end-to-end step spends its time: per-call host timings of the pipelined exporter loop
(kgx_reset_window, kgx_ingest, kgx_render_begin, kgx_render_end) on the bench workload."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kgx_loader
kgx_loader.load()
from kgx_b200 import kgx, abi
from kgx_b200.fleet import Fleet
from kgx_b200.abi import MODE_X

f = Fleet.config(4)
t = f.table(MODE_X)
cols = f.scrape(0)
g = kgx.Kgx(MODE_X)
g.set_table(t)
host = {}
for k in ("pid", "gpu", "used", "node", "sm_util", "mem_util", "ts_us", "node_off", "gpu_total"):
    a = getattr(cols, k)
    if a is not None:
        host[k] = kgx.pinned_empty(len(a), a.dtype)
        host[k][:] = a
hcols = abi.Cols(**host)
g.ingest(hcols)
n = g.render_device()
outs = [kgx.pinned_empty(n + 4096, np.uint8) for _ in range(2)]
pending = False
T0 = time.perf_counter()
for k in range(12):
    t0 = time.perf_counter(); g.reset_window()
    t1 = time.perf_counter(); g.ingest(hcols)
    t2 = time.perf_counter(); g.render_begin(outs[k % 2])
    t3 = time.perf_counter()
    if pending:
        g.render_end()
    t4 = time.perf_counter()
    pending = True
    print("step %2d: reset %.2f  ingest %.2f  render_begin %.2f  render_end %.2f  total %.2f ms" %
          (k, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3, (t4 - t0) * 1e3))
g.render_end()
print("12 steps + drain: %.2f ms per step" % ((time.perf_counter() - T0) * 1e3 / 12))
