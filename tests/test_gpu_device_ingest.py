"""kgx_ingest_device: sample columns already resident on the GPU.  The call returns with the join in flight; the next
call settles it (statistics, churn).  A render queued right behind an unsettled batch checks on the device that the join
met no new series and otherwise gives way to the churn path — the sequence bench.py times.  Every body is compared with
the CPU oracle's."""
import ctypes as C

import numpy as np
import pytest
import torch

from kgx_b200 import abi, build, kgx
from kgx_b200.abi import Cols, MODE_R_INTENDED, MODE_X
from kgx_b200.fleet import Fleet

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _built():
    build.build()


def _device_cols(cols):
    keep = {}
    ds = abi.kgx_cols()
    ds.n = cols.n
    for k in ("pid", "gpu", "used", "node", "sm_util", "mem_util", "ts_us", "node_off", "gpu_total"):
        a = getattr(cols, k)
        if a is None:
            continue
        t = torch.from_numpy(a.view(np.int64) if a.dtype == np.uint64 else a.view(np.int32)).cuda()
        keep[k] = t
        setattr(ds, k, C.cast(t.data_ptr(), abi.u64p if t.dtype == torch.int64 else abi.u32p))
    return ds, keep


def _body(g):
    n = g.render_device()
    out = np.empty(max(1, n), dtype=np.uint8)
    assert g.render_into(out) == n
    return out[:n].tobytes()


@pytest.mark.parametrize("mode", [MODE_R_INTENDED, MODE_X])
@pytest.mark.parametrize("grouped", [True, False])
def test_device_ingest_render_sequences(kgo, mode, grouped):
    f = Fleet(n_nodes=11, pods=60, procs=500, seed=0xB200 + 51, scrapes=5)
    t = f.table(mode)
    g, o = kgx.Kgx(mode), kgo.Oracle(mode, kgo.JOIN_HASH, 2)
    g.set_table(t); o.set_table(t)

    def scrape(k, part=None):
        s = f.scrape(k, shuffle_nodes=not grouped)
        if part is not None:
            keep = (s.pid % part) != 0
            s = Cols(pid=s.pid[keep], gpu=s.gpu[keep], used=s.used[keep], node=s.node[keep], sm_util=s.sm_util[keep],
                     mem_util=s.mem_util[keep], ts_us=s.ts_us[keep], gpu_total=s.gpu_total)
        return s

    alive = []
    # 1. first batch brings every series (churn): the render queued behind it must give way to the index rebuild
    s = scrape(0, part=3)
    ds, k = _device_cols(s); alive.append(k)
    g.ingest_device(ds); o.ingest(s)
    assert _body(g) == o.render()
    st = g.stats()
    assert st["n_matched"] + st["n_dropped"] == s.n and st["n_series"] == o.series_count()
    # 2. steady state: reset (X), ingest, render — nothing new, no host round trip between join and writer
    for rnd in (1, 2):
        if mode == MODE_X:
            g.reset_window(); o.reset_window()
        s = scrape(rnd, part=3)
        ds, k = _device_cols(s); alive.append(k)
        g.ingest_device(ds); o.ingest(s)
        assert _body(g) == o.render()
    # 3. two batches back to back (the second call settles the first), the second one with new series again
    s1, s2 = scrape(3, part=3), scrape(4)
    d1, k1 = _device_cols(s1); d2, k2 = _device_cols(s2); alive += [k1, k2]
    g.ingest_device(d1); g.ingest_device(d2)
    o.ingest(s1); o.ingest(s2)
    st = g.stats()                                   # settles the second batch
    assert st["n_matched"] + st["n_dropped"] == s2.n and st["n_series"] == o.series_count()
    assert _body(g) == o.render()
    for fam in range(g.family_count()):
        assert np.array_equal(g.read_values(fam), o.read_values(fam))
    # 4. an empty batch, then a reset that nothing follows: the render itself has to carry it out
    empty = Cols(pid=[], gpu=[], used=[], node=[], sm_util=[], mem_util=[], ts_us=[], gpu_total=s2.gpu_total)
    de, ke = _device_cols(empty); alive.append(ke)
    g.ingest_device(de); o.ingest(empty)
    assert _body(g) == o.render()
    if mode == MODE_X:
        g.reset_window(); o.reset_window()
        assert _body(g) == o.render() == b""
