"""Parity at BASELINE.json's full sizes, whole bodies byte for byte: the rack
replay (1 M samples), the 10 M-sample fleet (3.3 GB of text; the oracle on all
host threads takes a few seconds; sampled nodes and size-independent
properties — line counts, linearity of the sums — are checked as well), the
60-scrape window folded on both sides, and a single node under churn."""
import os

import numpy as np
import pytest

from kgx_b200 import build, kgx
from kgx_b200.abi import MODE_R_INTENDED, MODE_X
from kgx_b200.fleet import Fleet

pytestmark = pytest.mark.gpu
THREADS = min(16, os.cpu_count() or 1)


@pytest.fixture(scope="module", autouse=True)
def _built():
    build.build()


@pytest.mark.parametrize("mode", [MODE_R_INTENDED, MODE_X])
def test_rack_replay_1m_samples_bit_exact(kgo, mode):
    """configs[2]: 128 nodes, 128k pods, 1M process samples."""
    f = Fleet.config(3)
    t, s = f.table(mode), f.scrape(0)
    g, o = kgx.Kgx(mode), kgo.Oracle(mode, kgo.JOIN_HASH, THREADS)
    g.set_table(t); o.set_table(t)
    g.ingest(s); o.ingest(s)
    gb, ob = g.render(), o.render()
    assert len(gb) == len(ob) and gb == ob
    assert list(g.body_offsets()) == list(o.body_offsets())
    st = g.stats()
    assert st["n_matched"] + st["n_dropped"] == s.n and st["n_dropped"] == f.n_nodes * (f.procs // 100)  # the orphans


def test_fleet_replay_10m_samples_whole_body_and_invariants(kgo):
    """configs[3]: 2k nodes, 10M samples, full label set."""
    f = Fleet.config(4)
    g = kgx.Kgx(MODE_X)
    g.set_table(f.table(MODE_X))
    g.ingest(f.scrape(0))
    body = np.empty(g.render_device() + 16, dtype=np.uint8)
    n = g.render_into(body)
    body = body[:n]
    off = g.body_offsets()
    assert off[0] == 0 and off[-1] == n and (np.diff(off.astype(np.int64)) > 0).all()
    # every node's body is independent: the oracle renders a sample of nodes from a fleet of their own
    for lo, cnt in ((0, 3), (997, 2), (1998, 2)):
        sub = Fleet.config(4, n_nodes=cnt, node_base=lo)
        o = kgo.Oracle(MODE_X, kgo.JOIN_HASH, 2)
        o.set_table(sub.table(MODE_X))
        o.ingest(sub.scrape(0))
        want = o.render()
        got = body[int(off[lo]):int(off[lo + cnt])].tobytes()
        assert got == want, "bodies of nodes %d..%d differ" % (lo, lo + cnt)
    # size-independent properties of the whole 3.3 GB text
    st = g.stats()
    n_lines = int(np.count_nonzero(body == 10))
    assert n_lines == 6 * st["n_series"] + 2 * 6 * 2000            # 6 families x (series + HELP + TYPE per node)
    assert st["n_matched"] == 9_900_000 and st["n_dropped"] == 100_000
    starts = off[:-1].astype(np.int64)
    assert (body[starts] == ord("#")).all() and body[-1] == 10
    vals = g.read_values(3)                                         # pod_gpu_memory_usage = sum of used bytes
    assert len(vals) == st["n_series"]
    s = f.scrape(0)
    orphan_free = float(vals.sum())
    # linearity: the per-series sums add up to the sum over all joined samples (exact in u64, < 2^53 here)
    pid_ok = np.ones(s.n, dtype=bool)
    j = np.arange(f.procs)
    # orphans are every 100th process index of each node before the shuffle: recover them by their PID not being in the table
    tbl = f.table(MODE_X)
    key_tbl = tbl.podkey_node[tbl.entry_podkey].astype(np.uint64) << np.uint64(32) | tbl.entry_pid.astype(np.uint64)
    key_s = s.node.astype(np.uint64) << np.uint64(32) | s.pid.astype(np.uint64)
    joined = np.isin(key_s, key_tbl)
    assert int(joined.sum()) == st["n_matched"]
    assert orphan_free == float(s.used[joined].sum(dtype=np.uint64))
    # ... and the whole 3.3 GB body against the oracle on the full fleet (all host threads: a few seconds per pass)
    o = kgo.Oracle(MODE_X, kgo.JOIN_HASH, os.cpu_count() or 1)
    o.set_table(tbl)
    o.ingest(s)
    want = o.render_view()
    assert len(want) == n, "body sizes differ: gpu %d cpu %d" % (n, len(want))
    assert np.array_equal(g.body_offsets(), o.body_offsets())
    step = 1 << 28
    for lo in range(0, n, step):
        if not np.array_equal(body[lo:lo + step], want[lo:lo + step]):
            bad = lo + int(np.nonzero(body[lo:lo + step] != want[lo:lo + step])[0][0])
            raise AssertionError("bodies differ at byte %d: gpu %r cpu %r" % (bad, body[max(0, bad - 60):bad + 20].tobytes(),
                                                                            want[max(0, bad - 60):bad + 20].tobytes()))


def test_time_window_60_scrapes_sum_max_last(kgo):
    """configs[4]: 60 scrapes x 1M samples folded into one window, rendered once."""
    f = Fleet.config(5)
    t = f.table(MODE_X)
    g, o = kgx.Kgx(MODE_X), kgo.Oracle(MODE_X, kgo.JOIN_HASH, THREADS)
    g.set_table(t); o.set_table(t)
    for k in range(f.scrapes):
        s = f.scrape(k)
        g.ingest(s); o.ingest(s)
    gb, ob = g.render(), o.render()
    assert gb == ob and len(gb) > 100_000_000
    # idempotence: rendering again changes nothing; a new window starts empty
    assert g.render() == gb
    g.reset_window()
    assert g.render() == b""


def test_single_node_high_churn_16k_processes(kgo):
    """configs[1]: 1 node, 1k pods, 16k processes, the pod->PID map rebuilt with churn every cycle
    (pods come and go, their GaugeVec children persist): whole body byte-exact every cycle."""
    import time
    f = Fleet.config(2)
    base = f.table(MODE_R_INTENDED)
    g, o = kgx.Kgx(MODE_R_INTENDED), kgo.Oracle(MODE_R_INTENDED, kgo.JOIN_HASH, 1)
    rng = np.random.default_rng(2)
    n_pods = len(base.podkey_node)
    cycle_ms = []
    for cyc in range(5):
        keep = rng.random(n_pods) > 0.1 * (cyc > 0)          # 10% of the pods are replaced each cycle
        names = [("pod-0-%d" % i) if keep[i] else ("pod-0-%d-gen%d" % (i, cyc)) for i in range(n_pods)]
        from kgx_b200.abi import Table
        t = Table(1, f.G, base.podkey_node, names, base.entry_podkey, base.entry_pid)
        s = f.scrape(cyc, x_cols=False)
        t0 = time.perf_counter()
        g.set_table(t); g.ingest(s); gb = g.render()
        cycle_ms.append((time.perf_counter() - t0) * 1e3)
        o.set_table(t); o.ingest(s)
        assert gb == o.render(), "cycle %d" % cyc
    assert g.stats()["n_series"] > 15840                      # replaced pods left their series behind
    print("churn cycle ms (set_table + ingest + render, 16k processes):", ["%.2f" % m for m in cycle_ms])
