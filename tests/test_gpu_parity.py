"""GPU parity tests: the CUDA path behind the C ABI against the CPU oracle on
the same seeded inputs.  Bar: /metrics bytes bit-exact."""
import os

import numpy as np
import pytest

from goformat_model import go_format
from kgx_b200.abi import Cols, Table, MODE_R_INTENDED, MODE_R_LITERAL, MODE_X
from kgx_b200.fleet import Fleet, GIB
from kgx_b200 import kgx, build
from test_oracle_format import KAT
from test_oracle_semantics import _cols, _tbl

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module", autouse=True)
def _built():
    build.build()


def _pair(kgo, mode, threads=1):
    return kgx.Kgx(mode), kgo.Oracle(mode, kgo.JOIN_HASH, threads)


def _same(g, o):
    gb, ob = g.render(), o.render()
    if gb != ob:
        gl, ol = gb.split(b"\n"), ob.split(b"\n")
        for i, (a, b) in enumerate(zip(gl, ol)):
            assert a == b, "line %d differs:\n gpu: %r\n cpu: %r" % (i, a, b)
        assert len(gl) == len(ol), "line counts differ: gpu %d cpu %d" % (len(gl), len(ol))
    assert list(g.body_offsets()) == list(o.body_offsets())
    return gb


def test_float_printer_known_answers():
    g = kgx.Kgx(MODE_R_INTENDED)
    vals = np.array([v for v, _ in KAT], dtype=np.float64)
    out = g.format_values(vals)
    for (v, want), rec in zip(KAT, out):
        assert bytes(rec[:rec[31]]).decode() == want, (v, bytes(rec[:rec[31]]), want)


def test_float_printer_vs_oracle_and_repr(kgo):
    g = kgx.Kgx(MODE_R_INTENDED)
    rng = np.random.default_rng(0xB200)
    bits = rng.integers(0, 2**64, size=2_000_000, dtype=np.uint64)
    vals = bits.view(np.float64).copy()
    used = rng.integers(1, 65537, size=200_000).astype(np.uint64) * np.uint64(2 << 20)
    dom = np.concatenate([used.astype(np.float64), used.astype(np.float64) / np.float64(192 * GIB) * 100,
                          rng.integers(0, 5000, 200_000) / rng.integers(1, 50, 200_000),
                          rng.integers(1_790_000_000_000_000, 1_790_000_010_000_000_000, 200_000).astype(np.float64) / 1e6,
                          np.ldexp(1.0, np.arange(-1074, 1024)), np.arange(0, 4096, dtype=np.float64)])
    vals = np.concatenate([vals, dom, -dom[:1000]])
    got = g.format_values(vals)
    want = kgo.format_values(vals)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert len(bad) == 0, (vals[bad[0]].hex(), bytes(got[bad[0]]), bytes(want[bad[0]]))
    for i in rng.integers(0, len(vals), 20000):  # and against CPython's repr directly
        assert bytes(got[i][:got[i][31]]).decode() == go_format(float(vals[i]))


def test_float_printer_golden_fixture():
    fx = np.load(os.path.join(GOLD, "float_text.npz"))
    g = kgx.Kgx(MODE_R_INTENDED)
    got = g.format_values(fx["bits"].view(np.float64))
    assert (got == fx["text"]).all()


def test_whole_body_golden(kgo):
    g = kgx.Kgx(MODE_R_INTENDED)
    g.set_table(_tbl([(0, "trainer-0", [1, 4242, 77]), (0, "idle", [5, 6])]))
    g.ingest(_cols([(4242, 0, 1 * GIB), (999, 1, 5 * GIB)]))
    assert g.render() == open(os.path.join(GOLD, "r_single.txt"), "rb").read()
    st = g.stats()
    assert st["n_matched"] == 1 and st["n_dropped"] == 1 and st["n_series"] == 1


def test_empty_inputs(kgo):
    g, o = _pair(kgo, MODE_R_INTENDED)
    t = _tbl([(0, "p", [10])])
    g.set_table(t); o.set_table(t)
    for rows in ([(11, 0, 1)], []):
        c = _cols(rows)
        g.ingest(c); o.ingest(c)
        assert _same(g, o) == b""
    e = _tbl([])
    g2 = kgx.Kgx(MODE_R_INTENDED)
    g2.set_table(e)
    g2.ingest(_cols([(1, 0, 1)]))
    assert g2.render() == b""


def test_literal_golden(kgo):
    g, o = _pair(kgo, MODE_R_LITERAL)
    t = _tbl([(0, "p", [1000, 2000, 3000]), (0, "q", [7])])
    g.set_table(t); o.set_table(t)
    c = _cols([(0, 0, 10), (2, 0, 30), (5, 0, 60)])
    g.ingest(c); o.ingest(c)
    body = _same(g, o).decode()
    lines = [l for l in body.splitlines() if l.startswith("pod_gpu_memory_usage{")]
    assert lines == ['pod_gpu_memory_usage{pid="0",pod="p"} 10', 'pod_gpu_memory_usage{pid="0",pod="q"} 10',
                     'pod_gpu_memory_usage{pid="2",pod="p"} 30']


def test_last_write_wins_persistence_and_table_change(kgo):
    g, o = _pair(kgo, MODE_R_INTENDED)
    t1 = _tbl([(0, "p", [7, 8]), (0, "gone", [9])])
    g.set_table(t1); o.set_table(t1)
    c = _cols([(7, 0, 100), (7, 1, 200), (8, 0, 5), (9, 2, 77)])
    g.ingest(c); o.ingest(c)
    assert b'pod_gpu_memory_usage{pid="7",pod="p"} 200\n' in _same(g, o)
    # table rebuilt (main.go:84 every cycle): pid 8 and pod "gone" vanish, their children keep their values
    t2 = _tbl([(0, "p", [7]), (0, "new", [8])])
    g.set_table(t2); o.set_table(t2)
    _same(g, o)
    c = _cols([(7, 0, 300), (8, 0, 6)], total=40 * GIB)
    g.ingest(c); o.ingest(c)
    body = _same(g, o)
    for want in (b'{pid="7",pod="p"} 300\n', b'{pid="8",pod="p"} 5\n', b'{pid="8",pod="new"} 6\n', b'{pid="9",pod="gone"} 77\n'):
        assert b"pod_gpu_memory_usage" + want in body


def test_escaping_special_values_and_duplicates(kgo):
    g, o = _pair(kgo, MODE_R_INTENDED)
    t = _tbl([(0, 'we"ird\\pod\nx', [1, 1, 1]), (0, "zero-total", [2]), (0, "héllo-✓", [1, 3]), (0, "", [4])])
    g.set_table(t); o.set_table(t)
    c = _cols([(1, 0, 2**64 - 1), (2, 1, 0), (3, 0, 12345), (4, 0, 1), (2, 9, 5)])  # last row: gpu out of range
    c.gpu_total[1] = 0
    g.ingest(c); o.ingest(c)
    body = _same(g, o).decode()
    assert 'pod="we\\"ird\\\\pod\\nx"} 1.8446744073709552e+19' in body
    assert 'docker_gpu_memory_perc_usage{pid="2",pod="zero-total"} NaN' in body


def test_invalid_inputs_rejected():
    g = kgx.Kgx(MODE_R_INTENDED)
    with pytest.raises(kgx.KgxError):
        g.ingest(_cols([(1, 0, 1)]))  # no table yet
    with pytest.raises(kgx.KgxError):
        g.set_table(_tbl([(0, b"bad\xff", [1])]))  # invalid UTF-8: WithLabelValues would panic
    with pytest.raises(kgx.KgxError):
        g.set_table(_tbl([(0, "dup", [1]), (0, "dup", [2])]))
    with pytest.raises(kgx.KgxError):
        g.reset_window()
    g.set_table(_tbl([(0, "p", [1])]))
    small = np.zeros(8, dtype=np.uint8)
    g.ingest(_cols([(1, 0, 1)]))
    assert g.render_into(small) < 0  # KGX_E_NOSPACE reports the needed size


@pytest.mark.parametrize("mode", [MODE_R_INTENDED, MODE_R_LITERAL, MODE_X])
@pytest.mark.parametrize("cfg", [1, 2])
def test_fleet_configs_bit_exact(kgo, mode, cfg):
    f = Fleet.config(cfg)
    t = f.table(mode)
    g, o = _pair(kgo, mode, threads=4)
    g.set_table(t); o.set_table(t)
    s = f.scrape(0)
    if mode == MODE_R_LITERAL:
        s.pid = (s.pid % 40).astype(np.uint32)
    g.ingest(s); o.ingest(s)
    body = _same(g, o)
    assert len(body) > 1000
    st = g.stats()
    assert st["n_series"] == o.series_count()
    for fam in range(g.family_count()):
        assert np.array_equal(g.read_values(fam), o.read_values(fam))


@pytest.mark.parametrize("mode", [MODE_R_INTENDED, MODE_X])
@pytest.mark.parametrize("shuffle", [False, True])
def test_multi_node_fleet(kgo, mode, shuffle):
    f = Fleet(n_nodes=37, pods=50, procs=400, seed=0xB200 + 3)
    t = f.table(mode)
    g, o = _pair(kgo, mode, threads=4)
    g.set_table(t); o.set_table(t)
    s = f.scrape(0, shuffle_nodes=shuffle)
    if not shuffle:
        s.node = None  # exercise the node_off grouping path
    g.ingest(s)
    s2 = f.scrape(0, shuffle_nodes=shuffle)
    o.ingest(s2)
    _same(g, o)


@pytest.mark.parametrize("mode", [MODE_R_INTENDED, MODE_X])
def test_every_depth_of_the_writer_pipeline(kgo, mode):
    """The writer keeps three tiles in flight (emitted, measured, descriptor fetched): one to a dozen 128-line tiles per
    launch, series counts on and around the tile size, on one node and on three."""
    for nodes, procs in [(1, 1), (1, 2), (1, 127), (1, 128), (1, 129), (1, 255), (1, 256), (1, 257), (1, 385), (1, 640),
                         (3, 1), (3, 128), (3, 129), (3, 300)]:
        f = Fleet(n_nodes=nodes, pods=max(1, procs // 2), procs=procs, seed=0xB200 + 77 + procs)
        t = f.table(mode)
        g, o = _pair(kgo, mode)
        g.set_table(t); o.set_table(t)
        s = f.scrape(0)
        g.ingest(s); o.ingest(s)
        _same(g, o)
        _same(g, o)  # a second launch: the other ticket, the next epoch
        g.close()


def test_x_window_of_scrapes_and_reset(kgo):
    f = Fleet(n_nodes=4, pods=64, procs=512, seed=0xB200 + 5, scrapes=6)
    t = f.table(MODE_X)
    g, o = _pair(kgo, MODE_X, threads=2)
    g.set_table(t); o.set_table(t)
    for k in range(6):
        s = f.scrape(k)
        g.ingest(s); o.ingest(s)
        if k in (0, 3, 5):
            _same(g, o)
    g.reset_window(); o.reset_window()
    assert _same(g, o) == b""
    s = f.scrape(7)
    keep = (s.pid % 3) != 0  # some series stay empty in the new window: they must not be printed
    s = Cols(pid=s.pid[keep], gpu=s.gpu[keep], used=s.used[keep], node=s.node[keep], sm_util=s.sm_util[keep],
             mem_util=s.mem_util[keep], ts_us=s.ts_us[keep], gpu_total=s.gpu_total)
    g.ingest(s); o.ingest(s)
    assert len(_same(g, o)) > 0


def test_x_explicit_expiry(kgo):
    """kgx_expire: X-mode series whose newest sample is older than the cutoff leave the exposition and the index;
    R modes refuse (the reference never deletes a child)."""
    f = Fleet(n_nodes=5, pods=48, procs=400, seed=0xB200 + 9, scrapes=4)
    t = f.table(MODE_X)
    g, o = _pair(kgo, MODE_X, threads=2)
    g.set_table(t); o.set_table(t)
    base = 1_790_000_000_000_000  # fleet.py: scrape k is stamped base + 30 s * k (+ < 1 s of jitter)

    def part(s, keep):
        return Cols(pid=s.pid[keep], gpu=s.gpu[keep], used=s.used[keep], node=s.node[keep], sm_util=s.sm_util[keep],
                    mem_util=s.mem_util[keep], ts_us=s.ts_us[keep], gpu_total=s.gpu_total)

    s = f.scrape(0)
    g.ingest(s); o.ingest(s)
    s = f.scrape(1)
    s = part(s, (s.pid % 4) != 0)  # a quarter of the processes is gone by the second scrape
    g.ingest(s); o.ingest(s)
    _same(g, o)
    n0 = g.stats()["n_series"]
    assert g.expire(0) == o.expire(0) == 0
    ng, no = g.expire(base + 30_000_000), o.expire(base + 30_000_000)
    assert ng == no and 0 < ng < n0
    assert len(_same(g, o)) > 0
    assert g.stats()["n_series"] == n0 - ng == o.series_count()
    s = f.scrape(2)  # the expired keys come back: new series, counted from zero
    g.ingest(s); o.ingest(s)
    _same(g, o)
    for fam in range(g.family_count()):
        assert np.array_equal(g.read_values(fam), o.read_values(fam))
    total = o.series_count()
    ng, no = g.expire(2**63), o.expire(2**63)
    assert ng == no == total
    assert _same(g, o) == b"" and g.stats()["n_series"] == 0
    s = f.scrape(3)
    g.ingest(s); o.ingest(s)
    assert len(_same(g, o)) > 0
    g.reset_window(); o.reset_window()
    assert g.expire(1) == o.expire(1) == 0  # nothing exported goes, but the emptied series leave the index
    assert g.stats()["n_series"] == 0 and _same(g, o) == b""
    gr = kgx.Kgx(MODE_R_INTENDED)
    gr.set_table(f.table(MODE_R_INTENDED))
    with pytest.raises(kgx.KgxError):
        gr.expire(1)


def test_concurrent_scrapes_while_polling(kgo):
    """The reference serves /metrics from HTTP goroutines while its poll loop writes (main.go:68-74).  Here: one
    writer thread ingesting cycle after cycle, three reader threads rendering all the time.  Every body a reader
    gets must be exactly the oracle's body after SOME completed cycle (the library serialises calls on a ctx), and
    the last one must be reachable."""
    import threading
    f = Fleet(n_nodes=3, pods=40, procs=300, seed=0xB200 + 11, scrapes=8)
    t = f.table(MODE_R_INTENDED)
    g, o = _pair(kgo, MODE_R_INTENDED)
    g.set_table(t); o.set_table(t)
    scrapes = [f.scrape(k, x_cols=False) for k in range(8)]
    want = {o.render()}
    for rep in range(20):  # the same sequence the writer thread below goes through
        for s in scrapes:
            o.ingest(s)
            want.add(o.render())
    final = o.render()
    seen, errors, stop = [set() for _ in range(3)], [], threading.Event()

    def reader(i):
        try:
            buf = np.empty(1 << 20, dtype=np.uint8)  # every scraper has its own response buffer (Kgx.render() shares one)
            while not stop.is_set():
                n = g.render_into(buf)
                assert n >= 0
                b = buf[:n].tobytes()
                if b not in want:
                    errors.append("reader %d got a body that matches no completed cycle (%d bytes)" % (i, len(b)))
                    return
                seen[i].add(b)
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=reader, args=(i,)) for i in range(3)]
    for th in threads:
        th.start()
    for rep in range(20):
        for s in scrapes:
            g.ingest(s)
    stop.set()
    for th in threads:
        th.join()
    assert not errors, errors
    assert g.render() == final
    assert sum(len(s) for s in seen) >= 3  # the readers did run


def test_scrapes_do_not_wait_for_a_batch_in_transit(kgo):
    """kgx_ingest copies its batch to the GPU before it takes the lock that guards the series state, so a scrape that
    arrives while a large batch is crossing PCIe is served at once from the last completed reduction (the reference's
    HTTP goroutines never wait for the poll loop's I/O either, main.go:68-74).  One writer thread ingests 1 M-sample
    batches from pageable memory (milliseconds of copying each); a reader renders on the device all the time
    (a fraction of a millisecond each): some renders must start AND finish inside a single ingest call, and every
    body length must be the oracle's after a completed cycle."""
    import threading
    import time
    f = Fleet.config(3)
    t = f.table(MODE_X)
    g, o = kgx.Kgx(MODE_X), kgo.Oracle(MODE_X, kgo.JOIN_HASH, min(16, os.cpu_count() or 1))
    g.set_table(t); o.set_table(t)
    scrapes = [f.scrape(k) for k in range(4)]
    want, bodies = set(), []
    for s in scrapes:  # one window, four cycles
        o.ingest(s)
        bodies.append(o.render())
        want.add(len(bodies[-1]))
    g.ingest(scrapes[0])
    assert g.render() == bodies[0]
    spans, renders, errors, stop = [], [], [], threading.Event()

    def reader():
        try:
            while not stop.is_set():
                t0 = time.perf_counter()
                n = g.render_device()
                t1 = time.perf_counter()
                if n not in want:
                    errors.append("a body of %d bytes matches no completed cycle" % n)
                    return
                renders.append((t0, t1))
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    th = threading.Thread(target=reader)
    th.start()
    time.sleep(0.05)
    for s in scrapes[1:]:
        t0 = time.perf_counter()
        g.ingest(s)
        spans.append((t0, time.perf_counter()))
    stop.set()
    th.join()
    assert not errors, errors
    assert g.render() == bodies[-1]
    inside = sum(1 for a, b in renders for lo, hi in spans if lo < a and b < hi)
    assert inside >= 1, "no scrape was served while a batch was in transit (%d renders of ~%.2f ms, ingest calls took %s)" % (
        len(renders), 1e3 * float(np.mean([b - a for a, b in renders])) if renders else 0.0, ["%.1f ms" % ((hi - lo) * 1e3) for lo, hi in spans])


def test_launch_epoch_wraps_cleanly(kgo):
    """The writer's look-back words and per-node body offsets carry a 14-bit launch epoch instead of being cleared before
    every launch, and two tile tickets take turns by the epoch's parity.  Render more often than the epoch can count
    (16 383): bodies and offsets must stay right across the wrap, with state changes in between."""
    f = Fleet(n_nodes=3, pods=30, procs=200, seed=0xB200 + 41, scrapes=4)
    g, o = _pair(kgo, MODE_X)
    t = f.table(MODE_X)
    g.set_table(t); o.set_table(t)
    g.ingest(f.scrape(0)); o.ingest(f.scrape(0))
    want = _same(g, o)
    for i in range(16400):
        n = g.render_device()
        assert n == len(want), "render %d: %d bytes instead of %d" % (i, n, len(want))
        if i in (8000, 16380, 16383, 16384, 16390):
            s = f.scrape(1 + (i % 3))
            g.ingest(s); o.ingest(s)
            want = _same(g, o)
            assert list(g.body_offsets()) == list(o.body_offsets())
    assert g.render() == want


def test_r_staleness_over_cycles(kgo):
    f = Fleet(n_nodes=3, pods=40, procs=300, seed=0xB200 + 5)
    g, o = _pair(kgo, MODE_R_INTENDED)
    t = f.table(MODE_R_INTENDED)
    g.set_table(t); o.set_table(t)
    for k in range(4):
        s = f.scrape(k)
        keep = ((s.pid + k) % 4) != 0  # processes come and go; vanished ones keep their last value
        s = Cols(pid=s.pid[keep], gpu=s.gpu[keep], used=s.used[keep], node=s.node[keep], gpu_total=s.gpu_total)
        g.ingest(s); o.ingest(s)
        _same(g, o)


def test_multimap_pid_in_several_pods(kgo):
    for mode in (MODE_R_INTENDED, MODE_X):
        if mode == MODE_X:
            t = Table(1, 2, [0, 0, 0], ["a", "b", "c"], [0, 1, 2, 2], [1, 1, 1, 5], namespace=["n", "n", "n"],
                      container=["x", "x", "y"], gpu_uuid=["GPU-1", "GPU-0"])
            c = Cols(pid=[1, 5, 1], gpu=[0, 1, 1], used=[10, 20, 30], sm_util=[1, 2, 3], mem_util=[4, 5, 6],
                     ts_us=[7, 8, 9], gpu_total=[100, 200])
        else:
            t = _tbl([(0, "a", [1]), (0, "b", [1, 2]), (0, "c", [1])])  # `ps` sees PID 1 in every container
            c = _cols([(1, 0, 10), (2, 0, 20)])
        g, o = _pair(kgo, mode)
        g.set_table(t); o.set_table(t)
        g.ingest(c); o.ingest(c)
        assert len(_same(g, o)) > 0


def test_long_labels_and_many_lines_per_tile(kgo):
    pods = [(0, ("p%03d-" % i) + "x" * (i * 7 % 250), [1000 + i]) for i in range(300)]
    t = _tbl(pods)
    g, o = _pair(kgo, MODE_R_INTENDED)
    g.set_table(t); o.set_table(t)
    c = _cols([(1000 + i, i % 8, (i + 1) * 12345678901) for i in range(300)])
    g.ingest(c); o.ingest(c)
    _same(g, o)


def test_ingest_from_nvml_process_info_arrays(kgo):
    """The NVML wire format (nvmlProcessInfo_t arrays per device, main.go:135) gives the same body as columns."""
    from kgx_b200 import abi
    f = Fleet(n_nodes=5, pods=30, procs=200, seed=0xB200 + 2)
    t = f.table(MODE_R_INTENDED)
    s = f.scrape(0, x_cols=False)
    # regroup the samples the way the reference's loops see them: node, then device, then process order
    order = np.lexsort((np.arange(s.n), s.gpu, s.node))
    dev = s.node[order].astype(np.int64) * f.G + s.gpu[order]
    dev_off = np.zeros(f.n_nodes * f.G + 1, dtype=np.uint64)
    dev_off[1:] = np.cumsum(np.bincount(dev, minlength=f.n_nodes * f.G))
    infos = np.zeros(s.n, dtype=abi.NVML_PROCESS_INFO)
    infos["pid"], infos["used_gpu_memory"] = s.pid[order], s.used[order]
    infos["gpu_instance_id"] = 0xFFFFFFFF
    g, o = _pair(kgo, MODE_R_INTENDED)
    g.set_table(t); o.set_table(t)
    g.ingest_nvml(infos, dev_off, s.gpu_total)
    o.ingest(Cols(pid=s.pid[order], gpu=s.gpu[order], used=s.used[order], node=s.node[order], gpu_total=s.gpu_total))
    assert len(_same(g, o)) > 0
    gx = kgx.Kgx(MODE_X)
    with pytest.raises(kgx.KgxError):
        gx.ingest_nvml(infos, dev_off, s.gpu_total)


def test_x_ingest_from_nvml_info_and_utilization_arrays(kgo):
    """X mode from the two NVML wire formats (nvmlProcessInfo_t + nvmlProcessUtilizationSample_t per device): the GPU
    joins them by (device, pid); same body as the columns that join produces on the host."""
    from kgx_b200 import abi
    rng = np.random.default_rng(0xB200 + 77)
    f = Fleet(n_nodes=7, pods=40, procs=900, seed=0xB200 + 3)  # ~16 processes per device; one device made crowded below
    t = f.table(MODE_X)
    s = f.scrape(0)
    order = np.lexsort((np.arange(s.n), s.gpu, s.node))
    pid, gpu, node, used = s.pid[order], s.gpu[order], s.node[order], s.used[order]
    sm, mem, ts = s.sm_util[order].copy(), s.mem_util[order].copy(), s.ts_us[order].copy()
    gpu[node == 3] = 2  # every process of node 3 on one device; with the filler below it exceeds the shared-memory hash
    order2 = np.lexsort((np.arange(s.n), gpu, node))
    pid, gpu, node, used, sm, mem, ts = (a[order2] for a in (pid, gpu, node, used, sm, mem, ts))
    dev = node.astype(np.int64) * f.G + gpu
    n_dev = f.n_nodes * f.G
    dev_off = np.zeros(n_dev + 1, dtype=np.uint64)
    dev_off[1:] = np.cumsum(np.bincount(dev, minlength=n_dev))
    infos = np.zeros(s.n, dtype=abi.NVML_PROCESS_INFO)
    infos["pid"], infos["used_gpu_memory"] = pid, used
    # utilisation samples: per device, in another order; 10 % of the processes have none; some pids have two (last wins);
    # some samples belong to no process
    urows, want_sm, want_mem, want_ts = [], sm.copy(), mem.copy(), ts.copy()
    util_off = np.zeros(n_dev + 1, dtype=np.uint64)
    for d in range(n_dev):
        lo, hi = int(dev_off[d]), int(dev_off[d + 1])
        rows = []
        last = {}
        for i in rng.permutation(np.arange(lo, hi)):
            if rng.random() < 0.1:
                continue
            if rng.random() < 0.15:
                rows.append((pid[i], 1, 2, 3))       # an older sample of the same pid, superseded below
            rows.append((pid[i], ts[i], sm[i], mem[i]))
        for _ in range(3):
            rows.append((4_000_000_000 - d, 9, 9, 9))  # a process NVML has utilisation for but no memory record
        if d == 3 * f.G + 2:
            rows += [(5_000_000 + k, 1, 1, 1) for k in range(1100)]  # crowded device
        for j, r in enumerate(rows):
            last[int(r[0])] = r
        for i in range(lo, hi):
            r = last.get(int(pid[i]))
            want_ts[i], want_sm[i], want_mem[i] = (r[1], r[2], r[3]) if r else (0, 0, 0)
        urows += rows
        util_off[d + 1] = len(urows)
    utils = np.zeros(len(urows), dtype=abi.NVML_PROCESS_UTILIZATION_SAMPLE)
    utils["pid"] = [r[0] for r in urows]; utils["time_stamp"] = [r[1] for r in urows]
    utils["sm_util"] = [r[2] for r in urows]; utils["mem_util"] = [r[3] for r in urows]
    utils["enc_util"], utils["dec_util"] = 7, 8
    g, o = _pair(kgo, MODE_X)
    g.set_table(t); o.set_table(t)
    for _ in range(2):
        g.ingest_nvml_x(infos, dev_off, utils, util_off, s.gpu_total)
        o.ingest(Cols(pid=pid, gpu=gpu, used=used, node=node, sm_util=want_sm, mem_util=want_mem, ts_us=want_ts, gpu_total=s.gpu_total))
        assert len(_same(g, o)) > 0
    gr = kgx.Kgx(MODE_R_INTENDED)
    gr.set_table(f.table(MODE_R_INTENDED))
    with pytest.raises(kgx.KgxError):
        gr.ingest_nvml_x(infos, dev_off, utils, util_off, s.gpu_total)


def test_pipelined_render_begin_end(kgo):
    """kgx_render_begin/end: the body of scrape k is still being copied out while scrape k+1 is ingested."""
    f = Fleet(n_nodes=6, pods=60, procs=500, seed=0xB200 + 8, scrapes=4)
    t = f.table(MODE_X)
    g, o = _pair(kgo, MODE_X, threads=2)
    g.set_table(t); o.set_table(t)
    bufs = [kgx.pinned_empty(8 << 20, np.uint8) for _ in range(2)]
    want, got, lens = [], [], []
    for k in range(4):
        s = f.scrape(k)
        g.reset_window(); o.reset_window()
        g.ingest(s); o.ingest(s)            # overlaps the copy of body k-1
        if k:
            g.render_end()
            got.append(bufs[(k - 1) % 2][:lens[-1]].tobytes())
        lens.append(g.render_begin(bufs[k % 2]))
        want.append(o.render())
    g.render_end()
    got.append(bufs[3 % 2][:lens[-1]].tobytes())
    assert got == want
    # two bodies in flight: scrape k is rendered (into the library's second buffer) while body k-1 is still leaving;
    # kgx_render_end then delivers the older one.  A table change in between waits for both.
    want, got, lens = [], [], []
    for k in range(5):
        s = f.scrape(k % 4)
        if k == 3:
            g.set_table(t); o.set_table(t)
        g.reset_window(); o.reset_window()
        g.ingest(s); o.ingest(s)
        lens.append(g.render_begin(bufs[k % 2]))
        want.append(o.render())
        if k:
            g.render_end()
            got.append(bufs[(k - 1) % 2][:lens[k - 1]].tobytes())
    g.render_end()
    got.append(bufs[4 % 2][:lens[-1]].tobytes())
    assert got == want
    assert g.render() == want[-1]  # the synchronous call still works afterwards
    for b in bufs:
        kgx.pinned_free(b)
