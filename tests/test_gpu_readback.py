"""Read-back entry points beside the renderer, through the C ABI, against the CPU oracle:
kgx_body_node (one node's /metrics body, what main.go:69-71 serves per scrape) and kgx_read_labels (the label
text of every series at once, for Collector.Collect — GaugeVec.Collect behind main.go:41-42)."""
import time

import numpy as np
import pytest

from kgx_b200 import abi, build, kgx
from kgx_b200.abi import Cols, MODE_R_INTENDED, MODE_X
from kgx_b200.fleet import Fleet

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _built():
    build.build()


def _labels_of_body(body, first_family_name):
    """Label text of every line of the first family, in body order (= exposition order, node after node)."""
    out = []
    pre = first_family_name.encode() + b"{"
    for line in body.split(b"\n"):
        if line.startswith(pre):
            out.append(line[len(pre) - 1:line.rindex(b" ")])
    return out


def _part(s, keep):
    return Cols(pid=s.pid[keep], gpu=s.gpu[keep], used=s.used[keep], node=s.node[keep], sm_util=s.sm_util[keep],
                mem_util=s.mem_util[keep], ts_us=s.ts_us[keep], gpu_total=s.gpu_total)


@pytest.mark.parametrize("mode", [MODE_R_INTENDED, MODE_X])
def test_body_node_is_the_nodes_slice(kgo, mode):
    f = Fleet(n_nodes=23, pods=40, procs=300, seed=0xB200 + 21)
    t = f.table(mode)
    g, o = kgx.Kgx(mode), kgo.Oracle(mode, kgo.JOIN_HASH, 2)
    g.set_table(t); o.set_table(t)
    with pytest.raises(kgx.KgxError) as e:      # nothing rendered yet
        g.body_node(0)
    assert e.value.code == abi.KGX_E_STATE
    s = f.scrape(0)
    s = _part(s, (s.node != 7) & (s.node != 22))  # two nodes stay empty, one of them the last
    g.ingest(s); o.ingest(s)
    body, want = g.render(), o.render()
    assert body == want
    off = o.body_offsets()
    for node in range(23):
        got = g.body_node(node)
        assert got == want[int(off[node]):int(off[node + 1])], "node %d" % node
        if node in (7, 22):
            assert got == b""
        else:
            assert got.startswith(b"# HELP ")
    assert b"".join(g.body_node(n) for n in range(23)) == want
    with pytest.raises(kgx.KgxError) as e:
        g.body_node(23)
    assert e.value.code == abi.KGX_E_INVAL
    n = kgx.C.c_size_t()                          # too small a buffer: the size comes back
    small = np.zeros(4, dtype=np.uint8)
    rc = kgx.lib().kgx_body_node(g._h, 0, small.ctypes.data, 4, kgx.C.byref(n))
    assert rc == abi.KGX_E_NOSPACE and n.value == int(off[1] - off[0])


def test_body_read_streams_the_same_bytes(kgo):
    """kgx_body_read: the body chunk by chunk (a chunked-transfer handler) equals kgx_render's, for chunk sizes that do
    and do not divide it; ranges, the end of the body and the error cases."""
    f = Fleet(n_nodes=11, pods=50, procs=400, seed=0xB200 + 23)
    t = f.table(MODE_X)
    g, o = kgx.Kgx(MODE_X), kgo.Oracle(MODE_X, kgo.JOIN_HASH, 2)
    g.set_table(t); o.set_table(t)
    buf = np.zeros(64, dtype=np.uint8)
    with pytest.raises(kgx.KgxError) as e:      # nothing rendered yet
        g.body_read(0, buf)
    assert e.value.code == abi.KGX_E_STATE
    s = f.scrape(0)
    g.ingest(s); o.ingest(s)
    want = o.render()
    assert g.render() == want
    for chunk in (1 << 20, 4096, 1000, len(want), len(want) - 1):
        parts = list(g.stream_body(chunk))
        assert b"".join(parts) == want
        assert len(parts) == (len(want) + chunk - 1) // chunk and all(len(p) == chunk for p in parts[:-1])
    mid = len(want) // 3
    assert g.body_read(mid, buf) == 64 and buf.tobytes() == want[mid:mid + 64]
    assert g.body_read(len(want) - 10, buf) == 10 and buf[:10].tobytes() == want[-10:]
    assert g.body_read(len(want), buf) == 0       # the end: no bytes, no error
    with pytest.raises(kgx.KgxError) as e:
        g.body_read(len(want) + 1, buf)
    assert e.value.code == abi.KGX_E_INVAL
    g.reset_window(); o.reset_window()            # the next scrape: the reader sees the new body
    s = f.scrape(1)
    g.ingest(s); o.ingest(s)
    assert b"".join(g.stream_body(8192)) == o.render()


@pytest.mark.parametrize("mode", [MODE_R_INTENDED, MODE_X])
def test_read_labels_in_one_call(kgo, mode):
    f = Fleet(n_nodes=9, pods=60, procs=500, seed=0xB200 + 22)
    t = f.table(mode)
    g, o = kgx.Kgx(mode), kgo.Oracle(mode, kgo.JOIN_HASH, 2)
    g.set_table(t); o.set_table(t)
    blob, off = g.read_labels()                  # before any sample: nothing
    assert len(blob) == 0 and list(off) == [0]
    s = f.scrape(0)
    g.ingest(s); o.ingest(s)
    name0 = g.family_info(0)[0]

    def check():
        want = _labels_of_body(o.render(), name0)
        blob, off = g.read_labels()
        got = [blob[int(off[i]):int(off[i + 1])].tobytes() for i in range(len(off) - 1)]
        assert got == want
        vals, slots = g.read_values(0, with_slots=True)
        assert len(vals) == len(got)
        for i in np.linspace(0, len(slots) - 1, 50).astype(int):    # same text as the per-slot call
            assert g.slot_labels(int(slots[i])) == got[i]
        return len(got)

    n_all = check()
    if mode == MODE_X:                            # a window in which only some series carry a value
        g.reset_window(); o.reset_window()
        s = f.scrape(1)
        s = _part(s, (s.pid % 3) != 0)
        g.ingest(s); o.ingest(s)
        assert 0 < check() < n_all
    n, nb = kgx.C.c_size_t(), kgx.C.c_size_t()    # buffers too small: KGX_E_NOSPACE and the sizes
    tiny = np.zeros(8, dtype=np.uint8)
    o8 = np.zeros(2, dtype=np.uint64)
    rc = kgx.lib().kgx_read_labels(g._h, tiny.ctypes.data, 8, o8.ctypes.data_as(abi.u64p), 2, kgx.C.byref(n), kgx.C.byref(nb))
    assert rc == abi.KGX_E_NOSPACE and n.value > 2 and nb.value > 8


def test_bulk_collect_of_a_million_series_is_fast(kgo):
    """Collector.Collect needs values and labels of every series: BASELINE config 3 (1 M series, R mode) in well under
    100 ms with the bulk call (the per-slot call costs several stream synchronisations per series)."""
    f = Fleet.config(3)
    g = kgx.Kgx(MODE_R_INTENDED)
    g.set_table(f.table(MODE_R_INTENDED))
    g.ingest(f.scrape(0))
    g.read_labels()                               # warm-up: buffers
    t0 = time.perf_counter()
    blob, off = g.read_labels()
    v0 = g.read_values(0)
    v1 = g.read_values(1)
    dt = time.perf_counter() - t0
    assert len(off) - 1 == len(v0) == len(v1) > 1_000_000
    assert dt < 0.5, "bulk read-back took %.3f s" % dt   # generous bound for a loaded test box; measured value in profiles/
    print("bulk collect of %d series: %.1f ms" % (len(v0), dt * 1e3))
    o = kgo.Oracle(MODE_R_INTENDED, kgo.JOIN_HASH, 8)
    o.set_table(f.table(MODE_R_INTENDED))
    o.ingest(f.scrape(0))
    want = _labels_of_body(o.render(), g.family_info(0)[0])
    idx = np.linspace(0, len(want) - 1, 5000).astype(int)
    for i in idx:
        assert blob[int(off[i]):int(off[i + 1])].tobytes() == want[i]
    assert np.array_equal(v0, o.read_values(0)) and np.array_equal(v1, o.read_values(1))
