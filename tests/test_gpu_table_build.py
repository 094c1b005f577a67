"""kgx_set_table builds its plan on the GPU (csrc/kgx_build.cuh); the host builder
(csrc/kgx_table.cpp, KGX_HOST_TABLE_BUILD=1) is kept beside it.  One scenario
script — awkward names, decimal pid order, duplicates, empty nodes, pods that
only carried series still name, literal join, X-mode label pieces, rejected
tables — runs once per builder in a fresh process.  Each run checks every body
against the CPU oracle; the two runs must also print identical digests of the
bodies, of every slot's label text and of the values read back."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import hashlib, sys
sys.path.insert(0, %r)
sys.path.insert(0, %r)
import numpy as np
import kgx_loader
kgx_loader.load()
kgo = kgx_loader.load_oracle()
from kgx_b200 import kgx
from kgx_b200.abi import Cols, Table, MODE_R_INTENDED, MODE_R_LITERAL, MODE_X
from test_oracle_semantics import _cols, _tbl

GIB = 1 << 30
digest = hashlib.sha256()

def check(g, o, tag):
    gb, ob = g.render(), o.render()
    assert gb == ob, (tag, gb[:400], ob[:400])
    assert list(g.body_offsets()) == list(o.body_offsets()), tag
    digest.update(gb)
    st = g.stats()
    for s in range(st["n_slots"]):
        digest.update(g.slot_labels(s) + b"|")
    for fam in range(g.family_count()):
        v, sl = g.read_values(fam, with_slots=True)
        digest.update(v.tobytes()); digest.update(sl.tobytes())
    print(tag, len(gb), st["n_slots"], st["n_series"])

def rejected(g, t, tag):
    try:
        g.set_table(t)
    except kgx.KgxError as e:
        print(tag, "rejected", e.code)
        return
    raise AssertionError(tag + ": table accepted")

# ---- R, intended join: names and pids whose byte order differs from their numeric / length order
for mode in (MODE_R_INTENDED, MODE_R_LITERAL):
    g, o = kgx.Kgx(mode), kgo.Oracle(mode, kgo.JOIN_HASH, 2)
    N = 4
    t1 = _tbl([(0, "a", [9, 10, 100, 5]), (0, "ab", [1000000, 4294967295, 5]), (0, "a-", [1, 1, 1, 5]),
               (0, 'we"ird\\pod\nx', [7]), (0, "héllo-✓", [8, 9]), (0, "", [11]), (0, "zz", [12, 13]),
               (2, "lonely", []), (2, "a", [9, 10]), (3, "p3", [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11])], n_nodes=N)
    rows1 = [(9, 0, 1 * GIB, 0), (10, 1, 2 * GIB, 0), (100, 2, 3 * GIB, 0), (5, 3, 4 * GIB, 0), (1000000, 0, 5, 0),
             (4294967295, 7, 2**64 - 1, 0), (1, 0, 6, 0), (7, 1, 7, 0), (8, 2, 8, 0), (11, 3, 9, 0), (12, 4, 10, 0), (13, 5, 11, 0),
             (9, 0, 12, 2), (10, 0, 13, 2), (3, 0, 14, 3), (11, 0, 15, 3), (2, 1, 16, 0), (0, 0, 17, 0), (77, 0, 18, 1)]
    g.set_table(t1); o.set_table(t1)
    check(g, o, "r%%d empty" %% mode)
    c = _cols(rows1, n_nodes=N)
    g.ingest(c); o.ingest(c)
    check(g, o, "r%%d t1" %% mode)
    # pods "a-" and "zz" leave the table: their series stay, named by pods only the old state knows
    t2 = _tbl([(0, "a", [9, 100]), (0, "ab", [5, 6]), (0, 'we"ird\\pod\nx', [7, 70]), (0, "b", [1, 12]), (0, "", [11]),
               (1, "new-on-1", [77]), (2, "a", [10]), (3, "p3", [0, 1])], n_nodes=N)
    g.set_table(t2); o.set_table(t2)
    check(g, o, "r%%d t2 carried" %% mode)
    c = _cols([(9, 0, 100, 0), (6, 0, 101, 0), (12, 1, 102, 0), (1, 2, 103, 0), (77, 0, 104, 1), (10, 0, 105, 2), (13, 0, 106, 0),
               (70, 0, 107, 0), (0, 0, 108, 3), (1, 0, 109, 0)], n_nodes=N)
    g.ingest(c); o.ingest(c)
    check(g, o, "r%%d t2 ingest" %% mode)
    # "zz" comes back with another pid list; "a" disappears from node 0 but not from node 2
    t3 = _tbl([(0, "zz", [13, 14]), (0, "ab", [5]), (2, "a", [9]), (3, "p3", [0, 1, 2])], n_nodes=N)
    g.set_table(t3); o.set_table(t3)
    c = _cols([(13, 0, 200, 0), (14, 0, 201, 0), (12, 0, 202, 0), (9, 0, 203, 2), (2, 0, 204, 3), (1, 0, 205, 0)], n_nodes=N)
    g.ingest(c); o.ingest(c)
    check(g, o, "r%%d t3" %% mode)
    # rejected tables leave everything as it was
    before = g.render()
    rejected(g, _tbl([(0, b"bad\xff", [1])], n_nodes=N), "r%%d utf8" %% mode)
    rejected(g, _tbl([(0, "dup", [1]), (0, "dup", [2])], n_nodes=N), "r%%d dup pod" %% mode)
    rejected(g, _tbl([(N, "node-out-of-range", [1])], n_nodes=N), "r%%d node" %% mode)
    rejected(g, Table(N, 8, [0], ["p"], [1], [5]), "r%%d podkey" %% mode)
    assert g.render() == before
    c = _cols([(14, 0, 300, 0)], n_nodes=N)
    g.ingest(c); o.ingest(c)
    check(g, o, "r%%d after rejects" %% mode)
    # an empty table: nothing joins any more, every series stays
    t4 = _tbl([], n_nodes=N)
    g.set_table(t4); o.set_table(t4)
    c = _cols([(14, 0, 400, 0)], n_nodes=N)
    g.ingest(c); o.ingest(c)
    check(g, o, "r%%d empty table" %% mode)

# ---- X: (container, gpu, namespace, pod) order, escapes in every label, gpu order by uuid text
N, G = 4, 3
pods = [(0, "p-1", "ns-b", "c1", [10, 11]), (0, "p-1", "ns-a", "c1", [12]), (0, "p-10", "ns-a", "c1", [13, 13]), (0, "p", "ns-a", "c0", [14]),
        (0, 'q"x', "n\\s", "c\n2", [15]), (0, "p-1", "ns-a", "c", [10]), (2, "p-1", "ns-a", "c1", [10]), (2, "solo", "ns", "c9", []),
        (3, "é", "éé", "é", [1, 2, 3])]
uu = ["GPU-b", "GPU-a", 'GPU-"c', "GPU-0", "GPU-00", "GPU-1", "z", "y", "x", "GPU-\\", "GPU-\n", "GPU-"]
def xtable(pods):
    ent_pk, ent_pid = [], []
    for i, p in enumerate(pods):
        ent_pk += [i] * len(p[4]); ent_pid += list(p[4])
    return Table(N, G, [p[0] for p in pods], [p[1] for p in pods], ent_pk, ent_pid, namespace=[p[2] for p in pods],
                 container=[p[3] for p in pods], gpu_uuid=uu)
def xcols(rows):
    return Cols(pid=[r[0] for r in rows], gpu=[r[1] for r in rows], used=[r[2] for r in rows], node=[r[3] for r in rows],
                sm_util=[r[4] for r in rows], mem_util=[r[5] for r in rows], ts_us=[r[6] for r in rows],
                gpu_total=np.full(N * G, 192 * GIB, dtype=np.uint64))
g, o = kgx.Kgx(MODE_X), kgo.Oracle(MODE_X, kgo.JOIN_HASH, 2)
tx = xtable(pods)
g.set_table(tx); o.set_table(tx)
rows = [(10, 0, 1 * GIB, 0, 50, 40, 1000), (10, 1, 2 * GIB, 0, 60, 30, 2000), (11, 2, 3, 0, 1, 2, 3000), (12, 0, 4, 0, 3, 4, 4000),
        (13, 1, 5, 0, 5, 6, 5000), (14, 2, 6, 0, 7, 8, 6000), (15, 0, 7, 0, 9, 10, 7000), (10, 0, 8, 2, 11, 12, 8000),
        (1, 1, 9, 3, 13, 14, 9000), (2, 2, 10, 3, 15, 16, 10000), (3, 0, 11, 3, 17, 18, 11000), (99, 0, 12, 0, 0, 0, 1), (10, 0, 13, 1, 0, 0, 2)]
c = xcols(rows)
g.ingest(c); o.ingest(c)
check(g, o, "x t1")
g.ingest(c); o.ingest(c)
check(g, o, "x t1 twice")
tx2 = xtable(pods[2:] + [(1, "late", "ns", "c1", [10])])
g.set_table(tx2); o.set_table(tx2)
g.ingest(c); o.ingest(c)
check(g, o, "x t2")
before = g.render()
rejected(g, xtable(pods + [(0, "p-1", "ns-b", "c1", [1])]), "x dup key")
uu_ok = list(uu); uu[4] = "GPU-0"
rejected(g, xtable(pods), "x dup uuid")
uu[:] = uu_ok; uu[5] = b"\xc0\x80"
rejected(g, xtable(pods), "x uuid utf8")
uu[:] = uu_ok
rejected(g, xtable([(0, "p", b"\xe2\x82", "c", [1])]), "x ns utf8")
assert g.render() == before
g.ingest(c); o.ingest(c)
check(g, o, "x after rejects")
print("digest", digest.hexdigest())
""" % (ROOT, os.path.join(ROOT, "tests"))


def _run(env):
    e = dict(os.environ)
    e.pop("KGX_HOST_TABLE_BUILD", None)
    e.update(env)
    r = subprocess.run([sys.executable, "-c", SCRIPT], env=e, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "digest" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
    # KGX_TRACE names the builder on stderr: make sure each run used the one it was meant to
    assert ("device build" in r.stderr) == ("KGX_HOST_TABLE_BUILD" not in env), r.stderr[-2000:]
    return r.stdout


def test_device_and_host_builders_agree_with_the_oracle_and_each_other():
    dev = _run({"KGX_TRACE": "1"})
    host = _run({"KGX_TRACE": "1", "KGX_HOST_TABLE_BUILD": "1"})
    assert dev == host, "\n".join(a + "   |   " + b for a, b in zip(dev.splitlines(), host.splitlines()) if a != b)
