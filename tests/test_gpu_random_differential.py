"""Randomised differential test: small random clusters (random pod / namespace /
container / uuid strings with quotes, backslashes, newlines and UTF-8; PIDs
shared between pods; duplicate list entries; empty pods; GPUs out of range;
zero totals; all-ones used bytes) driven through random sequences of
set_table / ingest / reset_window / render on the GPU and on the oracle."""
import numpy as np
import pytest

from kgx_b200 import build, kgx
from kgx_b200.abi import Cols, Table, MODE_R_INTENDED, MODE_R_LITERAL, MODE_X

pytestmark = pytest.mark.gpu
ALPHABET = list("abcdefghijklmnopqrstuvwxyz0123456789-._") + ['"', "\\", "\n", " ", "é", "✓", "{", "}", "=", ","]


def _name(rng, lo=0, hi=24):
    return "".join(rng.choice(ALPHABET, size=int(rng.integers(lo, hi))))


def _table(rng, mode, n_nodes, G):
    pods = []
    for n in range(n_nodes):
        seen = set()
        for _ in range(int(rng.integers(0, 7))):
            key = (_name(rng, 0, 20), _name(rng, 1, 8), _name(rng, 0, 6))
            if mode != MODE_X:
                key = (key[0], "", "")
            if key in seen:
                continue
            seen.add(key)
            pids = list(rng.integers(0, 40, size=int(rng.integers(0, 9))))
            pods.append((n, key, pids))
    ent_pk, ent_pid = [], []
    for i, (_, _, pids) in enumerate(pods):
        ent_pk += [i] * len(pids)
        ent_pid += pids
    kw = {}
    if mode == MODE_X:
        kw = dict(namespace=[p[1][1] for p in pods], container=[p[1][2] for p in pods],
                  gpu_uuid=["GPU-%d-%s" % (i, _name(rng, 0, 5)) for i in range(n_nodes * G)])
    return Table(n_nodes, G, [p[0] for p in pods], [p[1][0] for p in pods], ent_pk, ent_pid, **kw)


def _cols(rng, n_nodes, G, grouped):
    n = int(rng.integers(0, 120))
    node = np.sort(rng.integers(0, n_nodes, size=n)).astype(np.uint32) if grouped else rng.integers(0, n_nodes + 1, size=n).astype(np.uint32)
    used = rng.integers(0, 1 << 40, size=n).astype(np.uint64)
    if n:
        used[rng.integers(0, n)] = np.uint64(2**64 - 1)  # NVML_VALUE_NOT_AVAILABLE
    total = rng.integers(1, 1 << 41, size=n_nodes * G).astype(np.uint64)
    total[rng.integers(0, n_nodes * G)] = 0
    kw = dict(pid=rng.integers(0, 45, size=n), gpu=rng.integers(0, G + 1, size=n), used=used, node=node,
              sm_util=rng.integers(0, 101, size=n), mem_util=rng.integers(0, 101, size=n),
              ts_us=rng.integers(0, 1 << 50, size=n).astype(np.uint64), gpu_total=total)
    if grouped:
        off = np.searchsorted(node, np.arange(n_nodes + 1)).astype(np.uint64)
        kw["node_off"] = off
    return Cols(**kw)


@pytest.mark.parametrize("mode", [MODE_R_INTENDED, MODE_R_LITERAL, MODE_X])
def test_random_sequences_match_the_oracle(kgo, mode):
    build.build()
    rng = np.random.default_rng(0xB200 + mode)
    for scenario in range(150):
        n_nodes, G = int(rng.integers(1, 5)), int(rng.integers(1, 5))
        g, o = kgx.Kgx(mode), kgo.Oracle(mode, kgo.JOIN_LOOPS if scenario % 2 else kgo.JOIN_HASH)
        t = _table(rng, mode, n_nodes, G)
        g.set_table(t); o.set_table(t)
        for step in range(int(rng.integers(1, 7))):
            op = rng.integers(0, 10)
            if op == 0:
                t = _table(rng, mode, n_nodes, G)      # the pod->PID map is rebuilt (main.go:84, every cycle)
                g.set_table(t); o.set_table(t)
            elif op == 1 and mode == MODE_X:
                g.reset_window(); o.reset_window()
            elif op == 2 and mode == MODE_X:
                cut = int(rng.integers(0, 1 << 50))   # explicit expiry: series whose newest sample is older go
                assert g.expire(cut) == o.expire(cut)
            else:
                c = _cols(rng, n_nodes, G, grouped=bool(rng.integers(0, 2)))
                g.ingest(c); o.ingest(c)
            gb, ob = g.render(), o.render()
            assert gb == ob, "scenario %d step %d:\n%r\n%r" % (scenario, step, gb[:400], ob[:400])
            assert list(g.body_offsets()) == list(o.body_offsets())
        for fam in range(g.family_count()):
            a, b = g.read_values(fam), o.read_values(fam)
            assert a.tobytes() == b.tobytes()   # NaN-safe bit comparison
        g.close()
