"""Pins the oracle's join / upsert / gather semantics against the hand-derived
goldens of SURVEY.md section 8c and against itself (nested loops of
main.go:141-145 vs hash join)."""
import os

import numpy as np
import pytest

from kgx_b200.abi import Cols, Table, MODE_R_INTENDED, MODE_R_LITERAL, MODE_X
from kgx_b200.fleet import Fleet, GIB

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _tbl(pods, n_nodes=1, G=8):
    """pods: list of (node, name, [pids])"""
    ent_pk, ent_pid = [], []
    for i, (_, _, pids) in enumerate(pods):
        ent_pk += [i] * len(pids)
        ent_pid += list(pids)
    return Table(n_nodes, G, [p[0] for p in pods], [p[1] for p in pods], ent_pk, ent_pid)


def _cols(rows, n_nodes=1, G=8, total=80 * GIB):
    """rows: list of (pid, gpu, used[, node])"""
    return Cols(pid=[r[0] for r in rows], gpu=[r[1] for r in rows], used=[r[2] for r in rows],
                node=[r[3] if len(r) > 3 else 0 for r in rows],
                gpu_total=np.full(n_nodes * G, total, dtype=np.uint64))


def test_whole_body_golden(kgo):
    for join in (kgo.JOIN_LOOPS, kgo.JOIN_HASH):
        o = kgo.Oracle(MODE_R_INTENDED, join)
        o.set_table(_tbl([(0, "trainer-0", [1, 4242, 77]), (0, "idle", [5, 6])]))
        o.ingest(_cols([(4242, 0, 1 * GIB), (999, 1, 5 * GIB)]))
        assert o.render() == open(os.path.join(GOLD, "r_single.txt"), "rb").read()


def test_empty_body(kgo):
    o = kgo.Oracle(MODE_R_INTENDED)
    o.set_table(_tbl([(0, "p", [10])]))
    o.ingest(_cols([(11, 0, 1)]))
    assert o.render() == b""
    o.ingest(_cols([]))
    assert o.render() == b""


def test_literal_golden(kgo):
    # main.go:144-145: `for pid := range pids` yields indices 0..len-1
    for join in (kgo.JOIN_LOOPS, kgo.JOIN_HASH):
        o = kgo.Oracle(MODE_R_LITERAL, join)
        o.set_table(_tbl([(0, "p", [1000, 2000, 3000])]))
        o.ingest(_cols([(0, 0, 10), (2, 0, 30), (5, 0, 60)]))
        body = o.render().decode()
        lines = [l for l in body.splitlines() if l.startswith("pod_gpu_memory_usage{")]
        assert lines == ['pod_gpu_memory_usage{pid="0",pod="p"} 10', 'pod_gpu_memory_usage{pid="2",pod="p"} 30']


def test_series_order_is_string_order(kgo):
    o = kgo.Oracle(MODE_R_INTENDED)
    o.set_table(_tbl([(0, "b", [9, 10, 100]), (0, "a", [9])]))
    o.ingest(_cols([(9, 0, 1), (10, 0, 2), (100, 0, 3)]))
    lines = [l for l in o.render().decode().splitlines() if l.startswith("pod_gpu_memory_usage{")]
    assert [l.split(" ")[0] for l in lines] == [
        'pod_gpu_memory_usage{pid="10",pod="b"}', 'pod_gpu_memory_usage{pid="100",pod="b"}',
        'pod_gpu_memory_usage{pid="9",pod="a"}', 'pod_gpu_memory_usage{pid="9",pod="b"}']


def test_last_write_wins_and_persistence(kgo):
    o = kgo.Oracle(MODE_R_INTENDED)
    o.set_table(_tbl([(0, "p", [7, 8])]))
    o.ingest(_cols([(7, 0, 100), (7, 1, 200), (8, 0, 5)]))  # same pid on two devices: later device wins
    assert b'pod_gpu_memory_usage{pid="7",pod="p"} 200\n' in o.render()
    # next cycle: pid 8 gone, table rebuilt without it: its child keeps its value (no Delete in main.go)
    o.set_table(_tbl([(0, "p", [7])]))
    o.ingest(_cols([(7, 0, 300)]))
    body = o.render()
    assert b'pod_gpu_memory_usage{pid="7",pod="p"} 300\n' in body
    assert b'pod_gpu_memory_usage{pid="8",pod="p"} 5\n' in body


def test_escaping_and_special_values(kgo):
    o = kgo.Oracle(MODE_R_INTENDED)
    o.set_table(_tbl([(0, 'we"ird\\pod\nx', [1]), (0, "zero-total", [2])]))
    c = _cols([(1, 0, 2**64 - 1), (2, 1, 0)])
    c.gpu_total[1] = 0
    o.ingest(c)
    body = o.render().decode()
    assert 'pod_gpu_memory_usage{pid="1",pod="we\\"ird\\\\pod\\nx"} 1.8446744073709552e+19\n' in body
    assert 'docker_gpu_memory_perc_usage{pid="2",pod="zero-total"} NaN\n' in body


@pytest.mark.parametrize("mode", [MODE_R_INTENDED, MODE_R_LITERAL, MODE_X])
def test_loops_equals_hash_on_fleet(kgo, mode):
    f = Fleet(n_nodes=2, pods=16, procs=64, seed=0xB200 + 1)
    t = f.table(mode)
    bodies = []
    for join in (kgo.JOIN_LOOPS, kgo.JOIN_HASH):
        o = kgo.Oracle(mode, join, threads=2 if join else 1)
        o.set_table(t)
        s = f.scrape(0)
        if mode == MODE_R_LITERAL:
            s.pid = (s.pid % 7).astype(np.uint32)  # small Pids so the literal predicate fires
        o.ingest(s)
        bodies.append((o.render(), list(o.body_offsets())))
    assert bodies[0] == bodies[1]
    assert len(bodies[0][0]) > 0


def test_x_mode_values(kgo):
    t = Table(1, 2, [0, 0], ["web", "web"], [0, 0, 1], [10, 11, 12], namespace=["prod", "prod"],
              container=["app", "side"], gpu_uuid=["GPU-b", "GPU-a"])
    o = kgo.Oracle(MODE_X)
    o.set_table(t)
    c = Cols(pid=[10, 11, 12, 99], gpu=[0, 0, 1, 0], used=[100, 300, 50, 1], sm_util=[10, 21, 7, 1],
             mem_util=[1, 2, 3, 4], ts_us=[1_000_000, 2_500_000, 3_000_000, 9], gpu_total=[1000, 2000])
    o.ingest(c)
    body = o.render().decode()
    assert 'pod_gpu_memory_usage{container="app",gpu="GPU-b",namespace="prod",pod="web"} 400\n' in body
    assert 'docker_gpu_memory_perc_usage{container="app",gpu="GPU-b",namespace="prod",pod="web"} 40\n' in body
    assert 'pod_gpu_memory_usage_max{container="app",gpu="GPU-b",namespace="prod",pod="web"} 300\n' in body
    assert 'pod_gpu_sm_utilization{container="app",gpu="GPU-b",namespace="prod",pod="web"} 15.5\n' in body
    assert 'pod_gpu_last_sample_timestamp_seconds{container="app",gpu="GPU-b",namespace="prod",pod="web"} 2.5\n' in body
    assert 'pod_gpu_memory_usage{container="side",gpu="GPU-a",namespace="prod",pod="web"} 50\n' in body
    fams = [l.split(" ")[2] for l in body.splitlines() if l.startswith("# TYPE")]
    assert fams == sorted(fams) and len(fams) == 6
    o.reset_window()
    assert o.render() == b""
