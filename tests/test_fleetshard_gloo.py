"""The N>1 host logic on CPU: world_size-2 (and 3) gloo groups.  Each rank
"renders" its node shard with the CPU oracle; the ordered concatenation on
rank 0 must equal the oracle's body of the whole fleet."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, q):
    sys.path.insert(0, ROOT)
    import kgx_loader
    kgx_loader.load()
    kgo = kgx_loader.load_oracle()
    from kgx_b200 import fleetshard
    from kgx_b200.fleet import Fleet
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n_nodes = 7
        lo, hi = fleetshard.node_range(rank, world, n_nodes)
        # the shard is a fleet of its own whose node ids start at `lo`
        f = Fleet(n_nodes=hi - lo, pods=20, procs=150, seed=0xB200 + 3, node_base=lo)
        o = kgo.Oracle(mode)
        o.set_table(f.table(mode))
        o.ingest(f.scrape(0))
        body = o.render()
        local = torch.frombuffer(bytearray(body), dtype=torch.uint8) if body else torch.empty(0, dtype=torch.uint8)
        out, off = fleetshard.gather_bodies(local, dst=0)
        slowest = fleetshard.max_over_ranks(float(rank + 1))
        if rank == 0:
            q.put((out.numpy().tobytes(), off.tolist(), slowest))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_ordered_gather_equals_whole_fleet(kgo, world):
    from kgx_b200.abi import MODE_X
    from kgx_b200.fleet import Fleet
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, MODE_X, q)) for r in range(world)]
    for p in procs:
        p.start()
    got, off, slowest = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    whole = Fleet(n_nodes=7, pods=20, procs=150, seed=0xB200 + 3)
    o = kgo.Oracle(MODE_X)
    o.set_table(whole.table(MODE_X))
    o.ingest(whole.scrape(0))
    want = o.render()
    assert got == want
    assert off[0] == 0 and off[-1] == len(want) and len(off) == world + 1
    assert slowest == float(world)


def test_node_range_partitions():
    from kgx_b200.fleetshard import node_range
    for n in (1, 7, 128, 2000):
        for w in (1, 2, 3, 8):
            parts = [node_range(r, w, n) for r in range(w)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1


def test_nccl_gather_env_sets_defaults_and_respects_the_environment(monkeypatch):
    from kgx_b200 import fleetshard
    for k in ("NCCL_MIN_P2P_NCHANNELS", "NCCL_MAX_P2P_NCHANNELS", "NCCL_MAX_NCHANNELS"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("NCCL_MAX_NCHANNELS", "16")      # the operator's choice wins
    fleetshard.nccl_gather_env()
    assert os.environ["NCCL_MIN_P2P_NCHANNELS"] == "64" and os.environ["NCCL_MAX_P2P_NCHANNELS"] == "64"
    assert os.environ["NCCL_MAX_NCHANNELS"] == "16"
