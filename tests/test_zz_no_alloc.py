"""kgx_render_noalloc: the exposition writer for callers that cannot allocate
GPU memory at scrape time — a /metrics handler sharing its GPU with a job that
already owns all of the device memory, or one whose latency budget rules out
the device-wide stall of cudaMalloc/cudaFree.  Its scratch is a fixed array
inside the library and the text is written into a device buffer the caller
brings along.

Checked here, through the C ABI, against the ordinary kgx_render on the same
context: byte-identical bodies on two calls in a row (the second after the
state changed, so leftovers of the first call in the shared scratch array
would show); that the second call really allocates nothing (the library's own
accounting and the driver's free-memory figure both stay put); two contexts
taking turns on the one shared array; the generic writer; and the documented
size limits.  Collected last on purpose."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from kgx_b200 import abi, build, kgx
from kgx_b200.abi import MODE_R_INTENDED, MODE_X
from kgx_b200.fleet import Fleet
from test_oracle_semantics import _cols, _tbl

pytestmark = pytest.mark.gpu


def _noalloc(g, buf):
    fn = kgx.lib().kgx_render_noalloc
    n = C.c_size_t()
    rc = fn(g._h, buf.data_ptr() if buf is not None else None, buf.numel() if buf is not None else 0, C.byref(n))
    return rc, n.value


@pytest.mark.parametrize("mode", [MODE_R_INTENDED, MODE_X])
def test_noalloc_matches_render_twice_in_a_row(mode):
    build.build()
    f = Fleet(n_nodes=5, pods=80, procs=700, seed=0xB200 + 9)
    g = kgx.Kgx(mode)
    g.set_table(f.table(mode))
    g.ingest(f.scrape(0))
    rc, need = _noalloc(g, None)                      # too small: reports the worst-case size
    assert rc == abi.KGX_E_NOSPACE and need > 0
    buf = torch.empty(need, dtype=torch.uint8, device="cuda")   # the caller's own device buffer
    for k in (0, 1):
        if k == 1:                                    # change the state between the two calls
            s = f.scrape(1)
            keep = (s.pid % 5) != 0
            s = abi.Cols(pid=s.pid[keep], gpu=s.gpu[keep], used=s.used[keep], node=s.node[keep], sm_util=s.sm_util[keep],
                         mem_util=s.mem_util[keep], ts_us=s.ts_us[keep], gpu_total=s.gpu_total)
            if mode == MODE_X:
                g.reset_window()
            g.ingest(s)
        want = g.render()                             # existing entry point
        buf.fill_(0x5a)
        torch.cuda.synchronize()
        held, free = g.stats()["device_bytes"], torch.cuda.mem_get_info()[0]
        rc, n = _noalloc(g, buf)
        assert rc == 0, kgx.lib().kgx_last_error(g._h)
        if k == 1:                                    # the promise itself: nothing was allocated (the first call on a
            assert g.stats()["device_bytes"] == held  # GPU may still make the driver load the module's data)
            assert torch.cuda.mem_get_info()[0] == free
        got = buf[:n].cpu().numpy().tobytes()
        assert n == len(want) and got == want         # bytes: exact
        assert g.render() == want                     # and the ordinary path is undisturbed


def test_noalloc_two_contexts_take_turns_on_the_shared_array():
    """Every ctx on a GPU uses the same scratch array.  Two of them, of different shapes, called one after the other
    (serialised, as the header asks): each call must clear what the other left behind."""
    build.build()
    fa = Fleet(n_nodes=7, pods=90, procs=900, seed=0xB200 + 31)
    fb = Fleet(n_nodes=3, pods=20, procs=150, seed=0xB200 + 32)
    ga, gb = kgx.Kgx(MODE_X), kgx.Kgx(MODE_R_INTENDED)
    ga.set_table(fa.table(MODE_X)); gb.set_table(fb.table(MODE_R_INTENDED))
    ga.ingest(fa.scrape(0)); gb.ingest(fb.scrape(0))
    wa, wb = ga.render(), gb.render()
    _, need_a = _noalloc(ga, None)
    _, need_b = _noalloc(gb, None)
    buf = torch.empty(max(need_a, need_b), dtype=torch.uint8, device="cuda")
    for g, want in ((ga, wa), (gb, wb), (gb, wb), (ga, wa), (gb, wb)):
        buf.fill_(0xa5)
        rc, n = _noalloc(g, buf)
        torch.cuda.synchronize()
        assert rc == 0 and buf[:n].cpu().numpy().tobytes() == want


def test_noalloc_generic_writer_path():
    """KGX_RENDER_V1 (read when the table is installed) selects the generic writer, which schemas with unusual metric
    names take; the allocation-free entry point must serve that path too."""
    build.build()
    f = Fleet(n_nodes=4, pods=30, procs=260, seed=0xB200 + 33)
    os.environ["KGX_RENDER_V1"] = "1"
    try:
        g = kgx.Kgx(MODE_X)
        g.set_table(f.table(MODE_X))
    finally:
        del os.environ["KGX_RENDER_V1"]
    g.ingest(f.scrape(0))
    want = g.render()
    _, need = _noalloc(g, None)
    buf = torch.empty(need, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        buf.fill_(0)
        rc, n = _noalloc(g, buf)
        assert rc == 0 and buf[:n].cpu().numpy().tobytes() == want


def test_noalloc_rejects_sizes_beyond_its_scratch():
    build.build()
    g = kgx.Kgx(MODE_R_INTENDED)
    n_nodes = 4097                                    # KGX_NOALLOC_MAX_NODES is 4096
    g.set_table(_tbl([(0, "p", [1])], n_nodes=n_nodes))
    g.ingest(_cols([(1, 0, 7)], n_nodes=n_nodes))
    buf = torch.empty(1 << 20, dtype=torch.uint8, device="cuda")
    rc, _ = _noalloc(g, buf)
    assert rc == abi.KGX_E_LIMIT
    assert b"KGX_NOALLOC_MAX_NODES" in kgx.lib().kgx_last_error(g._h)
    assert g.render().startswith(b"# HELP")           # the allocating entry point still serves it
    # within the node limit but more render tiles than KGX_NOALLOC_MAX_TILES (65 536): 4 000 nodes x 2 families x 11
    # tiles of up to 128 series
    f = Fleet(n_nodes=4000, pods=100, procs=1400, seed=0xB200 + 34)
    g = kgx.Kgx(MODE_R_INTENDED)
    g.set_table(f.table(MODE_R_INTENDED))
    g.ingest(f.scrape(0))
    rc, _ = _noalloc(g, buf)
    assert rc == abi.KGX_E_LIMIT
    assert b"KGX_NOALLOC_MAX_TILES" in kgx.lib().kgx_last_error(g._h)
    assert len(g.render()) > (1 << 20)
