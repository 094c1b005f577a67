"""kgx_render_noalloc: the exposition writer for callers that cannot allocate
GPU memory at scrape time — a /metrics handler sharing its GPU with a job that
already owns all of the device memory, or one whose latency budget rules out
the device-wide stall of cudaMalloc/cudaFree.  Its scratch is a fixed array
inside the library and the text is written into a device buffer the caller
brings along.

Checked here, through the C ABI, against the ordinary kgx_render on the same
context: byte-identical bodies on two calls in a row (the second after the
state changed, so leftovers of the first call in the shared scratch array
would show), plus the documented size limits.  Collected last on purpose."""
import ctypes as C

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
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
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
        rc, n = _noalloc(g, buf)
        assert rc == 0, kgx.lib().kgx_last_error(g._h)
        got = buf[:n].cpu().numpy().tobytes()
        assert n == len(want) and got == want         # bytes: exact
        assert g.render() == want                     # and the ordinary path is undisturbed


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
