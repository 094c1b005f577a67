"""The generic kernels that serve inputs the fast paths do not take — k_render
(metric names outside 15..64 bytes) and k_ingest (rows not grouped by node) —
stay byte-exact too.  They are forced through environment switches in a fresh
process, and the tile-size variants of the writer are swept the same way."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import sys
sys.path.insert(0, %r)
import kgx_loader
kgx_loader.load()
kgo = kgx_loader.load_oracle()
from kgx_b200 import kgx
from kgx_b200.abi import MODE_R_INTENDED, MODE_R_LITERAL, MODE_X
from kgx_b200.fleet import Fleet
import numpy as np
for mode in (MODE_R_INTENDED, MODE_R_LITERAL, MODE_X):
    f = Fleet(n_nodes=9, pods=70, procs=600, seed=0xB200 + 4, scrapes=2)
    t = f.table(mode)
    g, o = kgx.Kgx(mode), kgo.Oracle(mode, kgo.JOIN_HASH, 2)
    g.set_table(t); o.set_table(t)
    for k in range(2):
        s = f.scrape(k)
        if mode == MODE_R_LITERAL:
            s.pid = (s.pid %% 50).astype(np.uint32)
        g.ingest(s); o.ingest(s)
        assert g.render() == o.render(), (mode, k)
        assert list(g.body_offsets()) == list(o.body_offsets())
print("ok")
""" % ROOT


@pytest.mark.parametrize("env", [{"KGX_RENDER_V1": "1"}, {"KGX_INGEST_V1": "1"}, {"KGX_RENDER_THREADS": "32"},
                                 {"KGX_RENDER_THREADS": "128"}, {"KGX_RENDER_THREADS": "256"}, {"KGX_HOST_TABLE_BUILD": "1"}])
def test_forced_path_is_byte_exact(env):
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, "-c", SCRIPT], env=e, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout[-2000:] + r.stderr[-2000:]
