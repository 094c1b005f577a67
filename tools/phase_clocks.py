#!/usr/bin/env python3
"""Where does a writer tile spend its time?  Builds the library with -DKGX_PHASE_CLOCKS (thread 0 of every writer CTA
accumulates clock64() deltas per phase of the tile loop), runs the bench workload, and prints cycles per tile and phase.

  python tools/phase_clocks.py build      (here, no GPU needed)
  python tools/phase_clocks.py run        (on the GPU box)
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "tools", "exp", "libkgx_phase.so")
PHASES = ["-", "measure next tile", "-", "look-back", "header+label wait", "stream line (+descriptor)", "end barrier",
          "label copy issue", "-"]

if __name__ == "__main__":
    sys.path.insert(0, ROOT)
    if sys.argv[1] == "build":
        import kgx_loader
        kgx_loader.load()
        from kgx_b200 import build
        os.makedirs(os.path.dirname(SO), exist_ok=True)
        build.build_variant(SO, ["-DKGX_PHASE_CLOCKS"])
        print("built", SO)
        sys.exit(0)
    os.environ.setdefault("KGX_LIB", SO)
    import kgx_loader
    kgx_loader.load()
    import numpy as np
    import torch
    from kgx_b200 import abi, kgx
    from kgx_b200.abi import MODE_X
    from kgx_b200.fleet import Fleet
    nodes = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
    f = Fleet.config(4, n_nodes=nodes)
    g = kgx.Kgx(MODE_X, profile=True)
    g.set_table(f.table(MODE_X))
    s = f.scrape(0)
    for _ in range(3):
        g.reset_window()
        g.ingest(s)
        g.render_device()
    fn = kgx.lib().kgx_debug_phase_clocks
    fn.argtypes = [C.c_void_p, C.c_int]
    out = (C.c_uint64 * 16)()
    fn(None, 1)
    g.profile_reset()
    K = 5
    for _ in range(K):
        g.render_device()
    fn(out, 0)
    prof = g.profile()
    tiles = out[10]
    print("render %.4f ms/launch, %d tiles/launch" % (prof["render"][0] / K, tiles // K))
    tot = sum(out[i] for i in range(8))
    for i, name in enumerate(PHASES[:8]):
        print("%-20s %8.0f cycles/tile  %5.1f %%" % (name, out[i] / tiles, 100.0 * out[i] / tot))
    print("%-20s %8.0f cycles/tile" % ("total", tot / tiles))
    print("look-back: %.2f steps of 32 predecessors per tile, %.2f re-polls of an unpublished word per tile (thread 0)" % (out[8] / tiles, out[9] / tiles))
