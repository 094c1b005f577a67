#!/usr/bin/env python3
"""All five BASELINE.json configs on one B200, each checked for parity against
the CPU oracle and timed beside it:

  GPU    kgx_ingest + kgx_render through the C ABI, host buffers (e2e) and the
         device pipeline alone (per-kernel CUDA-event times)
  CPU    oracle-literal  the nested loops of main.go:141-154, 1 thread (cfg1, cfg2 only:
                         cfg3+ are O(10^12) compares)
         oracle-hash     same results via a hash join, 1 thread and all host threads
         ("C++ restatement of the Go path, not Go": no Go toolchain on the box)

Writes profiles/r01_configs.json and prints a markdown table."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kgx_loader  # noqa: E402

kgx_loader.load()
from kgx_b200 import kgx  # noqa: E402
from kgx_b200.abi import MODE_R_INTENDED, MODE_X  # noqa: E402
from kgx_b200.fleet import CONFIGS, Fleet  # noqa: E402

kgo = kgx_loader.load_oracle()
NPROC = os.cpu_count() or 1
CASES = [(1, MODE_R_INTENDED), (2, MODE_R_INTENDED), (3, MODE_R_INTENDED), (3, MODE_X), (4, MODE_X), (5, MODE_X)]
MODE_NAME = {MODE_R_INTENDED: "R-intended", MODE_X: "X"}


def cpu_time(mode, join, threads, table, scrapes, reps=1, copy=True):
    o = kgo.Oracle(mode, join, threads)
    o.set_table(table)
    best = None
    for _ in range(reps):
        if mode == MODE_X:
            o.reset_window()
        t0 = time.perf_counter()
        for s in scrapes:
            o.ingest(s)
        body = o.render(copy=copy)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return best, body


def pin_cols(s):
    from kgx_b200.abi import Cols
    kw = {}
    for k in ("pid", "gpu", "used", "node", "sm_util", "mem_util", "ts_us", "node_off", "gpu_total"):
        a = getattr(s, k)
        if a is not None:
            kw[k] = kgx.pinned_empty(len(a), a.dtype)
            kw[k][:] = a
    return Cols(**kw)


def main():
    rows = []
    for cfg, mode in CASES:
        f = Fleet.config(cfg)
        table = f.table(mode)
        scrapes = [f.scrape(k, x_cols=(mode == MODE_X)) for k in range(f.scrapes)]
        pinned = [pin_cols(s) for s in scrapes]  # e2e: pinned host buffers in, pinned host buffer out
        n_samples = f.n_nodes * f.procs * f.scrapes
        g = kgx.Kgx(mode, profile=True)
        g.set_table(table)

        out = [None]

        def gpu_pass():
            if mode == MODE_X:
                g.reset_window()
            for s in pinned:
                g.ingest(s)
            if out[0] is None:
                out[0] = kgx.pinned_empty(g.render_device() + 4096, np.uint8)
            n = g.render_into(out[0])
            return out[0][:n]

        body = gpu_pass().tobytes()  # includes the series-index build
        g.profile_reset()
        t0 = time.perf_counter()
        reps = 1 if cfg == 5 else 3
        for _ in range(reps):
            gpu_pass()
        e2e = (time.perf_counter() - t0) / reps
        prof = g.profile()
        dev_ms = sum(v[0] for v in prof.values()) / reps
        row = {"config": cfg, "mode": MODE_NAME[mode], "nodes": f.n_nodes, "samples": n_samples, "series": g.stats()["n_series"],
               "body_bytes": len(body), "gpu_device_ms": dev_ms, "gpu_e2e_ms": e2e * 1e3,
               "kernel_ms": {k: v[0] / reps for k, v in prof.items()}}
        # CPU legs + parity
        scr = scrapes
        if cfg in (1, 2):
            t_lit, b_lit = cpu_time(mode, kgo.JOIN_LOOPS, 1, table, scr)
            row["cpu_literal_1t_ms"] = t_lit * 1e3
            assert b_lit == body
        if cfg != 4:
            t1, b1 = cpu_time(mode, kgo.JOIN_HASH, 1, table, scr)
            row["cpu_hash_1t_ms"] = t1 * 1e3
            assert b1 == body, "parity broken on config %d" % cfg
            row["parity"] = "whole body byte-identical"
        else:
            row["parity"] = "see tests/test_gpu_full_configs.py (sampled nodes + invariants)"
        tn, bn = cpu_time(mode, kgo.JOIN_HASH, NPROC, table, scr, copy=(cfg != 4))
        row["cpu_hash_%dt_ms" % NPROC] = tn * 1e3
        assert (bn == len(body)) if cfg == 4 else (bn == body)
        row["speedup_e2e_vs_cpu_all_threads"] = tn / e2e
        row["speedup_device_vs_cpu_all_threads"] = tn * 1e3 / dev_ms
        rows.append(row)
        g.close()
        print(json.dumps(row), flush=True)
    out = os.path.join(ROOT, "profiles", "r01_configs.json")
    json.dump({"host_threads": NPROC, "rows": rows}, open(out, "w"), indent=1)
    print("\n| cfg | mode | samples | series | body MB | GPU device ms | GPU e2e ms | CPU literal 1t ms | CPU hash 1t ms | CPU hash %dt ms |" % NPROC)
    print("|---|---|---|---|---|---|---|---|---|---|")
    for r in rows:
        print("| %d | %s | %d | %d | %.1f | %.3f | %.2f | %s | %s | %.1f |" % (
            r["config"], r["mode"], r["samples"], r["series"], r["body_bytes"] / 1e6, r["gpu_device_ms"], r["gpu_e2e_ms"],
            "%.1f" % r["cpu_literal_1t_ms"] if "cpu_literal_1t_ms" in r else "—",
            "%.1f" % r["cpu_hash_1t_ms"] if "cpu_hash_1t_ms" in r else "—", r["cpu_hash_%dt_ms" % NPROC]))


if __name__ == "__main__":
    main()
