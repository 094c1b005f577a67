"""Host-side pieces of bench.py that run without a GPU: the stale-capture guard on `roofline.traffic`, the workload
description of the reference arm, and the checksum that certifies a gathered body."""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
bench = importlib.import_module("bench")


def test_traffic_is_withheld_when_the_capture_is_of_another_kernel_version():
    import csv
    rows = list(csv.reader(open(os.path.join(ROOT, "profiles", bench.TRAFFIC_CSV))))
    hdr, units, vals = rows[0], rows[1], rows[-1]
    i = hdr.index("gpu__time_duration.sum")
    cap_ms = float(vals[i]) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0}[units[i]]
    traffic, src = bench.ncu_traffic(bench.TRAFFIC_CSV, cap_ms * 1.05)   # this run's kernel agrees with the capture
    assert traffic and 3e9 < traffic < 5e9 and "profiles/" in src        # 3.34 GB of text + label reads that miss L2
    traffic, src = bench.ncu_traffic(bench.TRAFFIC_CSV, cap_ms * 1.25)   # the kernel changed since the capture
    assert traffic is None and "stale" in src
    traffic, src = bench.ncu_traffic("no_such_capture.csv", 1.0)
    assert traffic is None and "no capture" in src


def test_workload_description_names_config_and_sharding():
    c = bench.workload_config(2000)
    assert "configs[3]" in c["workload"] and c["samples_per_gpu"] == 10_000_000 and "by node" in c["sharding"]


def test_body_checksum_detects_shift_truncation_and_flip():
    from kgx_b200 import fleetshard
    t = torch.from_numpy(np.random.default_rng(7).integers(0, 256, 300_001, dtype=np.uint8))
    c = fleetshard.body_checksum(t)
    assert c == fleetshard.body_checksum(t, chunk=4099)              # independent of the chunking
    assert c != fleetshard.body_checksum(torch.roll(t, 1))
    assert c != fleetshard.body_checksum(t[:-1])
    u = t.clone(); u[123_456] ^= 4
    assert c != fleetshard.body_checksum(u)
    offs = [0, 100_000, 300_001]
    sums = [fleetshard.body_checksum(t[:100_000]), fleetshard.body_checksum(t[100_000:])]
    assert fleetshard.verify_gather(t, offs, sums) and not fleetshard.verify_gather(u, offs, sums)
