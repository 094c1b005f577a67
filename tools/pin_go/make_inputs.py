#!/usr/bin/env python3
"""Writes tests/golden/go_inputs.json: the fixed inputs tools/pin_go/main.go feeds to the reference's own encoder
stack.  Deterministic (no clock, seeded).  Each case lists the GaugeVec.Set calls Go makes (label values plus the two
float64 values as hex bit patterns, so nothing is lost in JSON) and, for the cases whose values come out of a join, the
raw table and sample rows the parity test replays through the oracle and the GPU path.

Cases: the SURVEY 8c whole-body golden; SURVEY's known-answer values; label escaping; decimal-string series order
("10" < "9"); 1 000 random doubles; a small fleet (BASELINE config 1) joined by PID."""
import json
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import kgx_loader  # noqa: E402

kgx_loader.load()
from kgx_b200.fleet import Fleet, GIB  # noqa: E402
from kgx_b200.abi import MODE_R_INTENDED  # noqa: E402
from test_oracle_format import KAT  # noqa: E402


def bits(v):
    return "%016x" % struct.unpack("<Q", struct.pack("<d", float(v)))[0]


def joined(pods, rows, total):
    """main.go:141-151 in the intended reading: every (process, pod whose PID list holds its PID) -> one Set each,
    in row order (last write wins, which Go's repeated Set does by itself)."""
    out = []
    for pid, gpu, used in rows:
        for _, name, pids in pods:
            if pid in pids:
                out.append({"Pid": str(pid), "Pod": name, "Used": bits(float(used)),
                            "Perc": bits((float(used) / float(total[gpu])) * 100)})
    return out


def replay(pods, rows, total):
    return {"pods": [[n, name, list(map(int, pids))] for n, name, pids in pods], "rows": [list(map(int, r)) for r in rows],
            "total": list(map(int, total))}


def main():
    cases = []
    total = [80 * GIB] * 8
    pods = [(0, "trainer-0", [1, 4242, 77]), (0, "idle", [5, 6])]
    rows = [(4242, 0, 1 * GIB), (999, 1, 5 * GIB)]
    cases.append({"Name": "r_single", "Samples": joined(pods, rows, total), "Replay": replay(pods, rows, total)})
    cases.append({"Name": "kat_values", "Samples": [{"Pid": str(i), "Pod": "kat", "Used": bits(v), "Perc": bits(v)}
                                                    for i, (v, _) in enumerate(KAT)], "Replay": None})
    pods = [(0, "back\\slash", [11]), (0, 'quo"te', [12]), (0, "new\nline", [13]), (0, "utf8-é世", [14]),
            (0, "plain", [15, 11])]
    rows = [(11, 0, 3 * GIB), (12, 1, 2**64 - 1), (13, 2, 0), (14, 3, 80 * GIB), (15, 4, 12345)]
    cases.append({"Name": "escaping", "Samples": joined(pods, rows, total), "Replay": replay(pods, rows, total)})
    pods = [(0, "b", [9, 10, 100, 1, 2, 19]), (0, "a", [10, 9, 1000000, 99999])]
    rows = [(p, i % 8, (i + 1) * 1000003) for i, p in enumerate([9, 10, 100, 1, 2, 19, 1000000, 99999])]
    cases.append({"Name": "ordering", "Samples": joined(pods, rows, total), "Replay": replay(pods, rows, total)})
    rng = np.random.default_rng(0xB200)
    rnd = rng.integers(0, 2**64, size=1000, dtype=np.uint64)
    cases.append({"Name": "random_values", "Samples": [{"Pid": str(i), "Pod": "rnd", "Used": "%016x" % int(b), "Perc": "%016x" % int(b ^ np.uint64(1 << 63))}
                                                       for i, b in enumerate(rnd)], "Replay": None})
    f = Fleet.config(1)
    t, s = f.table(MODE_R_INTENDED), f.scrape(0)
    blob, off = t.pod
    names = [bytes(blob[int(off[i]):int(off[i + 1])]).decode() for i in range(len(off) - 1)]
    pmap = {}
    for pk, pid in zip(t.entry_podkey, t.entry_pid):
        pmap.setdefault(int(pk), []).append(int(pid))
    pods = [(0, names[k], v) for k, v in sorted(pmap.items())]
    rows = [(int(p), int(g), int(u)) for p, g, u in zip(s.pid, s.gpu, s.used)]
    total = [int(x) for x in s.gpu_total]
    cases.append({"Name": "fleet_cfg1", "Samples": joined(pods, rows, total), "Replay": replay(pods, rows, total)})
    path = os.path.join(ROOT, "tests", "golden", "go_inputs.json")
    with open(path, "w") as fh:
        fh.write("[\n" + ",\n".join(json.dumps(c, separators=(",", ":"), ensure_ascii=True) for c in cases) + "\n]\n")
    print("wrote %s: %s" % (path, ", ".join("%s (%d sets)" % (c["Name"], len(c["Samples"])) for c in cases)))


if __name__ == "__main__":
    main()
