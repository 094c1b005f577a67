"""Parity against Go itself, when Go-made goldens are present.

tools/pin_go/ runs the reference's own encoder stack (GaugeVecs as at main.go:22-35, registry, promhttp.HandlerFor as
at main.go:69, at go.mod's module versions) on the fixed inputs of tests/golden/go_inputs.json and writes
tests/golden/go/<case>.txt.  No box this repository was built on has a Go toolchain, so those files are absent from
the tree and the tests below SKIP with "parity unpinned"; anyone with Go runs tools/pin_go/run.sh once and they turn
into real checks of the oracle (CPU) and of the CUDA path (-m gpu) against Go's bytes.

What runs regardless: the inputs file is in sync with its generator, and every replayable case goes through the oracle
(and the GPU) and yields exactly the series Go is told to Set — so the harness itself is known to work."""
import json
import os
import struct
import subprocess
import sys

import numpy as np
import pytest

from kgx_b200.abi import MODE_R_INTENDED
from test_oracle_semantics import _cols, _tbl

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GO_DIR = os.path.join(HERE, "golden", "go")
CASES = json.load(open(os.path.join(HERE, "golden", "go_inputs.json")))
FAMILIES = ("docker_gpu_memory_perc_usage", "pod_gpu_memory_usage")  # sorted: Gather's order (O1)
UNPINNED = "parity unpinned: no Go-made golden for this case (run tools/pin_go/run.sh on a box with Go)"


def _golden(name):
    p = os.path.join(GO_DIR, name + ".txt")
    return open(p, "rb").read() if os.path.exists(p) else None


def _f64(hexbits):
    return struct.unpack("<d", struct.pack("<Q", int(hexbits, 16)))[0]


def _replay(obj, case):
    r = case["Replay"]
    obj.set_table(_tbl([(n, name, pids) for n, name, pids in r["pods"]]))
    c = _cols([tuple(row) for row in r["rows"]])
    c.gpu_total = np.array(r["total"], dtype=np.uint64)
    obj.ingest(c)
    return obj.render()


def _sample_lines(body):
    """{(family, pid, raw label text): value text} of a body."""
    out = {}
    for line in body.split(b"\n"):
        if not line or line.startswith(b"#"):
            continue
        fam, rest = line.split(b"{", 1)
        labels, value = rest.rsplit(b"} ", 1)
        out[(fam.decode(), labels)] = value.decode()
    return out


def test_inputs_file_matches_its_generator(tmp_path):
    before = open(os.path.join(HERE, "golden", "go_inputs.json"), "rb").read()
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "pin_go", "make_inputs.py")], stdout=subprocess.DEVNULL)
    assert open(os.path.join(HERE, "golden", "go_inputs.json"), "rb").read() == before


@pytest.mark.parametrize("case", [c for c in CASES if c["Replay"]], ids=lambda c: c["Name"])
def test_oracle_replays_every_case(kgo, case):
    """The harness: replaying a case through the oracle yields one line per (family, series Go is told to Set), with the
    value text of exactly the doubles Go is handed; and, when the Go golden exists, the same bytes as Go."""
    body = _replay(kgo.Oracle(MODE_R_INTENDED, kgo.JOIN_LOOPS), case)
    lines = _sample_lines(body)
    last = {}
    for s in case["Samples"]:  # repeated Set: the last one wins
        last[(s["Pid"], s["Pod"])] = s
    assert len(lines) == 2 * len(last)
    for (pid, pod), s in last.items():
        lbl = b'pid="%s",pod="%s"' % (pid.encode(), kgo.escape(pod))
        assert lines[(FAMILIES[0], lbl)] == kgo.format_value(_f64(s["Perc"]))
        assert lines[(FAMILIES[1], lbl)] == kgo.format_value(_f64(s["Used"]))
    gold = _golden(case["Name"])
    if gold is None:
        pytest.skip(UNPINNED)
    assert body == gold


@pytest.mark.parametrize("case", [c for c in CASES if not c["Replay"]], ids=lambda c: c["Name"])
def test_oracle_value_text_vs_go(kgo, case):
    gold = _golden(case["Name"])
    if gold is None:
        pytest.skip(UNPINNED)
    lines = _sample_lines(gold)
    assert len(lines) == 2 * len(case["Samples"])
    for s in case["Samples"]:
        lbl = ('pid="%s",pod="%s"' % (s["Pid"], s["Pod"])).encode()
        assert kgo.format_value(_f64(s["Perc"])) == lines[(FAMILIES[0], lbl)], s
        assert kgo.format_value(_f64(s["Used"])) == lines[(FAMILIES[1], lbl)], s
    # series order inside a family: byte-wise by (pid, pod) label values (O3), families sorted (O1)
    order = [k for k in lines if k[0] == FAMILIES[1]]
    assert order == sorted(order, key=lambda k: k[1])
    assert gold.startswith(b"# HELP " + FAMILIES[0].encode())


@pytest.mark.gpu
@pytest.mark.parametrize("case", [c for c in CASES if c["Replay"]], ids=lambda c: c["Name"])
def test_gpu_replays_every_case(kgo, case):
    from kgx_b200 import build, kgx
    build.build()
    body = _replay(kgx.Kgx(MODE_R_INTENDED), case)
    assert body == _replay(kgo.Oracle(MODE_R_INTENDED, kgo.JOIN_LOOPS), case)
    gold = _golden(case["Name"])
    if gold is None:
        pytest.skip(UNPINNED)
    assert body == gold


@pytest.mark.gpu
@pytest.mark.parametrize("case", [c for c in CASES if not c["Replay"]], ids=lambda c: c["Name"])
def test_gpu_value_text_vs_go(case):
    from kgx_b200 import build, kgx
    build.build()
    gold = _golden(case["Name"])
    if gold is None:
        pytest.skip(UNPINNED)
    lines = _sample_lines(gold)
    g = kgx.Kgx(MODE_R_INTENDED)
    vals = np.array([_f64(s[k]) for s in case["Samples"] for k in ("Perc", "Used")], dtype=np.float64)
    out = g.format_values(vals)
    for i, s in enumerate(case["Samples"]):
        lbl = ('pid="%s",pod="%s"' % (s["Pid"], s["Pod"])).encode()
        for j, fam in enumerate(FAMILIES):
            rec = out[2 * i + j]
            assert bytes(rec[:rec[31]]).decode() == lines[(fam, lbl)], s
