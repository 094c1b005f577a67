"""The C++ host mirror of the reference interface (Exporter::Cycle, Describe /
Collect, ServeHTTP) compiles against the C ABI; on a GPU box it is run against
the golden body."""
import os
import subprocess

import pytest

from kgx_b200 import build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "host_mirror_test.cpp")


def _compile(tmp_path):
    so = build.build()
    exe = str(tmp_path / "host_mirror_test")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", SRC, "-o", exe, "-L" + os.path.dirname(so), "-lkgx",
                           "-Wl,-rpath," + os.path.dirname(so), "-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64"])
    return exe


def test_host_mirror_compiles_and_links(tmp_path):
    assert os.path.exists(_compile(tmp_path))


@pytest.mark.gpu
def test_host_mirror_serves_the_golden_body(tmp_path):
    exe = _compile(tmp_path)
    r = subprocess.run([exe, os.path.join(ROOT, "tests", "golden", "r_single.txt")], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == "ok", r.stdout + r.stderr


@pytest.mark.gpu
def test_host_mirror_collects_a_million_series_in_bulk(tmp_path):
    """prometheus.Collector.Collect over 1 M series (BASELINE config 3's size): the read-back through the C ABI (labels
    in one call, values one call per family) stays under 100 ms; values and label values are checked by recomputation."""
    exe = _compile(tmp_path)
    r = subprocess.run([exe, "big"], capture_output=True, text=True, timeout=300)
    print(r.stdout)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout + r.stderr


@pytest.mark.gpu
def test_host_mirror_small_config_latency(tmp_path):
    """BASELINE config 1's size through the C ABI from C++ (no interpreter in the timed loop): prints the cost of one
    poll + scrape when the step is launch-bound."""
    exe = _compile(tmp_path)
    r = subprocess.run([exe, "small"], capture_output=True, text=True, timeout=120)
    print(r.stdout)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout + r.stderr
