"""Pins the oracle's value/label encoding: SURVEY.md 8c known-answer vectors,
CPython repr() as an independent shortest-digits implementation, and an exact
big-integer model of the power-of-ten table."""
import math
import struct

import numpy as np

from goformat_model import go_format

# SURVEY.md section 8c "Known-answer vectors"
KAT = [
    (0.0, "0"), (1.0, "1"), (100.0, "100"), (12.5, "12.5"), (99999.0, "99999"), (100000.0, "100000"),
    (999999.0, "999999"), (1000000.0, "1e+06"), (1048576.0, "1.048576e+06"), (1073741824.0, "1.073741824e+09"),
    (85899345920.0, "8.589934592e+10"), (205621559296.0, "2.05621559296e+11"),
    (float(2**64 - 1), "1.8446744073709552e+19"), (0.0001, "0.0001"), (0.00001, "1e-05"),
    ((1.0 / 3.0) * 100, "33.33333333333333"),  # SURVEY writes "100/3": only (1/3)*100, main.go:149's order, gives this
    (100 / 3, "33.333333333333336"), (536870912 / 85899345920 * 100, "0.625"), (123456.789, "123456.789"),
    (1234567.891, "1.234567891e+06"), (1e21, "1e+21"),
    ((2147483648.0 / 206158430208.0) * 100, "1.0416666666666665"),
    (-1.0, "-1"), (-0.0, "0"), (float("nan"), "NaN"), (float("inf"), "+Inf"), (float("-inf"), "-Inf"),
    (5e-324, "5e-324"), (1.7976931348623157e308, "1.7976931348623157e+308"), (-2.5, "-2.5"),
    (1e100, "1e+100"), (1.5e-7, "1.5e-07"), (123456.0, "123456"), (0.3, "0.3"), (2.0**53, "9.007199254740992e+15"),
]


def test_known_answers(kgo):
    for v, want in KAT:
        assert kgo.format_value(v) == want, (v, kgo.format_value(v), want)
        if not math.isnan(v):
            assert go_format(v) == want


def test_operation_order_vector():
    # main.go:149 divides first, then multiplies
    used, total = 2 * 2**30, 192 * 2**30
    assert (float(used) / float(total)) * 100 == 1.0416666666666665
    assert float(used) * 100 / float(total) != (float(used) / float(total)) * 100


def test_pow10_table_exact(kgo):
    for k in range(-292, 325):
        if k >= 0:
            n = 10**k
            bits = n.bit_length()
            want = n << (128 - bits) if bits <= 128 else -((-n) >> (bits - 128))
        else:
            d = 10**(-k)
            s = d.bit_length() + 127
            want = (1 << s) // d + 1
        assert want.bit_length() == 128
        assert kgo.pow10_entry(k) == want, k


def _check(kgo, vals):
    got = kgo.format_values(vals)
    for v, rec in zip(vals, got):
        s = bytes(rec[:rec[31]]).decode()
        assert s == go_format(float(v)), (float(v).hex(), s, go_format(float(v)))


def test_random_bit_patterns_vs_repr(kgo):
    rng = np.random.default_rng(0xB200)
    bits = rng.integers(0, 2**64, size=300_000, dtype=np.uint64)
    vals = bits.view(np.float64)
    vals = vals[np.isfinite(vals)]
    _check(kgo, vals)


def test_domain_values_vs_repr(kgo):
    rng = np.random.default_rng(7)
    used = rng.integers(1, 65537, size=100_000).astype(np.uint64) * np.uint64(2 << 20)
    total = np.uint64(192 << 30)
    _check(kgo, used.astype(np.float64))
    _check(kgo, used.astype(np.float64) / np.float64(total) * 100)
    cnt = rng.integers(1, 50, size=100_000)
    sm = np.array([rng.integers(0, 100 * c + 1) for c in cnt[:50_000]], dtype=np.float64)
    _check(kgo, sm / cnt[:50_000])
    ts = rng.integers(1_790_000_000_000_000, 1_790_000_000_000_000 + 10**10, size=50_000).astype(np.float64)
    _check(kgo, ts / 1e6)


def test_boundaries_vs_repr(kgo):
    vals = []
    for e in range(-1074, 1024, 7):
        vals += [math.ldexp(1.0, e), math.ldexp(1.0, e) * (1 + 2**-52), math.ldexp(1.0, e) * (2 - 2**-52)]
    for k in range(-320, 309, 3):
        for m in (1.0, 9.999999999999999, 5.0):
            try:
                v = float("%re%d" % (m, k))
            except OverflowError:
                continue
            if v != 0 and math.isfinite(v):
                vals.append(v)
                b = struct.unpack("<Q", struct.pack("<d", v))[0]
                vals.append(struct.unpack("<d", struct.pack("<Q", b + 1))[0])
                vals.append(struct.unpack("<d", struct.pack("<Q", b - 1))[0])
    vals += [float(i) for i in range(2, 3000)] + [i / 8 for i in range(1, 2000)] + [10.0**i for i in range(-10, 23)]
    _check(kgo, np.array([v for v in vals if math.isfinite(v)], dtype=np.float64))


def test_escape(kgo):
    assert kgo.escape('a\\b"c\nd') == b'a\\\\b\\"c\\nd'
    assert kgo.escape('a\\b"c\nd', is_label=False) == b'a\\\\b"c\\nd'
    assert kgo.escape("héllo ✓") == "héllo ✓".encode()
