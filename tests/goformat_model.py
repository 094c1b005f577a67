"""Independent model of expfmt's writeFloat built on CPython's repr(), which is
a shortest-round-trip printer of its own (David Gay's algorithm).  Used to pin
the oracle's digit generation; the Go %g layout rules are
applied on top of repr's digits."""
import math
from decimal import Decimal


def go_format(v):
    if v == 1:
        return "1"
    if v == 0:
        return "0"
    if v == -1:
        return "-1"
    if math.isnan(v):
        return "NaN"
    if math.isinf(v):
        return "+Inf" if v > 0 else "-Inf"
    sign = "-" if v < 0 else ""
    t = Decimal(repr(abs(v))).as_tuple()
    digits = "".join(map(str, t.digits)).lstrip("0")
    exp = t.exponent
    stripped = digits.rstrip("0")
    exp += len(digits) - len(stripped)
    digits = stripped
    n = len(digits)
    dp = n + exp
    x = dp - 1
    if x < -4 or x >= 6:
        s = digits[0] + ("." + digits[1:] if n > 1 else "")
        s += "e" + ("-" if x < 0 else "+") + ("%02d" % abs(x))
        return sign + s
    if dp <= 0:
        return sign + "0." + "0" * (-dp) + digits
    if dp >= n:
        return sign + digits + "0" * (dp - n)
    return sign + digits[:dp] + "." + digits[dp:]
