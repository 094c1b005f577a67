"""ctypes binding of the CPU ORACLE (oracle/libkgo.so).  TEST INFRASTRUCTURE:
imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  Never imported by the product package."""
import ctypes as C
import os
import subprocess

import numpy as np

import kgx_b200.abi as abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

JOIN_LOOPS, JOIN_HASH = 0, 1


def build(force=False):
    so = os.path.join(_HERE, "libkgo.so")
    srcs = [os.path.join(_HERE, f) for f in ("kgo.cpp", "kgo_fmt.cpp", "kgo.h")] + \
           [os.path.join(_HERE, "..", "include", "kgx.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libkgo.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "libkgo.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        L.kgo_create.restype = C.c_void_p
        L.kgo_create.argtypes = [C.c_int, C.c_int, C.c_int]
        L.kgo_destroy.argtypes = [C.c_void_p]
        L.kgo_set_table.argtypes = [C.c_void_p, C.POINTER(abi.kgx_table)]
        L.kgo_ingest.argtypes = [C.c_void_p, C.POINTER(abi.kgx_cols)]
        L.kgo_reset_window.argtypes = [C.c_void_p]
        L.kgo_expire.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.kgo_render.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
        L.kgo_body_offsets.argtypes = [C.c_void_p, abi.u64p, C.c_size_t]
        L.kgo_read_values.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.c_size_t, C.POINTER(C.c_size_t)]
        L.kgo_series_count.restype = C.c_uint64
        L.kgo_series_count.argtypes = [C.c_void_p]
        L.kgo_format_value.argtypes = [C.c_double, C.c_char_p]
        L.kgo_format_values.argtypes = [C.POINTER(C.c_double), C.c_size_t, C.c_void_p]
        L.kgo_pow10_entry.argtypes = [C.c_int, abi.u64p, abi.u64p]
        L.kgo_escape.restype = C.c_size_t
        L.kgo_escape.argtypes = [C.c_char_p, C.c_size_t, C.c_int, C.c_char_p, C.c_size_t]
        _LIB = L
    return _LIB


def format_value(v):
    buf = C.create_string_buffer(40)
    n = lib().kgo_format_value(float(v), buf)
    return buf.raw[:n].decode()


def format_values(values):
    v = np.ascontiguousarray(values, dtype=np.float64)
    out = np.zeros((len(v), 32), dtype=np.uint8)
    lib().kgo_format_values(v.ctypes.data_as(C.POINTER(C.c_double)), len(v), out.ctypes.data)
    return out


def pow10_entry(k):
    hi, lo = C.c_uint64(), C.c_uint64()
    lib().kgo_pow10_entry(k, C.byref(hi), C.byref(lo))
    return (hi.value << 64) | lo.value


def escape(s, is_label=True):
    b = s.encode() if isinstance(s, str) else s
    buf = C.create_string_buffer(2 * len(b) + 1)
    n = lib().kgo_escape(b, len(b), int(is_label), buf, len(buf))
    return buf.raw[:n]


class Oracle:
    def __init__(self, mode, join=JOIN_HASH, threads=1):
        self._h = lib().kgo_create(mode, join, threads)
        if not self._h:
            raise ValueError("kgo_create failed")
        self.mode = mode
        self._keep = None

    def close(self):
        if self._h:
            lib().kgo_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def _chk(self, rc):
        if rc != 0:
            raise RuntimeError("oracle error %d" % rc)

    def set_table(self, table):
        t = table.as_struct()
        self._chk(lib().kgo_set_table(self._h, C.byref(t)))
        self.n_nodes = table.n_nodes

    def ingest(self, cols):
        s = cols.as_struct()
        self._chk(lib().kgo_ingest(self._h, C.byref(s)))

    def reset_window(self):
        self._chk(lib().kgo_reset_window(self._h))

    def expire(self, older_than_ts_us):
        n = C.c_uint64()
        self._chk(lib().kgo_expire(self._h, int(older_than_ts_us), C.byref(n)))
        return n.value

    def render(self, copy=True):
        """The body as bytes; with copy=False only its length (bodies beyond 2 GB)."""
        p, n = C.c_void_p(), C.c_size_t()
        self._chk(lib().kgo_render(self._h, C.byref(p), C.byref(n)))
        if not copy:
            return n.value
        if n.value >= 1 << 31:
            return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value,)).tobytes()
        return C.string_at(p.value, n.value) if n.value else b""

    def render_view(self):
        """The body as a uint8 array over the oracle's own buffer (no copy; valid until the next render)."""
        p, n = C.c_void_p(), C.c_size_t()
        self._chk(lib().kgo_render(self._h, C.byref(p), C.byref(n)))
        if not n.value:
            return np.empty(0, dtype=np.uint8)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value,))

    def body_offsets(self):
        off = np.zeros(self.n_nodes + 1, dtype=np.uint64)
        self._chk(lib().kgo_body_offsets(self._h, off.ctypes.data_as(abi.u64p), len(off)))
        return off

    def series_count(self):
        return int(lib().kgo_series_count(self._h))

    def read_values(self, family):
        n = C.c_size_t()
        lib().kgo_read_values(self._h, family, None, 0, C.byref(n))
        out = np.zeros(n.value, dtype=np.float64)
        self._chk(lib().kgo_read_values(self._h, family, out.ctypes.data_as(C.POINTER(C.c_double)), len(out), C.byref(n)))
        return out
