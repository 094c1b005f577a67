"""ctypes binding of csrc/libkgx.so (the C ABI of include/kgx.h).

This is the product path: it fails loudly if the CUDA library is missing or no
GPU is present — there is no CPU fallback and the oracle is never imported."""
import ctypes as C
import os

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("KGX_LIB") or os.path.join(_HERE, "csrc", "libkgx.so")  # KGX_LIB: an experiment build (tools/)
_LIB = None

EXPORTS = [
    "kgx_abi_version", "kgx_strerror", "kgx_last_error", "kgx_create", "kgx_destroy", "kgx_set_table", "kgx_ingest",
    "kgx_ingest_device", "kgx_ingest_nvml", "kgx_ingest_nvml_x", "kgx_reset_window", "kgx_expire", "kgx_render_device", "kgx_render", "kgx_render_begin", "kgx_render_end", "kgx_body_device", "kgx_body_offsets", "kgx_body_node", "kgx_body_read",
    "kgx_read_values", "kgx_read_labels", "kgx_render_noalloc", "kgx_family_count", "kgx_family_info", "kgx_slot_labels", "kgx_format_values", "kgx_host_alloc",
    "kgx_host_free", "kgx_get_stats", "kgx_get_profile", "kgx_profile_reset", "kgx_set_profile", "kgx_sync",
]


class KgxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("kgx error %d (%s): %s" % (code, lib().kgx_strerror(code).decode(), msg))
        self.code = code


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(SO_PATH):
            raise ImportError("libkgx.so is not built (%s): run __graft_entry__.build(); there is no CPU fallback" % SO_PATH)
        L = C.CDLL(SO_PATH)
        vp = C.c_void_p
        L.kgx_strerror.restype = C.c_char_p
        L.kgx_strerror.argtypes = [C.c_int]
        L.kgx_last_error.restype = C.c_char_p
        L.kgx_last_error.argtypes = [vp]
        L.kgx_create.argtypes = [C.c_int, C.POINTER(abi.kgx_opts), C.POINTER(vp)]
        L.kgx_destroy.argtypes = [vp]
        L.kgx_destroy.restype = None
        L.kgx_set_table.argtypes = [vp, C.POINTER(abi.kgx_table)]
        L.kgx_ingest.argtypes = [vp, C.POINTER(abi.kgx_cols)]
        L.kgx_ingest_device.argtypes = [vp, C.POINTER(abi.kgx_cols)]
        L.kgx_reset_window.argtypes = [vp]
        L.kgx_expire.argtypes = [vp, C.c_uint64, C.POINTER(C.c_uint64)]
        L.kgx_ingest_nvml.argtypes = [vp, vp, abi.u64p, abi.u64p]
        L.kgx_ingest_nvml_x.argtypes = [vp, vp, abi.u64p, vp, abi.u64p, abi.u64p]
        L.kgx_render_device.argtypes = [vp, C.POINTER(C.c_size_t)]
        L.kgx_render.argtypes = [vp, vp, C.c_size_t, C.POINTER(C.c_size_t)]
        L.kgx_render_begin.argtypes = [vp, vp, C.c_size_t, C.POINTER(C.c_size_t)]
        L.kgx_render_end.argtypes = [vp]
        L.kgx_body_device.argtypes = [vp, C.POINTER(vp), C.POINTER(C.c_size_t)]
        L.kgx_body_offsets.argtypes = [vp, abi.u64p, C.c_size_t]
        L.kgx_body_node.argtypes = [vp, C.c_uint32, vp, C.c_size_t, C.POINTER(C.c_size_t)]
        L.kgx_body_read.argtypes = [vp, C.c_uint64, vp, C.c_size_t, C.POINTER(C.c_size_t)]
        L.kgx_read_labels.argtypes = [vp, vp, C.c_size_t, abi.u64p, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]
        L.kgx_render_noalloc.argtypes = [vp, vp, C.c_size_t, C.POINTER(C.c_size_t)]
        L.kgx_read_values.argtypes = [vp, C.c_int, C.POINTER(C.c_double), abi.u32p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.kgx_family_count.argtypes = [vp]
        L.kgx_family_info.argtypes = [vp, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p),
                                      C.POINTER(C.POINTER(C.c_char_p)), C.POINTER(C.c_int)]
        L.kgx_slot_labels.argtypes = [vp, C.c_uint32, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.kgx_format_values.argtypes = [vp, C.POINTER(C.c_double), C.c_size_t, vp]
        L.kgx_host_alloc.restype = vp
        L.kgx_host_alloc.argtypes = [C.c_size_t]
        L.kgx_host_free.argtypes = [vp]
        L.kgx_host_free.restype = None
        L.kgx_get_stats.argtypes = [vp, C.POINTER(abi.kgx_stats)]
        L.kgx_get_profile.argtypes = [vp, C.POINTER(abi.kgx_profile)]
        L.kgx_profile_reset.argtypes = [vp]
        L.kgx_set_profile.argtypes = [vp, C.c_int]
        L.kgx_sync.argtypes = [vp]
        _LIB = L
    return _LIB


def pinned_empty(n, dtype):
    """numpy array over pinned host memory from kgx_host_alloc (kept alive by the array)."""
    dtype = np.dtype(dtype)
    nbytes = max(1, int(n) * dtype.itemsize)
    p = lib().kgx_host_alloc(nbytes)
    if not p:
        raise MemoryError("kgx_host_alloc(%d) failed" % nbytes)
    buf = (C.c_uint8 * nbytes).from_address(p)
    arr = np.frombuffer(buf, dtype=dtype, count=int(n))
    _PINNED[arr.ctypes.data] = p
    return arr


_PINNED = {}


def pinned_free(arr):
    p = _PINNED.pop(arr.ctypes.data, None)
    if p:
        lib().kgx_host_free(p)


class Kgx:
    """One context on one GPU."""

    def __init__(self, mode, device=0, stream=None, profile=False):
        self._h = C.c_void_p()
        self.mode = mode
        o = abi.kgx_opts(abi.KGX_ABI_VERSION, mode, stream, int(profile), 0)
        rc = lib().kgx_create(device, C.byref(o), C.byref(self._h))
        if rc != 0:
            self._h = C.c_void_p()
            raise KgxError(rc, "kgx_create(device=%d) failed — a CUDA device is required" % device)
        self.n_nodes = 0
        self._out = None

    def close(self):
        if self._h:
            lib().kgx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc):
        if rc != 0:
            raise KgxError(rc, lib().kgx_last_error(self._h).decode())

    def set_table(self, table):
        t = table.as_struct()
        self._chk(lib().kgx_set_table(self._h, C.byref(t)))
        self.n_nodes = table.n_nodes

    def ingest(self, cols):
        s = cols.as_struct()
        self._chk(lib().kgx_ingest(self._h, C.byref(s)))

    def ingest_device(self, cols_struct):
        self._chk(lib().kgx_ingest_device(self._h, C.byref(cols_struct)))

    def ingest_nvml(self, infos, dev_off, gpu_total):
        """infos: structured array with the nvmlProcessInfo_t layout (abi.NVML_PROCESS_INFO); R modes."""
        infos = np.ascontiguousarray(infos, dtype=abi.NVML_PROCESS_INFO)
        dev_off = np.ascontiguousarray(dev_off, dtype=np.uint64)
        gpu_total = np.ascontiguousarray(gpu_total, dtype=np.uint64)
        self._chk(lib().kgx_ingest_nvml(self._h, infos.ctypes.data, dev_off.ctypes.data_as(abi.u64p),
                                        gpu_total.ctypes.data_as(abi.u64p)))

    def ingest_nvml_x(self, infos, dev_off, utils, util_off, gpu_total):
        """X mode: nvmlProcessInfo_t rows joined on the GPU with nvmlProcessUtilizationSample_t rows (same pid, same device)."""
        infos = np.ascontiguousarray(infos, dtype=abi.NVML_PROCESS_INFO)
        utils = np.ascontiguousarray(utils, dtype=abi.NVML_PROCESS_UTILIZATION_SAMPLE)
        dev_off = np.ascontiguousarray(dev_off, dtype=np.uint64)
        util_off = np.ascontiguousarray(util_off, dtype=np.uint64)
        gpu_total = np.ascontiguousarray(gpu_total, dtype=np.uint64)
        self._chk(lib().kgx_ingest_nvml_x(self._h, infos.ctypes.data, dev_off.ctypes.data_as(abi.u64p), utils.ctypes.data,
                                          util_off.ctypes.data_as(abi.u64p), gpu_total.ctypes.data_as(abi.u64p)))

    def reset_window(self):
        self._chk(lib().kgx_reset_window(self._h))

    def expire(self, older_than_ts_us):
        """X mode: drop the series whose newest sample is older than the cutoff; returns how many exported series went."""
        n = C.c_uint64()
        self._chk(lib().kgx_expire(self._h, int(older_than_ts_us), C.byref(n)))
        return n.value

    def render_device(self):
        n = C.c_size_t()
        self._chk(lib().kgx_render_device(self._h, C.byref(n)))
        return n.value

    def render_into(self, out):
        """Render into a caller-owned uint8 array (host); returns the length."""
        n = C.c_size_t()
        rc = lib().kgx_render(self._h, out.ctypes.data, out.nbytes, C.byref(n))
        if rc == abi.KGX_E_NOSPACE:
            return -int(n.value)
        self._chk(rc)
        return n.value

    def render_begin(self, out):
        """Render and start copying the body into `out` (pinned); returns the length.  Pair with render_end()."""
        n = C.c_size_t()
        self._chk(lib().kgx_render_begin(self._h, out.ctypes.data, out.nbytes, C.byref(n)))
        return n.value

    def render_end(self):
        self._chk(lib().kgx_render_end(self._h))

    def render(self):
        """The body as bytes.  Convenience for single-threaded callers: it reuses one host buffer of this object;
        concurrent scrapers call render_into() with a buffer each (the C ABI itself is safe to call concurrently)."""
        if self._out is None:
            self._out = np.empty(1 << 16, dtype=np.uint8)
        n = self.render_into(self._out)
        if n < 0:
            self._out = np.empty(-n, dtype=np.uint8)
            n = self.render_into(self._out)
        return self._out[:n].tobytes()

    def body_device(self):
        p, n = C.c_void_p(), C.c_size_t()
        self._chk(lib().kgx_body_device(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def body_offsets(self):
        off = np.zeros(self.n_nodes + 1, dtype=np.uint64)
        self._chk(lib().kgx_body_offsets(self._h, off.ctypes.data_as(abi.u64p), len(off)))
        return off

    def body_node(self, node):
        """One node's /metrics body out of the last rendered text (kgx_body_node)."""
        n = C.c_size_t()
        rc = lib().kgx_body_node(self._h, int(node), None, 0, C.byref(n))
        if rc not in (0, abi.KGX_E_NOSPACE):
            self._chk(rc)
        out = np.empty(max(1, n.value), dtype=np.uint8)
        self._chk(lib().kgx_body_node(self._h, int(node), out.ctypes.data, out.nbytes, C.byref(n)))
        return out[:n.value].tobytes()

    def body_read(self, offset, out):
        """Up to len(out) bytes of the last rendered text from `offset` into the uint8 array `out` (kgx_body_read);
        returns how many were copied (0 at the end of the body)."""
        n = C.c_size_t()
        self._chk(lib().kgx_body_read(self._h, int(offset), out.ctypes.data, out.nbytes, C.byref(n)))
        return n.value

    def stream_body(self, chunk_bytes=1 << 20):
        """Render on the device, then yield the body chunk by chunk (what a chunked-transfer handler writes)."""
        self.render_device()
        buf = np.empty(chunk_bytes, dtype=np.uint8)
        at = 0
        while True:
            n = self.body_read(at, buf)
            if n == 0:
                return
            yield buf[:n].tobytes()
            at += n

    def family_count(self):
        return lib().kgx_family_count(self._h)

    def family_info(self, f):
        name, help_, names, nl = C.c_char_p(), C.c_char_p(), C.POINTER(C.c_char_p)(), C.c_int()
        self._chk(lib().kgx_family_info(self._h, f, C.byref(name), C.byref(help_), C.byref(names), C.byref(nl)))
        return name.value.decode(), help_.value.decode(), [names[i].decode() for i in range(nl.value)]

    def read_values(self, family, with_slots=False):
        n = C.c_size_t()
        rc = lib().kgx_read_values(self._h, family, None, None, 0, C.byref(n))
        if rc not in (0, abi.KGX_E_NOSPACE):
            self._chk(rc)
        vals = np.zeros(n.value, dtype=np.float64)
        slots = np.zeros(n.value, dtype=np.uint32)
        self._chk(lib().kgx_read_values(self._h, family, vals.ctypes.data_as(C.POINTER(C.c_double)),
                                        slots.ctypes.data_as(abi.u32p), len(vals), C.byref(n)))
        return (vals, slots) if with_slots else vals

    def read_labels(self):
        """(blob, offsets): label text of every series with a value, exposition order (kgx_read_labels)."""
        n, nb = C.c_size_t(), C.c_size_t()
        self._chk(lib().kgx_read_labels(self._h, None, 0, None, 0, C.byref(n), C.byref(nb)))
        blob = np.empty(max(1, nb.value), dtype=np.uint8)
        off = np.zeros(n.value + 1, dtype=np.uint64)
        self._chk(lib().kgx_read_labels(self._h, blob.ctypes.data, nb.value, off.ctypes.data_as(abi.u64p), len(off),
                                        C.byref(n), C.byref(nb)))
        return blob[:nb.value], off

    def slot_labels(self, slot):
        n = C.c_size_t()
        buf = C.create_string_buffer(4096)
        self._chk(lib().kgx_slot_labels(self._h, int(slot), buf, len(buf), C.byref(n)))
        return buf.raw[:n.value]

    def format_values(self, values):
        v = np.ascontiguousarray(values, dtype=np.float64)
        out = np.zeros((len(v), 32), dtype=np.uint8)
        self._chk(lib().kgx_format_values(self._h, v.ctypes.data_as(C.POINTER(C.c_double)), len(v), out.ctypes.data))
        return out

    def stats(self):
        s = abi.kgx_stats()
        self._chk(lib().kgx_get_stats(self._h, C.byref(s)))
        return {n: int(getattr(s, n)) for n, _ in abi.kgx_stats._fields_}

    def profile(self):
        p = abi.kgx_profile()
        self._chk(lib().kgx_get_profile(self._h, C.byref(p)))
        return {abi.K_NAMES[i]: (p.ms[i], int(p.launches[i])) for i in range(abi.K_COUNT)}

    def set_profile(self, on):
        self._chk(lib().kgx_set_profile(self._h, int(on)))

    def profile_reset(self):
        self._chk(lib().kgx_profile_reset(self._h))

    def sync(self):
        self._chk(lib().kgx_sync(self._h))
