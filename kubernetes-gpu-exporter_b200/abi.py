"""ctypes mirror of include/kgx.h plus numpy-backed holders for its two input
structs.  Shared by the product binding (kgx.py) and, in tests only, by the
oracle binding (oracle/kgo.py) so both sides are fed the very same buffers."""
import ctypes as C

import numpy as np

KGX_ABI_VERSION = 2
KGX_OK, KGX_E_INVAL, KGX_E_NOMEM, KGX_E_CUDA, KGX_E_NOSPACE, KGX_E_STATE, KGX_E_LIMIT = 0, -1, -2, -3, -4, -5, -6
MODE_R_INTENDED, MODE_R_LITERAL, MODE_X = 0, 1, 2
K_INGEST, K_FINALIZE, K_INDEX, K_RENDER, K_OTHER, K_COUNT = 0, 1, 2, 3, 4, 5
K_NAMES = ["ingest", "finalize", "index", "render", "other"]

# nvmlProcessInfo_t (nvml.h:294-303): 24 bytes
NVML_PROCESS_INFO = np.dtype([("pid", "<u4"), ("pad_", "<u4"), ("used_gpu_memory", "<u8"),
                              ("gpu_instance_id", "<u4"), ("compute_instance_id", "<u4")])

# nvmlProcessUtilizationSample_t (nvml.h:1277-1285): 32 bytes
NVML_PROCESS_UTILIZATION_SAMPLE = np.dtype([("pid", "<u4"), ("pad_", "<u4"), ("time_stamp", "<u8"), ("sm_util", "<u4"),
                                            ("mem_util", "<u4"), ("enc_util", "<u4"), ("dec_util", "<u4")])

u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)


class kgx_opts(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("mode", C.c_int32), ("stream", C.c_void_p),
                ("profile", C.c_int32), ("reserved", C.c_int32)]


class kgx_table(C.Structure):
    _fields_ = [("n_nodes", C.c_uint32), ("gpus_per_node", C.c_uint32), ("n_podkeys", C.c_uint64),
                ("podkey_node", u32p),
                ("pod_blob", C.c_void_p), ("pod_off", u64p),
                ("namespace_blob", C.c_void_p), ("namespace_off", u64p),
                ("container_blob", C.c_void_p), ("container_off", u64p),
                ("n_entries", C.c_uint64), ("entry_podkey", u32p), ("entry_pid", u32p),
                ("gpu_uuid_blob", C.c_void_p), ("gpu_uuid_off", u64p)]


class kgx_cols(C.Structure):
    _fields_ = [("n", C.c_uint64), ("pid", u32p), ("gpu", u32p), ("used", u64p), ("node", u32p),
                ("sm_util", u32p), ("mem_util", u32p), ("ts_us", u64p), ("node_off", u64p),
                ("gpu_total", u64p)]


class kgx_stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("n_series", "n_slots", "n_matched", "n_dropped", "rendered_bytes",
                                          "kernel_launches", "rebuilds", "device_bytes")]


class kgx_profile(C.Structure):
    _fields_ = [("ms", C.c_double * K_COUNT), ("launches", C.c_uint64 * K_COUNT)]


def _u32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.uint32)


def _u64(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.uint64)


def _ptr(a, typ):
    return None if a is None else a.ctypes.data_as(typ)


def pack_strings(strings):
    """list of str/bytes -> (blob bytes ndarray, offsets uint64[n+1])."""
    bs = [s.encode("utf-8") if isinstance(s, str) else bytes(s) for s in strings]
    off = np.zeros(len(bs) + 1, dtype=np.uint64)
    if bs:
        off[1:] = np.cumsum(np.fromiter((len(b) for b in bs), dtype=np.uint64, count=len(bs)))
    blob = np.frombuffer(b"".join(bs) + b"\0", dtype=np.uint8).copy()
    return blob, off


class Table:
    """The join table (podPIDMap of main.go:84 in flat form).  Strings are
    (blob, offsets) pairs from pack_strings."""

    def __init__(self, n_nodes, gpus_per_node, podkey_node, pod, entry_podkey, entry_pid,
                 namespace=None, container=None, gpu_uuid=None):
        self.n_nodes, self.gpus_per_node = int(n_nodes), int(gpus_per_node)
        self.podkey_node = _u32(podkey_node)
        self.pod = pod if isinstance(pod, tuple) else pack_strings(pod)
        self.namespace = namespace if (namespace is None or isinstance(namespace, tuple)) else pack_strings(namespace)
        self.container = container if (container is None or isinstance(container, tuple)) else pack_strings(container)
        self.gpu_uuid = gpu_uuid if (gpu_uuid is None or isinstance(gpu_uuid, tuple)) else pack_strings(gpu_uuid)
        self.entry_podkey, self.entry_pid = _u32(entry_podkey), _u32(entry_pid)
        assert len(self.entry_podkey) == len(self.entry_pid)
        assert len(self.pod[1]) == len(self.podkey_node) + 1

    def as_struct(self):
        t = kgx_table()
        t.n_nodes, t.gpus_per_node = self.n_nodes, self.gpus_per_node
        t.n_podkeys = len(self.podkey_node)
        t.podkey_node = _ptr(self.podkey_node, u32p)
        t.pod_blob, t.pod_off = self.pod[0].ctypes.data, _ptr(self.pod[1], u64p)
        if self.namespace is not None:
            t.namespace_blob, t.namespace_off = self.namespace[0].ctypes.data, _ptr(self.namespace[1], u64p)
        if self.container is not None:
            t.container_blob, t.container_off = self.container[0].ctypes.data, _ptr(self.container[1], u64p)
        if self.gpu_uuid is not None:
            t.gpu_uuid_blob, t.gpu_uuid_off = self.gpu_uuid[0].ctypes.data, _ptr(self.gpu_uuid[1], u64p)
        t.n_entries = len(self.entry_pid)
        t.entry_podkey, t.entry_pid = _ptr(self.entry_podkey, u32p), _ptr(self.entry_pid, u32p)
        return t


class Cols:
    """One batch of samples in column form (what main.go:123-138 reads)."""

    def __init__(self, pid, gpu, used, gpu_total, node=None, sm_util=None, mem_util=None, ts_us=None, node_off=None):
        self.pid, self.gpu, self.used = _u32(pid), _u32(gpu), _u64(used)
        self.node, self.sm_util, self.mem_util = _u32(node), _u32(sm_util), _u32(mem_util)
        self.ts_us, self.node_off, self.gpu_total = _u64(ts_us), _u64(node_off), _u64(gpu_total)
        self.n = len(self.pid)
        for a in (self.gpu, self.used, self.node, self.sm_util, self.mem_util, self.ts_us):
            assert a is None or len(a) == self.n

    def as_struct(self):
        s = kgx_cols()
        s.n = self.n
        s.pid, s.gpu, s.used = _ptr(self.pid, u32p), _ptr(self.gpu, u32p), _ptr(self.used, u64p)
        s.node, s.sm_util, s.mem_util = _ptr(self.node, u32p), _ptr(self.sm_util, u32p), _ptr(self.mem_util, u32p)
        s.ts_us, s.node_off, s.gpu_total = _ptr(self.ts_us, u64p), _ptr(self.node_off, u64p), _ptr(self.gpu_total, u64p)
        return s

    def nbytes(self):
        return sum(a.nbytes for a in (self.pid, self.gpu, self.used, self.node, self.sm_util, self.mem_util,
                                      self.ts_us) if a is not None)
