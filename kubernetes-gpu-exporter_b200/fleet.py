"""Deterministic synthetic fleets: what a fleet of nodes
running the reference exporter would hand to the hot path each cycle — a
pod->PID table (main.go:84-114) and per-process NVML samples (main.go:123-138).

seed = 0xB200 + config index; all randomness is counter-based SplitMix64, so a
fleet is reproducible from (config, scrape index) alone.  Layout rules:
  * per node: G=8 GPUs of 192 GiB; `pods` pods named pod-<node>-<i> in
    namespace ns-<i mod 16>; pod i has 1 + (i & 1) containers c0[, c1];
  * process j of a node belongs to pod j mod pods, container (j // pods) mod
    n_containers(pod); PIDs are unique per node in [300, 4194304) (an affine
    bijection of j with per-node random coefficients);
  * a container is pinned to one GPU (device-plugin style, uniform over 8);
    one process in 8 runs on the next GPU instead;
  * used = 2 MiB * U[1,65536]; sm/mem util U[0,100]; ts = scrape cadence 30 s
    plus U[0,1s) jitter;
  * every 100th sample of a node is an orphan: a PID no table entry has (it
    must be dropped, as the `if` at main.go:145 never fires for it) — the
    process it replaces simply has no GPU sample;
  * sample order is shuffled inside a node; nodes are concatenated in order
    (node_off given) unless shuffle_nodes=True.
"""
import math

import numpy as np

from .abi import Cols, Table, MODE_X, pack_strings

GIB = 1 << 30
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)
PID_LO, PID_HI = 300, 4194304

CONFIGS = {
    # idx: (n_nodes, pods_per_node, procs_per_node, scrapes)
    1: dict(n_nodes=1, pods=64, procs=256, scrapes=1),
    2: dict(n_nodes=1, pods=1000, procs=16000, scrapes=1),
    3: dict(n_nodes=128, pods=1000, procs=8192, scrapes=1),
    4: dict(n_nodes=2000, pods=1000, procs=5000, scrapes=1),
    5: dict(n_nodes=128, pods=1000, procs=8192, scrapes=60),
}


def splitmix64(x):
    """Vectorised SplitMix64 finaliser over uint64 counters."""
    with np.errstate(over="ignore"):
        z = (np.asarray(x, dtype=np.uint64) + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


class Fleet:
    def __init__(self, n_nodes, pods, procs, seed, gpus=8, scrapes=1, node_base=0, total_bytes=192 * GIB):
        self.n_nodes, self.pods, self.procs, self.G = int(n_nodes), int(pods), int(procs), int(gpus)
        self.seed, self.scrapes, self.node_base = int(seed), int(scrapes), int(node_base)
        self.total_bytes = int(total_bytes)
        self._stream = 0
        M = PID_HI - PID_LO
        nid = np.arange(self.n_nodes, dtype=np.uint64) + np.uint64(self.node_base)
        ra = self._rand(nid, 1)
        rb = self._rand(nid, 2)
        a = (ra % np.uint64(M)).astype(np.int64) | 1
        for i in range(len(a)):  # make a coprime to M so j -> a*j+b mod M is a bijection
            while math.gcd(int(a[i]), M) != 1:
                a[i] += 2
        self._a, self._b, self._M = a, (rb % np.uint64(M)).astype(np.int64), M

    @classmethod
    def config(cls, idx, n_nodes=None, node_base=0):
        c = CONFIGS[idx]
        return cls(n_nodes or c["n_nodes"], c["pods"], c["procs"], 0xB200 + idx, scrapes=c["scrapes"],
                   node_base=node_base)

    def _rand(self, counter, stream):
        """uint64 random per counter, independent per stream id."""
        with np.errstate(over="ignore"):
            base = splitmix64(np.uint64(self.seed) + np.uint64(stream) * np.uint64(0xD1B54A32D192ED03))
            return splitmix64(np.asarray(counter, dtype=np.uint64) * np.uint64(0x2545F4914F6CDD1D) + base)

    def _pid_of(self, node_local, j):
        """PID of process index j (may be >= procs: orphan space) on a node."""
        a, b = self._a[node_local], self._b[node_local]
        return (PID_LO + (a * j.astype(np.int64) + b) % self._M).astype(np.uint32)

    # ---- table ----------------------------------------------------------------
    def table(self, mode):
        nn, P, Q, G = self.n_nodes, self.pods, self.procs, self.G
        node_ids = np.arange(nn) + self.node_base
        pod_i = np.arange(P)
        j = np.arange(Q, dtype=np.int64)
        proc_pod = j % P
        ncont = 1 + (proc_pod & 1)
        proc_ct = (j // P) % ncont
        pids = np.stack([self._pid_of(n, j) for n in range(nn)])  # [nn, Q]
        if mode == MODE_X:
            # podkeys: per node, pod-major then container; podkey id = node*K + local
            pk_pod = np.repeat(pod_i, 1 + (pod_i & 1))
            pk_ct = np.concatenate([np.arange(1 + (i & 1)) for i in pod_i]) if P else np.zeros(0, int)
            K = len(pk_pod)
            first_pk = np.concatenate([[0], np.cumsum(1 + (pod_i & 1))])[:-1]
            proc_pk_local = first_pk[proc_pod] + proc_ct
            pod_names, ns_names, ct_names = [], [], []
            for n in node_ids:
                pod_names += ["pod-%d-%d" % (n, i) for i in pk_pod]
            ns_one = ["ns-%d" % (i % 16) for i in pk_pod]
            ct_one = ["c%d" % c for c in pk_ct]
            podkey_node = np.repeat(np.arange(nn), K)
            entry_podkey = (np.arange(nn)[:, None] * K + proc_pk_local[None, :]).reshape(-1)
            entry_pid = pids.reshape(-1)
            uuids = ["GPU-%08x-%04x-%04x-%04x-%012x" % (int(v) & 0xFFFFFFFF, (int(v) >> 32) & 0xFFFF, (int(v) >> 48) & 0xFFFF,
                                                         int(w) & 0xFFFF, (int(w) >> 16) & 0xFFFFFFFFFFFF)
                     for v, w in zip(self._rand(np.arange(nn * G) + self.node_base * G, 3),
                                     self._rand(np.arange(nn * G) + self.node_base * G, 4))]
            ns_blob, ns_off = pack_strings(ns_one)
            ct_blob, ct_off = pack_strings(ct_one)
            return Table(nn, G, podkey_node, pod_names, entry_podkey, entry_pid,
                         namespace=_tile_strings(ns_blob, ns_off, nn), container=_tile_strings(ct_blob, ct_off, nn),
                         gpu_uuid=uuids)
        # R modes: podkey = pod; the PID list of a pod is repeated once per
        # container, as main.go:101-109 execs `ps` in the default container
        # once per ContainerStatuses entry.
        pod_names = []
        for n in node_ids:
            pod_names += ["pod-%d-%d" % (n, i) for i in pod_i]
        podkey_node = np.repeat(np.arange(nn), P)
        rep = np.where(ncont == 2)[0]
        ent_pod_local = np.concatenate([proc_pod, proc_pod[rep]])
        ent_j = np.concatenate([j, j[rep]])
        entry_podkey = (np.arange(nn)[:, None] * P + ent_pod_local[None, :]).reshape(-1)
        entry_pid = pids[:, ent_j].reshape(-1)
        return Table(nn, G, podkey_node, pod_names, entry_podkey, entry_pid)

    # ---- samples --------------------------------------------------------------
    def scrape(self, k=0, shuffle_nodes=False, x_cols=True):
        nn, P, Q, G = self.n_nodes, self.pods, self.procs, self.G
        j = np.arange(Q, dtype=np.int64)
        gidx = (np.arange(nn, dtype=np.uint64)[:, None] + np.uint64(self.node_base)) * np.uint64(Q) + j[None, :].astype(np.uint64)
        orphan = (j % 100) == 99
        jj = np.where(orphan, j + Q, j)
        pid = np.stack([self._pid_of(n, jj) for n in range(nn)])
        proc_pod = j % P
        ncont = 1 + (proc_pod & 1)
        proc_ct = (j // P) % ncont
        ckey = (np.arange(nn, dtype=np.uint64)[:, None] + np.uint64(self.node_base)) * np.uint64(2 * P) + \
            (proc_pod * 2 + proc_ct)[None, :].astype(np.uint64)
        cgpu = self._rand(ckey, 5) % np.uint64(G)
        hop = (self._rand(gidx, 6) % np.uint64(8)) == 0
        gpu = ((cgpu + hop.astype(np.uint64)) % np.uint64(G)).astype(np.uint32)
        kk = np.uint64(k) * np.uint64(0x100000000)
        used = (np.uint64(2 << 20) * (np.uint64(1) + self._rand(gidx + kk, 7) % np.uint64(65536)))
        order = np.argsort(self._rand(gidx + kk, 8), axis=1, kind="stable")
        node = np.repeat(np.arange(nn, dtype=np.uint32)[:, None], Q, axis=1)
        take = lambda a: np.take_along_axis(a, order, axis=1).reshape(-1)
        cols = dict(pid=take(pid), gpu=take(gpu), used=take(used), node=node.reshape(-1))
        if x_cols:
            sm = (self._rand(gidx + kk, 9) % np.uint64(101)).astype(np.uint32)
            mem = (self._rand(gidx + kk, 10) % np.uint64(101)).astype(np.uint32)
            ts = np.uint64(1_790_000_000_000_000) + np.uint64(k) * np.uint64(30_000_000) + self._rand(gidx + kk, 11) % np.uint64(1_000_000)
            cols.update(sm_util=take(sm), mem_util=take(mem), ts_us=take(ts))
        node_off = np.arange(nn + 1, dtype=np.uint64) * np.uint64(Q)
        if shuffle_nodes:
            perm = np.argsort(self._rand(np.arange(nn * Q, dtype=np.uint64) + kk, 12), kind="stable")
            cols = {key: v[perm] for key, v in cols.items()}
            node_off = None
        gpu_total = np.full(nn * G, self.total_bytes, dtype=np.uint64)
        return Cols(gpu_total=gpu_total, node_off=node_off, **cols)


def _tile_strings(blob, off, reps):
    """Repeat a packed string list `reps` times without re-encoding."""
    body = blob[:-1]
    n = len(off) - 1
    lens = np.diff(off)
    new_off = np.zeros(n * reps + 1, dtype=np.uint64)
    new_off[1:] = np.cumsum(np.tile(lens, reps))
    return np.concatenate([np.tile(body, reps), np.zeros(1, np.uint8)]), new_off
