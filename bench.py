#!/usr/bin/env python3
"""bench.py — the headline benchmark of the join -> reduce -> render hot path.

  python bench.py --gpus N --steps K --warmup W            (N>1: under torchrun)
  python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json configs[3], "fleet replay"): 2 000 nodes, 10 M process
samples, full label set (X mode), per GPU.  One step = one pass of the hot path
over one scrape: reset window -> join+reduce (kgx_ingest) -> render.  The path
shards by node with no data-path collective, so with N GPUs every rank replays
its own 2 000-node shard (weak scaling) and `value` is the samples all ranks
processed divided by the slowest rank's device time.

Prints ONE JSON line (rank 0).  Keys beyond the base contract:
  roofline      the dominant kernel (k_render): algorithmic bytes / its mean
                CUDA-event duration, against MEASURED_PEAKS.json's hbm_gbs
  cpu_baseline  the CPU oracle (a C++ restatement of the Go path — the Go
                toolchain is absent) on a bounded sample of the same fleet
  e2e           same metric through kgx_ingest/kgx_render with HOST buffers
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import kgx_loader  # noqa: E402

kgx_loader.load()
from kgx_b200 import abi  # noqa: E402
from kgx_b200.abi import MODE_X  # noqa: E402
from kgx_b200.fleet import CONFIGS, Fleet  # noqa: E402

METRIC = "per-pod samples aggregated+rendered/sec"
UNIT = "samples/s"
CFG = 4
TRAFFIC_CSV = "r02_k_render_raw.csv"  # the `ncu --set full` raw page of the shipped writer (profiles/README.md)


def ncu_traffic(kernel_csv, measured_ms):
    """(dram read+write bytes per launch of the dominant kernel, note) from the committed `ncu --set full` raw page
    under profiles/.  A capture of a different kernel version would go stale silently, so the capture's own duration
    must agree with the duration measured in THIS run within 10 %, or the traffic is withheld (None)."""
    import csv
    path = os.path.join(ROOT, "profiles", kernel_csv)
    if not os.path.exists(path):
        return None, "no capture at profiles/%s" % kernel_csv
    rows = list(csv.reader(open(path)))
    hdr, units, vals = rows[0], rows[1], rows[-1]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    tscale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
    total = 0.0
    for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        if key not in hdr:
            return None, "capture lacks %s" % key
        i = hdr.index(key)
        total += float(vals[i].replace(",", "")) * scale.get(units[i], 1.0)
    if "gpu__time_duration.sum" in hdr and measured_ms > 0:
        i = hdr.index("gpu__time_duration.sum")
        cap_ms = float(vals[i].replace(",", "")) * tscale.get(units[i], 1.0)
        if abs(cap_ms - measured_ms) / measured_ms > 0.10:
            return None, "stale capture withheld: profiles/%s ran %.3f ms, this run's kernel %.3f ms" % (kernel_csv, cap_ms, measured_ms)
    return total, "profiles/%s (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, one launch)" % kernel_csv


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) < 9:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def build_fleet(rank, n_nodes):
    f = Fleet.config(CFG, n_nodes=n_nodes, node_base=rank * n_nodes)
    return f, f.table(MODE_X), f.scrape(0)


def cpu_leg(fleet_nodes, threads, steps, warmup, join_hash=True):
    """Times the oracle (CPU restatement of main.go:141-154 + Gather + expfmt) on
    a bounded sample of the workload: the first `fleet_nodes` nodes."""
    kgo = kgx_loader.load_oracle()
    f = Fleet.config(CFG, n_nodes=fleet_nodes)
    t, s = f.table(MODE_X), f.scrape(0)
    o = kgo.Oracle(MODE_X, kgo.JOIN_HASH if join_hash else kgo.JOIN_LOOPS, threads)
    o.set_table(t)
    times = []
    for k in range(warmup + steps):
        t0 = time.perf_counter()
        o.reset_window()
        o.ingest(s)
        body = o.render()
        dt = time.perf_counter() - t0
        if k >= warmup:
            times.append(dt)
    return s.n, float(np.mean(times)), len(body)


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  The Go
    program cannot be built here (no Go toolchain, modules not vendored, binary
    stripped), so this arm times the C++ restatement in oracle/ with all host
    threads, on a bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    nodes = args.ref_nodes
    n, sec, body_len = cpu_leg(nodes, threads, args.steps, args.warmup)
    v = n / sec
    go = subprocess.run("go version", shell=True, capture_output=True, text=True)
    sample = "first %d of %d nodes of the fleet (%d samples, %d rendered bytes) per step" % (
        nodes, CONFIGS[CFG]["n_nodes"], n, body_len)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64/f64", "data": "synthetic",
        "config": dict(workload_config(CONFIGS[CFG]["n_nodes"]),
                       ran="a bounded sample of that workload per step: the first %d of its %d nodes, %d samples (rates compare; "
                           "the whole fleet would take ~20x longer per step)" % (nodes, CONFIGS[CFG]["n_nodes"], n)),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                         "note": "C++ restatement of the Go path, not Go: " +
                                 ("go toolchain present but reference deps are not vendored" if go.returncode == 0
                                  else "Go path unavailable on this box (no go toolchain)")},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_config(n_nodes):
    c = CONFIGS[CFG]
    return {"workload": "fleet replay (BASELINE.json configs[3]): %d nodes x %d samples, %d pods/node, 8 GPUs/node, "
                        "X mode full label set (container,gpu,namespace,pod), 6 families" % (n_nodes, c["procs"], c["pods"]),
            "samples_per_gpu": n_nodes * c["procs"], "step": "reset window + join/reduce + render of one scrape",
            "l2": "inputs (360 MB) and outputs (GBs) per step exceed the 126 MB L2; no explicit flush",
            "sharding": "by node, one fleet shard per GPU, no data-path collective"}


class near_gpu:
    """Run the enclosed allocations on CPUs of the NUMA node the GPU hangs off, so that pinned host buffers are
    placed there (first touch) and the body does not cross the socket interconnect; the affinity is restored on exit
    (the CPU baseline legs keep every core).  A no-op when the topology cannot be read."""

    def __init__(self, dev):
        self.dev, self.saved, self.note = dev, None, "numa: not bound"

    def __enter__(self):
        try:
            import torch
            p = torch.cuda.get_device_properties(self.dev)
            bus = "%04x:%02x:%02x.0" % (getattr(p, "pci_domain_id", 0), p.pci_bus_id, p.pci_device_id)
            node = int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read())
            if node >= 0:
                cpus = set()
                for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
                    lo, _, hi = part.partition("-")
                    cpus.update(range(int(lo), int(hi or lo) + 1))
                cpus &= os.sched_getaffinity(0)
                if cpus:
                    self.saved = os.sched_getaffinity(0)
                    os.sched_setaffinity(0, cpus)
                    self.note = "numa: host buffers allocated on node %d (gpu %s)" % (node, bus)
        except Exception as e:  # noqa: BLE001
            self.note = "numa: not bound (%s)" % type(e).__name__
        return self

    def __exit__(self, *a):
        if self.saved is not None:
            os.sched_setaffinity(0, self.saved)
        return False


def copy_floor(dev, h2d_bytes, d2h_bytes, steps, barrier, world):
    """What the link alone allows: per step one pinned D2H copy of a body and, the other way at the same time, one
    pinned H2D copy of a scrape — no kernels, nothing else.  CUDA events on both streams, slower stream counts, max
    over ranks (all ranks copy concurrently, as they do in the e2e leg)."""
    import torch
    import torch.distributed as dist
    if d2h_bytes <= 0:
        return None
    with near_gpu(dev.index):
        h_out = torch.empty(d2h_bytes, dtype=torch.uint8, pin_memory=True)
        h_in = torch.empty(max(1, h2d_bytes), dtype=torch.uint8, pin_memory=True)
        h_out[::4096] = 0
        h_in[::4096] = 0
    d_out = torch.empty(d2h_bytes, dtype=torch.uint8, device=dev)
    d_in = torch.empty(max(1, h2d_bytes), dtype=torch.uint8, device=dev)
    s_out, s_in = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

    def run(k):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record(s_out); ev[2].record(s_in)
        for _ in range(k):
            with torch.cuda.stream(s_out):
                h_out.copy_(d_out, non_blocking=True)
            with torch.cuda.stream(s_in):
                d_in.copy_(h_in, non_blocking=True)
        ev[1].record(s_out); ev[3].record(s_in)
        torch.cuda.synchronize()
        return max(ev[0].elapsed_time(ev[1]), ev[2].elapsed_time(ev[3])) / k

    run(1)
    barrier()
    ms = run(steps)
    barrier()
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    del h_out, h_in, d_out, d_in
    return float(t.item())


def device_cols(cols, dev):
    """The sample columns resident in HBM (torch owns the buffers): (kgx_cols struct of device pointers, keep-alive)."""
    import torch

    def dev_t(a):
        return None if a is None else torch.from_numpy(a.view(np.int64) if a.dtype == np.uint64 else a.view(np.int32)).to(dev)

    keep = {k: dev_t(getattr(cols, k)) for k in ("pid", "gpu", "used", "node", "sm_util", "mem_util", "ts_us", "node_off", "gpu_total")}
    ds = abi.kgx_cols()
    ds.n = cols.n
    for k, t in keep.items():
        if t is not None:
            setattr(ds, k, C.cast(t.data_ptr(), abi.u64p if t.dtype == torch.int64 else abi.u32p))
    return ds, keep


def timed_steps(g, ds, stream, steps, warmup, barrier, x_mode=True, clocks=None):
    """W untimed + K timed passes of the hot path over one resident scrape; CUDA events on the ctx's stream."""
    import torch

    def step():
        if x_mode:
            g.reset_window()
        g.ingest_device(ds)
        return g.render_device()

    with torch.cuda.stream(stream):
        for _ in range(warmup):
            body_len = step()
        g.sync()
        g.profile_reset()
        launches0 = g.stats()["kernel_launches"]
        if clocks:
            clocks.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(steps):
            body_len = step()
        e1.record(stream)
        barrier()
    return e0.elapsed_time(e1) / steps, body_len, g.stats()["kernel_launches"] - launches0


def sharded_leg(cfg, mode, rank, world, dev, local, steps, warmup, barrier, oracle_nodes=3, whole_body_oracle=False):
    """Strong scaling: ONE fleet of BASELINE config `cfg` split by node over the ranks (fleetshard.node_range), each
    rank joins/reduces/renders its own nodes, then the bodies are concatenated in node order on rank 0 (the path's
    only cross-rank step).  Outside the timed regions the concatenation is checked: every rank's slice
    against that rank's own checksum, and sampled nodes (or the whole body) byte for byte against the CPU oracle."""
    import torch
    import torch.distributed as dist
    from kgx_b200 import fleetshard, kgx
    n_total = CONFIGS[cfg]["n_nodes"]
    lo, hi = fleetshard.node_range(rank, world, n_total)
    f = Fleet.config(cfg, n_nodes=hi - lo, node_base=lo)
    table, cols = f.table(mode), f.scrape(0)
    stream = torch.cuda.Stream(device=dev)
    g = kgx.Kgx(mode, device=local, stream=stream.cuda_stream, profile=True)
    g.set_table(table)
    ds, keep = device_cols(cols, dev)
    ms, body_len, launches = timed_steps(g, ds, stream, steps, warmup, barrier, x_mode=(mode == MODE_X))
    ms_max = fleetshard.max_over_ranks(ms, device=dev)
    prof = g.profile()
    total_samples = CONFIGS[cfg]["n_nodes"] * CONFIGS[cfg]["procs"]
    # ---- ordered concatenation on rank 0, timed on the device, max over ranks
    g.render_device()
    ptr_, n_ = g.body_device()
    body = torch.as_tensor(fleetshard.DeviceBytes(ptr_, n_), device=dev) if n_ else torch.empty(0, dtype=torch.uint8, device=dev)
    lengths, offs = fleetshard.gather_layout(n_, device=dev)
    root_out = torch.empty(int(offs[-1]), dtype=torch.uint8, device=dev) if rank == 0 else None
    for _ in range(2):
        fleetshard.gather_bodies(body, dst=0, out=root_out)
    barrier()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    g0.record()
    for _ in range(reps):
        fleetshard.gather_bodies(body, dst=0, out=root_out)
    g1.record()
    barrier()
    gms = fleetshard.max_over_ranks(g0.elapsed_time(g1) / reps, device=dev)
    # ---- parity of what arrived (untimed)
    if rank == 0:
        root_out.fill_(0)
    fleetshard.gather_bodies(body, dst=0, out=root_out)
    sums = fleetshard.gather_checksums(body)
    node_offs = [None] * world
    dist.all_gather_object(node_offs, [int(x) for x in g.body_offsets()])
    out = None
    if rank == 0:
        torch.cuda.synchronize()
        checks_ok = fleetshard.verify_gather(root_out, offs, sums)
        kgo = kgx_loader.load_oracle()  # the checker, outside every timed region
        checked, oracle_ok = 0, True
        if whole_body_oracle:
            whole = Fleet.config(cfg)
            o = kgo.Oracle(mode, kgo.JOIN_HASH, os.cpu_count() or 1)
            o.set_table(whole.table(mode))
            o.ingest(whole.scrape(0))
            oracle_ok = root_out.cpu().numpy().tobytes() == o.render()
            checked = n_total
        else:
            for r in range(world):
                rlo, rhi = fleetshard.node_range(r, world, n_total)
                for node in sorted({rlo, (rlo + rhi) // 2, rhi - 1})[:oracle_nodes]:
                    sub = Fleet.config(cfg, n_nodes=1, node_base=node)
                    o = kgo.Oracle(mode, kgo.JOIN_HASH, 1)
                    o.set_table(sub.table(mode))
                    o.ingest(sub.scrape(0))
                    a = int(offs[r]) + node_offs[r][node - rlo]
                    b = int(offs[r]) + node_offs[r][node - rlo + 1]
                    oracle_ok = oracle_ok and root_out[a:b].cpu().numpy().tobytes() == o.render()
                    checked += 1
        inbound = int(offs[-1]) - int(lengths[0])
        out = {"workload": "BASELINE.json configs[%d] split by node: %d nodes -> %s nodes per GPU" % (
                   cfg - 1, n_total, "/".join(str(fleetshard.node_range(r, world, n_total)[1] - fleetshard.node_range(r, world, n_total)[0])
                                              for r in sorted({0, world - 1}))),
               "mode": {MODE_X: "X", 0: "R-intended", 1: "R-literal"}[mode], "samples_total": total_samples,
               "ms_per_step": ms_max, "value": total_samples / (ms_max * 1e-3), "unit": UNIT, "gpu_launches_per_step": launches / steps,
               "kernel_ms": {k: v[0] / max(1, steps) for k, v in prof.items()},
               "gather": {"ms": gms, "bytes_on_root": int(offs[-1]), "inbound_bytes": inbound,
                          "inbound_gbs": inbound / (gms * 1e-3) / 1e9 if gms > 0 else None,
                          "frac_of_nvlink_inbound_900": inbound / (gms * 1e-3) / 1e9 / 900.0 if gms > 0 else None,
                          "how": "lengths all-gather + one batched group of NCCL send/recv into rank 0's buffer, bodies in node order",
                          "checksums_ok": bool(checks_ok), "oracle_ok": bool(oracle_ok),
                          "oracle_checked": ("whole body" if whole_body_oracle else "%d nodes (first/middle/last of every rank)" % checked),
                          "parity_ok": bool(checks_ok and oracle_ok)}}
    g.close()
    del keep
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="kgx", choices=["kgx", "reference"])
    ap.add_argument("--nodes", type=int, default=CONFIGS[CFG]["n_nodes"], help="nodes per GPU (default: the full config)")
    ap.add_argument("--ref-nodes", type=int, default=100, help="nodes in the CPU sample")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-strong", action="store_true",
                    help="N>1: skip the strong-scaling legs (split fleet, ordered gather on rank 0, parity check)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "kgx" else args.warmup
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from kgx_b200 import fleetshard, kgx

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        fleetshard.nccl_gather_env()
        dist.init_process_group("nccl", device_id=dev)

    fleet, table, cols = build_fleet(rank, args.nodes)
    stream = torch.cuda.Stream(device=dev)
    g = kgx.Kgx(MODE_X, device=local, stream=stream.cuda_stream, profile=True)
    g.set_table(table)

    ds, _keep = device_cols(cols, dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    clocks = ClockSampler(local)
    ms, body_len, launches = timed_steps(g, ds, stream, args.steps, args.warmup, barrier, clocks=clocks)
    clk = clocks.stop()
    prof = g.profile()
    st = g.stats()
    tmax = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms_max = float(tmax.item())
    total_samples = cols.n * world
    value = total_samples / (ms_max * 1e-3)

    # ---- roofline of the dominant kernel (render) -------------------------------
    peak, peak_src = peaks()
    n_fam = g.family_count()
    render_ms = prof["render"][0] / max(1, prof["render"][1])
    ingest_ms = prof["ingest"][0] / max(1, args.steps)  # join + its one-thread statistics hand-over
    series = st["n_series"]
    label_bytes = label_bytes_of(g, table, series)
    alg_render = n_fam * label_bytes + body_len
    achieved = alg_render / (render_ms * 1e-3) / 1e9 if render_ms > 0 else 0.0
    alg_pipeline = cols.n * 36 + len(table.entry_pid) * 12 + alg_render
    traffic, traffic_src = ncu_traffic(TRAFFIC_CSV, render_ms)
    roofline = {"bound": "hbm", "kernel": "k_render5<128> (exposition writer, kgx_writer5.cuh)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src, "traffic_source": traffic_src,
                "algorithmic_bytes_per_launch": alg_render, "kernel_ms": render_ms,
                "share_of_step": render_ms / ms if ms else None,
                "pipeline": {"algorithmic_bytes_per_step": alg_pipeline, "achieved_gbs": alg_pipeline / (ms * 1e-3) / 1e9,
                             "frac": alg_pipeline / (ms * 1e-3) / 1e9 / peak, "ingest_ms": ingest_ms, "render_ms": render_ms},
                "frac_of_nominal_8000": achieved / 8000.0}

    # ---- e2e: host buffers in, host bytes out, through the C ABI ------------------
    e2e = None
    if not args.no_e2e:
        host = {}
        with near_gpu(local) as numa_note:  # the exporter's buffers live on the GPU's own NUMA node
            for k in ("pid", "gpu", "used", "node", "sm_util", "mem_util", "ts_us", "node_off", "gpu_total"):
                a = getattr(cols, k)
                if a is not None:
                    host[k] = kgx.pinned_empty(len(a), a.dtype)
                    host[k][:] = a
            outs = [kgx.pinned_empty(body_len + 4096, np.uint8) for _ in range(2)]  # bodies alternate between two host buffers
            for o_ in outs:
                o_[::4096] = 0  # touch every page here
        hcols = abi.Cols(**host)
        h2d = hcols.nbytes() + hcols.gpu_total.nbytes + (hcols.node_off.nbytes if hcols.node_off is not None else 0)
        state = {"k": 0, "pending": False}

        def step_e2e():
            # the call sequence of a pipelined exporter: while body k-1 is still on its way to the host, scrape k is
            # copied in, joined and rendered (kgx_render_begin: into the library's second body buffer, its copy
            # queued behind the one in flight); then body k-1 is awaited (kgx_render_end).  Every body is fully
            # delivered, the last one inside the timed region (drain).
            g.reset_window()
            g.ingest(hcols)
            n = g.render_begin(outs[state["k"] % 2])
            if state["pending"]:
                g.render_end()
            state["k"] += 1
            state["pending"] = True
            return n

        def drain():
            if state["pending"]:
                g.render_end()
                state["pending"] = False

        with torch.cuda.stream(stream):
            for _ in range(2):
                n_out = step_e2e()
            drain()
            barrier()
            t0 = time.perf_counter()
            ee0, ee1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ee0.record(stream)
            k_e2e = max(3, min(args.steps, 50))  # the last body is delivered inside the timed region too: more steps amortise that tail
            for _ in range(k_e2e):
                n_out = step_e2e()
            drain()
            ee1.record(stream)
            barrier()
            wall = (time.perf_counter() - t0) / k_e2e
        e_ms = max(ee0.elapsed_time(ee1) / k_e2e, wall * 1e3)  # the host call blocks: take the larger clock
        tm = torch.tensor([e_ms], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        floor_ms = copy_floor(dev, int(h2d), int(n_out), max(3, min(args.steps, 10)), barrier, world)
        e2e = {"value": total_samples / (float(tm.item()) * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(n_out), "ms_per_step": float(tm.item()), "host_memory": "pinned (kgx_host_alloc); " + numa_note.note,
               "api": "kgx_reset_window + kgx_ingest + kgx_render_begin/kgx_render_end (body k-1 leaves while scrape k comes in)",
               "copy_floor_ms": floor_ms, "frac_of_floor": floor_ms / float(tm.item()) if floor_ms else None,
               "copy_floor": "bare copies of the same sizes in this process, no kernels: one pinned device->host copy of "
                             "d2h_bytes_per_step per step with one pinned host->device copy of h2d_bytes_per_step running the "
                             "other way on a second stream, all N ranks at once, max over ranks"}

    # ---- N>1: the 10 M-sample fleet SPLIT over the ranks (strong scaling), gathered in order and byte-checked;
    # and BASELINE config 3 (rack replay, 16 nodes per GPU at N=8) the same way ---------------------------------
    strong = cfg3 = gather = None
    if world > 1 and not args.no_strong:
        g.close()  # the weak leg's context (several GB) is not needed any more
        del _keep
        strong = sharded_leg(CFG, MODE_X, rank, world, dev, local, args.steps, args.warmup, barrier)
        cfg3 = sharded_leg(3, abi.MODE_R_INTENDED, rank, world, dev, local, args.steps, args.warmup, barrier, whole_body_oracle=True)
        if rank == 0:
            # the weak leg above IS the N=1 run of the same 10 M-sample fleet (every rank replays all 2 000 nodes)
            strong["n1_ms_per_step"] = ms_max
            strong["efficiency_vs_n1"] = ms_max / (world * strong["ms_per_step"])
            strong["note"] = ("efficiency = (N=1 time of the whole fleet, this run's weak leg) / (N x time of the split fleet); "
                              "kernel_ms vs ms_per_step shows the per-step host/launch share that does not shrink with the shard")
            gather = strong.pop("gather")

    # ---- CPU baseline on this box's host cores (rank 0, N=1 only) --------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu_nodes = min(4 * args.ref_nodes, args.nodes)  # ~10 s of single-core work at the default
        n, sec, blen = cpu_leg(cpu_nodes, 1, 2, 1)
        cpu = {"value": n / sec, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": "first %d nodes of the fleet (%d samples, %d rendered bytes), 1 warm-up + 2 timed passes" % (cpu_nodes, n, blen),
               "note": "C++ restatement of the Go path (oracle/), not Go: no Go toolchain on the box"}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_max, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64/f64",
                "data": "synthetic", "config": workload_config(args.nodes), "clocks": clk, "e2e": e2e,
                "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "strong": strong, "gather": gather,
                "cfg3_sharded": cfg3,
                "series_per_gpu": int(series), "rendered_bytes_per_gpu": int(body_len),
                "kernel_ms": {k: v[0] / max(1, args.steps) for k, v in prof.items()}}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def label_bytes_of(g, table, series):
    """Pre-escaped label bytes of the live series (read once per family by the render kernel): exact, from the
    library's bulk read-back of the label text (kgx_read_labels)."""
    _, off = g.read_labels()
    return int(off[-1])


if __name__ == "__main__":
    main()
