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


def ncu_traffic(kernel_csv):
    """dram read+write bytes per launch of the dominant kernel, from the committed
    `ncu --set full` raw page (profiles/); None if the capture is not there."""
    import csv
    path = os.path.join(ROOT, "profiles", kernel_csv)
    if not os.path.exists(path):
        return None
    rows = list(csv.reader(open(path)))
    hdr, units, vals = rows[0], rows[1], rows[-1]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    total = 0.0
    for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        if key not in hdr:
            return None
        i = hdr.index(key)
        total += float(vals[i].replace(",", "")) * scale.get(units[i], 1.0)
    return total


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
        "config": workload_config(CONFIGS[CFG]["n_nodes"]),
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
    ap.add_argument("--gather", action="store_true",
                    help="N>1: also time the ordered concatenation of all bodies on rank 0 (NCCL send/recv over NVLink)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "kgx" else args.warmup
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from kgx_b200 import kgx

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    fleet, table, cols = build_fleet(rank, args.nodes)
    stream = torch.cuda.Stream(device=dev)
    g = kgx.Kgx(MODE_X, device=local, stream=stream.cuda_stream, profile=True)
    g.set_table(table)

    # ---- inputs resident in HBM (torch owns the device buffers) ----------------
    def dev_t(a):
        return None if a is None else torch.from_numpy(a.view(np.int64) if a.dtype == np.uint64 else a.view(np.int32)).to(dev)

    dcols = {k: dev_t(getattr(cols, k)) for k in ("pid", "gpu", "used", "node", "sm_util", "mem_util", "ts_us", "node_off", "gpu_total")}
    ds = abi.kgx_cols()
    ds.n = cols.n
    for k, t in dcols.items():
        if t is not None:
            setattr(ds, k, C.cast(t.data_ptr(), abi.u64p if t.dtype == torch.int64 else abi.u32p))

    def step_device():
        g.reset_window()
        g.ingest_device(ds)
        return g.render_device()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            body_len = step_device()
        g.sync()
        g.profile_reset()
        launches0 = g.stats()["kernel_launches"]
        clocks = ClockSampler(local)
        clocks.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(args.steps):
            body_len = step_device()
        e1.record(stream)
        barrier()
        clk = clocks.stop()
    ms = e0.elapsed_time(e1) / args.steps
    prof = g.profile()
    st = g.stats()
    launches = st["kernel_launches"] - launches0
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
    ingest_ms = prof["ingest"][0] / max(1, prof["ingest"][1])
    series = st["n_series"]
    label_bytes = label_bytes_of(g, table, series)
    alg_render = n_fam * label_bytes + body_len
    achieved = alg_render / (render_ms * 1e-3) / 1e9 if render_ms > 0 else 0.0
    alg_pipeline = cols.n * 36 + len(table.entry_pid) * 12 + alg_render
    roofline = {"bound": "hbm", "kernel": "k_render3<64> (exposition writer, kgx_writer.cuh)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": ncu_traffic("r01_k_render3_raw.csv"), "peak_source": peak_src,
                "traffic_source": "profiles/r01_k_render3_raw.csv (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, one launch)",
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
        e2e = {"value": total_samples / (float(tm.item()) * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(n_out), "ms_per_step": float(tm.item()), "host_memory": "pinned (kgx_host_alloc); " + numa_note.note,
               "api": "kgx_reset_window + kgx_ingest + kgx_render_begin/kgx_render_end (body k-1 leaves while scrape k comes in)"}

    # ---- optional: ordered concatenation of the per-GPU bodies on rank 0 ------------------
    gather = None
    if args.gather and world > 1:
        from kgx_b200 import fleetshard
        g.render_device()
        ptr_, n_ = g.body_device()
        local = torch.as_tensor(fleetshard.DeviceBytes(ptr_, n_), device=dev)
        _, offs = fleetshard.gather_layout(n_, device=dev)
        root_out = torch.empty(int(offs[-1]), dtype=torch.uint8, device=dev) if rank == 0 else None
        for _ in range(2):
            fleetshard.gather_bodies(local, dst=0, out=root_out)
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for _ in range(5):
            fleetshard.gather_bodies(local, dst=0, out=root_out)
        g1.record()
        barrier()
        gms = fleetshard.max_over_ranks(g0.elapsed_time(g1) / 5, device=dev)
        gather = {"ms": gms, "bytes_on_root": int(offs[-1]), "inbound_gbs": (int(offs[-1]) - n_) / (gms * 1e-3) / 1e9 if rank == 0 else None,
                  "how": "lengths all-gather + NCCL send/recv into rank 0's buffer, bodies stay in node order"}

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
                "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "gather": gather,
                "series_per_gpu": int(series), "rendered_bytes_per_gpu": int(body_len),
                "kernel_ms": {k: v[0] / max(1, args.steps) for k, v in prof.items()}}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def label_bytes_of(g, table, series):
    """Pre-escaped label bytes of the live series (read once per family by the
    render kernel): taken from the library's own slot texts."""
    vals, slots = g.read_values(0, with_slots=True)
    if len(slots) == 0:
        return 0
    if len(slots) > 20000:  # sample: label lengths are near-uniform in the synthetic fleet
        idx = np.linspace(0, len(slots) - 1, 20000).astype(np.int64)
        return int(np.mean([len(g.slot_labels(int(slots[i]))) for i in idx]) * len(slots))
    return int(sum(len(g.slot_labels(int(s))) for s in slots))


if __name__ == "__main__":
    main()
