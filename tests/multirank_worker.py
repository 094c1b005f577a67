"""One rank of the multi-GPU parity check (launched by tests/test_gpu_multirank.py under torch.distributed.run,
one process per GPU, NCCL).  A fleet is split by node over the ranks; every rank joins, reduces and renders ITS nodes
on its own GPU through the C ABI; the bodies are concatenated in node order on rank 0 over NVLink; rank 0 then
compares the concatenation with the CPU oracle's body of the WHOLE fleet, byte for byte, and the per-rank checksums.
Covers BASELINE config 3 (rack replay: 128 nodes, 1 M samples, R-intended) and a 96-node slice of config 4 (X mode)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kgx_loader  # noqa: E402

kgx_loader.load()
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from kgx_b200 import fleetshard, kgx  # noqa: E402
from kgx_b200.abi import MODE_R_INTENDED, MODE_X  # noqa: E402
from kgx_b200.fleet import Fleet  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    fleetshard.nccl_gather_env()
    dist.init_process_group("nccl", device_id=dev)
    ok = True
    for cfg, mode, n_total in ((3, MODE_R_INTENDED, 128), (4, MODE_X, 96)):
        lo, hi = fleetshard.node_range(rank, world, n_total)
        f = Fleet.config(cfg, n_nodes=hi - lo, node_base=lo)
        g = kgx.Kgx(mode, device=local)
        g.set_table(f.table(mode))
        g.ingest(f.scrape(0))
        n = g.render_device()
        p, n = g.body_device()
        body = torch.as_tensor(fleetshard.DeviceBytes(p, n), device=dev)
        out, offs = fleetshard.gather_bodies(body, dst=0)
        sums = fleetshard.gather_checksums(body)
        node_offs = [None] * world
        dist.all_gather_object(node_offs, [int(x) for x in g.body_offsets()])
        if rank == 0:
            kgo = kgx_loader.load_oracle()
            whole = Fleet.config(cfg, n_nodes=n_total)
            o = kgo.Oracle(mode, kgo.JOIN_HASH, min(16, os.cpu_count() or 1))
            o.set_table(whole.table(mode))
            o.ingest(whole.scrape(0))
            want = o.render()
            got = out.cpu().numpy().tobytes()
            # per-node offsets of the concatenation = each rank's offsets shifted by where its body starts
            cat = [int(offs[r]) + x for r in range(world) for x in node_offs[r][:-1]] + [int(offs[-1])]
            checks = [got == want, fleetshard.verify_gather(out, offs, sums), cat == [int(x) for x in o.body_offsets()],
                      len(got) > 1_000_000]
            print("cfg%d mode %d: %d bytes on rank 0 from %d ranks; body==oracle %s, checksums %s, node offsets %s" % (
                cfg, mode, len(got), world, checks[0], checks[1], checks[2]), flush=True)
            ok = ok and all(checks)
        g.close()
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    if rank == 0:
        print("MULTIRANK PARITY %s" % ("OK" if ok else "FAILED"), flush=True)
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
