"""Multi-GPU host logic of the hot path: one process per GPU, the fleet sharded
by node (pods, PIDs and GPUs are node-local, so shards never exchange samples
or series).  The only cross-rank step is the optional ORDERED CONCATENATION of
the per-shard /metrics bodies on one rank: lengths are all-gathered, offsets
are their exclusive scan, and every rank's bytes travel once, point to point
(NCCL send/recv over NVLink on GPUs; gloo on CPU in the tests).  Nothing is
reduced: there is no reduction collective on this path.

The reference has no counterpart (one exporter per node, main.go is a single
process); this is SURVEY.md section 8(e)."""
import numpy as np
import torch
import torch.distributed as dist


def node_range(rank, world, n_nodes):
    """Contiguous node range [lo, hi) of `rank` when `n_nodes` are split over `world` ranks."""
    base, extra = divmod(int(n_nodes), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def nccl_gather_env():
    """Call before the NCCL process group is created.  The ordered gather is a handful of large point-to-point
    transfers into one GPU; with NCCL's default channel count one peer delivers ~345 GB/s over NVLink 5, with 64
    channels 540 GB/s (measured at 2 and 4 GPUs, profiles/README.md).  Values already in the environment win."""
    import os
    for k, v in (("NCCL_MIN_P2P_NCHANNELS", "64"), ("NCCL_MAX_P2P_NCHANNELS", "64"), ("NCCL_MAX_NCHANNELS", "64")):
        os.environ.setdefault(k, v)


def gather_layout(local_len, device=None, group=None):
    """All-gather the body lengths; returns (lengths[world], offsets[world+1])."""
    world = dist.get_world_size(group)
    mine = torch.tensor([int(local_len)], dtype=torch.int64, device=device)
    lens = [torch.zeros(1, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(lens, mine, group=group)
    lengths = np.array([int(t.item()) for t in lens], dtype=np.int64)
    offsets = np.zeros(world + 1, dtype=np.int64)
    offsets[1:] = np.cumsum(lengths)
    return lengths, offsets


def gather_bodies(local, dst=0, out=None, group=None):
    """Ordered concatenation of every rank's body (a 1-D uint8 tensor) on `dst`.

    Returns (tensor, offsets) on dst and (None, offsets) elsewhere.  `out`, if
    given on dst, is a preallocated uint8 tensor of at least the total size."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    lengths, offsets = gather_layout(local.numel(), device=local.device, group=group)
    # one batched group of point-to-point transfers: on NCCL they run concurrently over NVLink
    # (unbatched send/recv would be serialised one peer after the other)
    ops = []
    if rank == dst:
        total = int(offsets[-1])
        if out is None:
            out = torch.empty(total, dtype=torch.uint8, device=local.device)
        out[int(offsets[rank]):int(offsets[rank + 1])].copy_(local)
        ops = [dist.P2POp(dist.irecv, out[int(offsets[r]):int(offsets[r + 1])], r, group)
               for r in range(world) if r != dst and lengths[r] > 0]
    elif local.numel() > 0:
        ops = [dist.P2POp(dist.isend, local, dst, group)]
    if ops:
        for q in dist.batch_isend_irecv(ops):
            q.wait()
    if rank == dst:
        return out[:int(offsets[-1])], offsets
    return None, offsets


def max_over_ranks(value, device=None, group=None):
    """Device-timed durations are reported as the max over ranks."""
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def body_checksum(t, chunk=1 << 26):
    """64-bit position-weighted checksum of a 1-D uint8 tensor (any device): sum of byte[i] * (1 + i mod 65521) in
    wrapping int64 arithmetic, computed where the bytes live.  Two equal checksums over a few GB of text mean the
    bytes arrived in place; a shifted, truncated or reordered body changes it."""
    total = 0
    n = t.numel()
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        w = (torch.arange(lo, hi, device=t.device, dtype=torch.int64) % 65521) + 1
        total = (total + int((t[lo:hi].to(torch.int64) * w).sum().item())) & 0xFFFFFFFFFFFFFFFF
    return total


def gather_checksums(local, group=None):
    """Every rank's body checksum, on every rank (one small all-gather next to the lengths)."""
    world = dist.get_world_size(group)
    c = body_checksum(local)
    mine = torch.tensor([c - (1 << 64) if c >= (1 << 63) else c], dtype=torch.int64, device=local.device)
    got = [torch.zeros(1, dtype=torch.int64, device=local.device) for _ in range(world)]
    dist.all_gather(got, mine, group=group)
    return [int(x.item()) & 0xFFFFFFFFFFFFFFFF for x in got]


def verify_gather(out, offsets, checksums):
    """On the destination rank: each rank's slice of the concatenation carries that rank's own checksum."""
    return all(body_checksum(out[int(offsets[r]):int(offsets[r + 1])]) == checksums[r] for r in range(len(checksums)))


class DeviceBytes:
    """Zero-copy view of a raw device pointer (e.g. kgx_body_device) for torch.as_tensor."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False),
                                         "version": 2, "strides": None}
