// kgx_ingest2.cuh — node-grouped hash join + reduce (sm_100a).
//
// Same results as k_ingest (the loop of main.go:141-154: join, values, upsert),
// for batches whose rows are grouped by node (kgx_cols.node_off).  Pods, PIDs
// and GPUs are node-local, so one CTA takes one node at a time and
//   * stages that node's slice of the open-addressing table in shared memory
//     with one TMA bulk copy (cp.async.bulk + mbarrier) — the table is read
//     from HBM exactly once, contiguously, instead of one random 32-byte
//     sector per probe;
//   * probes in shared memory;
//   * folds into the node's series, which are contiguous in rank order, so the
//     integer atomics of one node stay inside a few L2-resident kilobytes.
// Nodes whose slice does not fit the shared-memory budget are probed in global
// memory by the same kernel (still node by node, for the L2 locality).
#pragma once
#include "kgx_kernels.cuh"
#include "kgx_async.cuh"

namespace kgx {

constexpr int kIngestThreads = 512;

struct DevNodeMeta {
  const uint32_t* node_a_base;  // X: [n_nodes+1] first sorted-podkey index of each node
  uint32_t smem_cells;          // table cells that fit the staging buffer
  uint32_t smem_pks;            // podkeys that fit the staging buffer
  const uint32_t* node_slot_base;  // [n_nodes+1] first slot of each node
  uint32_t smem_words;          // series-index words (32 slots each) that fit the staging buffer
};
constexpr uint32_t kMaxStagedGpus = 64;

template <bool XMODE, int PASS>
__global__ void __launch_bounds__(kIngestThreads) k_ingest_nodes(DevTable t, DevNodeMeta nm, DevIndex ix, DevAcc acc, DevCols c,
                                                                 uint32_t seq, unsigned long long* counters) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) unsigned long long s_bar;
  const uint32_t tid = threadIdx.x;
  const uint32_t bar = uint32_t(__cvta_generic_to_shared(&s_bar));
  Entry* s_tab = reinterpret_cast<Entry*>(smem);
  uint32_t* s_slot0 = reinterpret_cast<uint32_t*>(smem + size_t(nm.smem_cells) * sizeof(Entry));
  uint32_t* s_stride = s_slot0 + nm.smem_pks;
  uint2* s_bm = reinterpret_cast<uint2*>(s_stride + nm.smem_pks);  // the node's slice of the series index
  __shared__ uint32_t s_grank[kMaxStagedGpus];
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  uint32_t parity = 0;
  unsigned long long matched = 0, dropped = 0;
  bool any_new = false;
  for (uint32_t node = blockIdx.x; node < t.n_nodes; node += gridDim.x) {
    const unsigned long long lo = c.node_off[node], hi = c.node_off[node + 1];
    if (hi <= lo) continue;  // uniform over the CTA
    const NodeTab nt = t.node_tab[node];
    uint32_t a0 = 0, K = 0;
    if (XMODE) { a0 = nm.node_a_base[node]; K = nm.node_a_base[node + 1] - a0; }
    const uint32_t w0 = nm.node_slot_base[node] >> 5, w1 = (nm.node_slot_base[node + 1] + 31) >> 5;
    const bool staged = nt.cap <= nm.smem_cells && K <= nm.smem_pks && (w1 - w0) <= nm.smem_words && t.G <= kMaxStagedGpus;
    if (staged) {
      if (tid == 0) {  // the node's table slice: one contiguous, 64-byte aligned range
        mbar_expect_tx(bar, nt.cap * uint32_t(sizeof(Entry)));
        bulk_g2s(uint32_t(__cvta_generic_to_shared(s_tab)), t.tab + nt.off, nt.cap * uint32_t(sizeof(Entry)), bar);
      }
      if (XMODE)
        for (uint32_t i = tid; i < K; i += kIngestThreads) { s_slot0[i] = t.pk_slot0[a0 + i]; s_stride[i] = t.pk_stride[a0 + i]; }
      for (uint32_t i = tid; i < w1 - w0; i += kIngestThreads) s_bm[i] = ix.bm[w0 + i];
      if (XMODE && tid < t.G) s_grank[tid] = t.grank[node * t.G + tid];
      mbar_wait(bar, parity);
      parity ^= 1;
      __syncthreads();
    }
    bool marked_live = false;
    // the columns of my next row are fetched while the current one is probed and folded
    struct Row { uint32_t pid, gpu, sm, mem; unsigned long long used, ts; };
    auto load_row = [&](unsigned long long i) {
      Row w{};
      w.pid = c.pid[i]; w.gpu = c.gpu[i];
      if (XMODE) { w.used = c.used[i]; w.ts = c.ts_us[i]; w.sm = c.sm_util[i]; w.mem = c.mem_util[i]; }
      return w;
    };
    Row nxt{};
    if (lo + tid < hi) nxt = load_row(lo + tid);
    for (unsigned long long i = lo + tid; i < hi; i += kIngestThreads) {
      const Row row = nxt;
      if (i + kIngestThreads < hi) nxt = load_row(i + kIngestThreads);
      const uint32_t gpu = row.gpu;
      if (gpu >= t.G) { ++dropped; continue; }
      const uint32_t pid = row.pid;
      uint32_t pos = uint32_t((uint64_t(dev_hash_pid(pid)) * nt.cap) >> 32);
      bool hit = false;
      for (;;) {
        const Entry e = staged ? s_tab[pos] : t.tab[nt.off + pos];
        if (e.val == kEmpty) break;
        if (e.key == pid) {
          hit = true;
          const uint32_t target = e.val & ~kMore;
          uint32_t slot;
          if (XMODE) {
            const uint32_t gr = staged ? s_grank[gpu] : t.grank[node * t.G + gpu];
            slot = staged ? s_slot0[target - a0] + gr * s_stride[target - a0] : t.pk_slot0[target] + gr * t.pk_stride[target];
          } else {
            slot = target;
          }
          const uint32_t w = slot >> 5, bit = 1u << (slot & 31);
          const uint2 b = staged ? s_bm[w - w0] : ix.bm[w];
          bool fold;
          if (PASS == 0) {
            fold = (b.x & bit) != 0;
            if (!fold) { atomicOr(&ix.pending[w], bit); any_new = true; }
          } else {
            fold = (ix.pending[w] & bit) != 0;
          }
          if (fold) {
            const uint32_t r = b.y + __popc(b.x & (bit - 1));
            if (XMODE) {
              atomicAdd(&acc.count[r], 1ull);
              if (!marked_live) { acc.node_live[node] = 1; marked_live = true; }  // only "non-zero" matters: plain store
              atomicAdd(&acc.sum_used[r], row.used);
              atomicMax(&acc.max_used[r], row.used);
              atomicMax(&acc.last_ts[r], row.ts);
              atomicAdd(&acc.sum_sm[r], (unsigned long long)row.sm);
              atomicAdd(&acc.sum_mem[r], (unsigned long long)row.mem);
            } else {
              atomicMax(&acc.r_seq[r], (uint64_t(seq) << 32) | uint64_t(i));  // last write in sample order wins
            }
          }
          if (!(e.val & kMore)) break;
        }
        pos = pos + 1 == nt.cap ? 0 : pos + 1;
      }
      if (hit) ++matched; else ++dropped;
    }
    if (staged) __syncthreads();  // the staging buffer is rewritten for the next node
  }
  if (PASS == 0) {
    matched = __reduce_add_sync(0xffffffffu, uint32_t(matched));
    dropped = __reduce_add_sync(0xffffffffu, uint32_t(dropped));
    const bool warp_new = __any_sync(0xffffffffu, any_new);
    if ((threadIdx.x & 31) == 0) {
      if (matched) atomicAdd(&counters[C_MATCHED], matched);
      if (dropped) atomicAdd(&counters[C_DROPPED], dropped);
      if (warp_new) atomicAdd(&counters[C_NEW], 1ull);
    }
  }
}

}  // namespace kgx
