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
#include <type_traits>

namespace kgx {

constexpr int kIngestThreads = 512;

struct DevNodeMeta {
  const uint32_t* node_a_base;  // X: [n_nodes+1] first sorted-podkey index of each node
  uint32_t smem_cells;          // table cells that fit the staging buffer
  uint32_t smem_pks;            // podkeys that fit the staging buffer
  const uint32_t* node_slot_base;  // [n_nodes+1] first slot of each node
  uint32_t smem_words;          // series-index words (32 slots each) that fit the staging buffer
  const uint32_t* node_ser_off; // [n_nodes+1] first series rank of each node
  uint32_t reset;               // X, PASS 0: a window reset is pending — zero each node's series state before folding into it
  uint32_t parts;               // CTAs that share one node's rows (1 for fleets; > 1 only when nodes are fewer than CTAs)
};
constexpr uint32_t kMaxStagedGpus = 64;
constexpr uint32_t kNoRank = 0xffffffffu;

// X mode, one row into one series: the six integer reductions, lane-serial (fallback paths and repeated matches).
__device__ __forceinline__ void fold_x_scalar(const DevAcc& acc, uint32_t r, unsigned long long used, unsigned long long ts,
                                              uint32_t sm, uint32_t mem) {
  unsigned long long* xa = acc.xa + size_t(r) * kXAWords;
  unsigned long long* xb = acc.xb + size_t(r) * kXBWords;
  atomicAdd(&xa[XA_COUNT], 1ull);
  atomicAdd(&xa[XA_SUM_USED], used);
  atomicAdd(&xa[XA_SUM_SM], (unsigned long long)sm);
  atomicAdd(&xa[XA_SUM_MEM], (unsigned long long)mem);
  atomicMax(&xb[XB_MAX_USED], used);
  atomicMax(&xb[XB_LAST_TS], ts);
}

// X mode, the rows of a whole warp (lane's row -> series `r`, kNoRank: nothing to fold).  A series' state is
// xa[r] = {count, sum_used, sum_sm, sum_mem} and xb[r] = {max_used, last_ts}, so the four sums of ONE row are four
// adjacent words of one 32-byte sector and its two maxima two adjacent words: four lanes (two lanes) take one row and
// reduce it with a single instruction that touches a single sector.  Per 32 rows that is 4 + 2
// instructions and 64 memory requests instead of 6 instructions and 192 scattered ones — the join is bound by the
// number of reduction requests, not by their width (tools/micro/red_sector.cu).  Call with all 32 lanes.
__device__ __forceinline__ void fold_x_warp(const DevAcc& acc, uint32_t r, unsigned long long used, unsigned long long ts,
                                            uint32_t sm, uint32_t mem) {
  constexpr uint32_t full = 0xffffffffu;
  const uint32_t lane = threadIdx.x & 31;
  if (__all_sync(full, r == kNoRank)) return;
  const uint32_t u_lo = uint32_t(used), u_hi = uint32_t(used >> 32), t_lo = uint32_t(ts), t_hi = uint32_t(ts >> 32);
#pragma unroll
  for (int g = 0; g < 4; ++g) {  // sums: lanes 4k..4k+3 take row 8g+k
    const int src = g * 8 + int(lane >> 2);
    const uint32_t rr = __shfl_sync(full, r, src);
    const uint32_t a = __shfl_sync(full, u_lo, src), b = __shfl_sync(full, u_hi, src);
    const uint32_t c = __shfl_sync(full, sm, src), d = __shfl_sync(full, mem, src);
    const uint32_t f = lane & 3;
    const unsigned long long v = f == 0 ? 1ull : f == 1 ? ((unsigned long long)b << 32 | a) : f == 2 ? (unsigned long long)c : (unsigned long long)d;
    if (rr != kNoRank) atomicAdd(acc.xa + size_t(rr) * kXAWords + f, v);
  }
#pragma unroll
  for (int g = 0; g < 2; ++g) {  // maxima: lanes 2k, 2k+1 take row 16g+k
    const int src = g * 16 + int(lane >> 1);
    const uint32_t rr = __shfl_sync(full, r, src);
    const uint32_t a = __shfl_sync(full, u_lo, src), b = __shfl_sync(full, u_hi, src);
    const uint32_t c = __shfl_sync(full, t_lo, src), d = __shfl_sync(full, t_hi, src);
    const uint32_t f = lane & 1;
    const unsigned long long v = f ? ((unsigned long long)d << 32 | c) : ((unsigned long long)b << 32 | a);
    if (rr != kNoRank) atomicMax(acc.xb + size_t(rr) * kXBWords + f, v);
  }
}


template <bool XMODE, int PASS>
__global__ void __launch_bounds__(kIngestThreads) k_ingest_nodes(DevTable t, DevNodeMeta nm, DevIndex ix, DevAcc acc, DevCols c,
                                                                 uint32_t seq, unsigned long long* counters) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) unsigned long long s_bar;
  const uint32_t tid = threadIdx.x;
  const uint32_t bar = uint32_t(__cvta_generic_to_shared(&s_bar));
  Entry* s_tab = reinterpret_cast<Entry*>(smem);
  uint2* s_pk = reinterpret_cast<uint2*>(smem + size_t(nm.smem_cells) * sizeof(Entry));  // {first slot, stride} per podkey
  uint2* s_bm = s_pk + nm.smem_pks;  // the node's slice of the series index
  __shared__ uint32_t s_grank[kMaxStagedGpus];
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  uint32_t parity = 0;
  unsigned long long matched = 0, dropped = 0;
  bool any_new = false;
  // A fleet has more nodes than the GPU has CTAs; a single busy node (BASELINE configs 1-2) is split into nm.parts row
  // ranges that different CTAs take, each staging the node's table for itself.
  for (uint32_t vb = blockIdx.x; vb < t.n_nodes * nm.parts; vb += gridDim.x) {
    const uint32_t node = vb / nm.parts, part = vb - node * nm.parts;
    unsigned long long lo = c.node_off[node], hi = c.node_off[node + 1];
    if (nm.parts > 1) {
      const unsigned long long chunk = (hi - lo + nm.parts - 1) / nm.parts;
      lo = min(hi, lo + part * chunk);
      hi = min(hi, lo + chunk);
    }
    if (XMODE && PASS == 0 && nm.reset) {  // (only with parts == 1: the CTA owns the node)
      // kgx_reset_window folded into the join: the CTA that owns the node clears its series first.  The zeros are still
      // in L2 when the reductions arrive, instead of a separate pass writing them out and the join fetching them back.
      const uint32_t s0 = nm.node_ser_off[node], s1 = nm.node_ser_off[node + 1];
      ulonglong2* a2 = reinterpret_cast<ulonglong2*>(acc.xa + size_t(s0) * kXAWords);
      ulonglong2* b2 = reinterpret_cast<ulonglong2*>(acc.xb + size_t(s0) * kXBWords);
      for (uint32_t k = tid; k < (s1 - s0) * (kXAWords / 2); k += kIngestThreads) a2[k] = make_ulonglong2(0, 0);
      for (uint32_t k = tid; k < (s1 - s0) * (kXBWords / 2); k += kIngestThreads) b2[k] = make_ulonglong2(0, 0);
      if (tid == 0) acc.node_live[node] = 0;
      __syncthreads();
    }
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
        for (uint32_t i = tid; i < K; i += kIngestThreads) s_pk[i] = make_uint2(t.pk_slot0[a0 + i], t.pk_stride[a0 + i]);
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
    // The trip count is the same for all lanes of a warp (the fold below shuffles): lanes past the end idle.
    // Two copies of the loop, table in shared or in global memory: the choice is made once per node, not per access.
    auto rows = [&](auto staged_c) {
    constexpr bool STAGED = decltype(staged_c)::value;
    Row nxt{};
    if (lo + tid < hi) nxt = load_row(lo + tid);
    for (unsigned long long i = lo + tid; i - (tid & 31) < hi; i += kIngestThreads) {
      const bool valid = i < hi;
      const Row row = nxt;
      if (i + kIngestThreads < hi) nxt = load_row(i + kIngestThreads);
      const uint32_t gpu = row.gpu;
      uint32_t r_first = kNoRank;  // X mode: the series my row folds into (a second match of the same row is rare)
      if (valid && gpu >= t.G) ++dropped;
      if (valid && gpu < t.G) {
        const uint32_t pid = row.pid;
        uint32_t pos = uint32_t((uint64_t(dev_hash_pid(pid)) * nt.cap) >> 32);
        bool hit = false;
        for (;;) {
          const Entry e = STAGED ? s_tab[pos] : t.tab[nt.off + pos];
          if (e.val == kEmpty) break;
          if (e.key == pid) {
            hit = true;
            const uint32_t target = e.val & ~kMore;
            uint32_t slot;
            if (XMODE) {
              if (STAGED) { const uint2 pk = s_pk[target - a0]; slot = pk.x + s_grank[gpu] * pk.y; }
              else slot = t.pk_slot0[target] + t.grank[node * t.G + gpu] * t.pk_stride[target];
            } else {
              slot = target;
            }
            const uint32_t w = slot >> 5, bit = 1u << (slot & 31);
            const uint2 bw = STAGED ? s_bm[w - w0] : ix.bm[w];
            bool fold;
            if (PASS == 0) {
              fold = (bw.x & bit) != 0;
              if (!fold) { atomicOr(&ix.pending[w], bit); any_new = true; }
            } else {
              fold = (ix.pending[w] & bit) != 0;
            }
            if (fold) {
              const uint32_t r = bw.y + __popc(bw.x & (bit - 1));
              if (XMODE) {
                if (!marked_live) { acc.node_live[node] = 1; marked_live = true; }  // only "non-zero" matters: plain store
                if (r_first == kNoRank) r_first = r;
                else fold_x_scalar(acc, r, row.used, row.ts, row.sm, row.mem);  // the PID sits in several pods' lists
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
      if (XMODE) fold_x_warp(acc, r_first, row.used, row.ts, row.sm, row.mem);
    }
    };
    if (staged) rows(std::true_type{}); else rows(std::false_type{});
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
