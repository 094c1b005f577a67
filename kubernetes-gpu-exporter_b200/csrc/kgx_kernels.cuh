// kgx_kernels.cuh — sm_100a kernels of the join -> reduce -> render path.
//
//   k_ingest      hash join (open addressing, main.go:141-145) fused with the
//                 per-series integer reduction (main.go:147-150's upsert; X
//                 mode: sum / max / last-timestamp / count)
//   k_finalize_r  R modes: fetch the values of each series' last writer
//   k_render      exposition writer (main.go:69-71: Gather + expfmt): exact
//                 line lengths -> block scan -> inter-tile decoupled look-back
//                 -> byte scatter, all in one pass over the series
//   index kernels series-index rebuild on churn (cold path)
#pragma once
#include <cstdint>
#include "kgx_fmt.cuh"
#include "kgx_table.hpp"

namespace kgx {

// ---- device views ------------------------------------------------------------
struct DevTable {
  const NodeTab* node_tab;
  const Entry* tab;
  const uint32_t* pk_slot0;   // X
  const uint32_t* pk_stride;  // X
  const uint32_t* grank;      // X
  uint32_t n_nodes, G;
};

struct DevIndex {       // which potential series (slots) exist, and their rank
  uint2* bm;            // [n_words] {bits, exclusive prefix popcount}
  uint32_t* pending;    // [n_words] slots seen for the first time in this ingest
  uint32_t n_words;
};

// X mode: the state of a series is two small records, indexed by rank: xa = {count, sum_used, sum_sm, sum_mem} — the
// four sums of one row land in ONE 32-byte sector — and xb = {max_used, last_ts}, 16 bytes (see fold_x_warp in
// kgx_ingest2.cuh).  One 64-byte record per series would serve the join as well, but the writer reads one or two fields
// of 64 consecutive series per tile and pays per 128-byte line its lanes touch: 32-byte and 16-byte strides halve and
// quarter that (measured: writer 1.43 -> 1.51 ms with the 64-byte record).
enum { XA_COUNT = 0, XA_SUM_USED = 1, XA_SUM_SM = 2, XA_SUM_MEM = 3, kXAWords = 4, XB_MAX_USED = 0, XB_LAST_TS = 1, kXBWords = 2 };

struct DevAcc {  // per-series state, indexed by rank
  unsigned long long *xa, *xb;                                                   // X: [n_series][kXAWords], [n_series][kXBWords]
  unsigned long long *r_seq, *r_used, *r_total;                                  // R
  uint32_t* node_live;  // [n_nodes] series with a value in the current window
};

struct DevCols {
  const uint32_t *pid, *gpu, *node, *sm_util, *mem_util;
  const unsigned long long *used, *ts_us, *node_off;
  uint64_t n;
};

// counters[] (device): 0 matched samples, 1 dropped samples, 2 new slots seen — accumulated by one join launch and
// handed over by its last CTA (posted to mapped host memory, C_CHURN := new slots, then zeroed for the next launch);
// 3 total body length, 4 and 5 the two tile tickets (see ticket_of);
// 6 C_CHURN: new slots seen by the latest join (the writer refuses to run on a stale series index).
// Steady state therefore needs no memset around either hot kernel.
enum { C_MATCHED = 0, C_DROPPED = 1, C_NEW = 2, C_TOTAL_LEN = 3, C_TICKET = 4, C_TICKET_B = 5, C_CHURN = 6, C_SPARE = 7, C_COUNT = 8 };

// Hand-over of the join statistics, a one-thread kernel queued right behind a PASS 0 join: posts them to mapped host
// memory, records whether new series were seen (C_CHURN) and zeroes the accumulators for the next launch.  (Doing this
// in the join's last CTA needs a __threadfence per CTA, which waits for that CTA's reductions to drain: 0.41 -> 0.69 ms.)
__global__ void k_ingest_epilogue(unsigned long long* counters, unsigned long long* host_post) {
  if (threadIdx.x == 0) {
    host_post[C_MATCHED] = counters[C_MATCHED]; host_post[C_DROPPED] = counters[C_DROPPED]; host_post[C_NEW] = counters[C_NEW];
    counters[C_CHURN] = counters[C_NEW];
    counters[C_MATCHED] = 0; counters[C_DROPPED] = 0; counters[C_NEW] = 0;
  }
}

__device__ __forceinline__ uint32_t dev_hash_pid(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}

// ---- ingest: join + reduce ---------------------------------------------------
// PASS 0: normal pass.  Matches whose slot is not in the series index yet only
//         mark it pending (the host then rebuilds the index and runs PASS 1).
// PASS 1: churn pass, after the rebuild: folds exactly the matches PASS 0 had
//         to skip (slot pending).
template <bool XMODE, int PASS>
__global__ void __launch_bounds__(256) k_ingest(DevTable t, DevIndex ix, DevAcc acc, DevCols c, uint32_t seq,
                                                 unsigned long long* counters) {
  unsigned long long matched = 0, dropped = 0;
  bool any_new = false;
  for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < c.n; i += uint64_t(gridDim.x) * blockDim.x) {
    uint32_t node = 0;
    if (c.node) node = c.node[i];
    else if (c.node_off) {  // rows grouped by node: find the group
      uint32_t lo = 0, hi = t.n_nodes;
      while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (c.node_off[mid] <= i) lo = mid; else hi = mid; }
      node = lo;
    }
    const uint32_t gpu = c.gpu[i];
    if (node >= t.n_nodes || gpu >= t.G) { ++dropped; continue; }
    const uint32_t pid = c.pid[i];
    const NodeTab nt = t.node_tab[node];
    uint32_t pos = uint32_t((uint64_t(dev_hash_pid(pid)) * nt.cap) >> 32);
    bool hit = false;
    for (;;) {
      const Entry e = t.tab[nt.off + pos];
      if (e.val == kEmpty) break;
      if (e.key == pid) {
        hit = true;
        const uint32_t target = e.val & ~kMore;
        uint32_t slot;
        if (XMODE) slot = t.pk_slot0[target] + t.grank[node * t.G + gpu] * t.pk_stride[target];
        else slot = target;
        const uint32_t w = slot >> 5, bit = 1u << (slot & 31);
        const uint2 b = ix.bm[w];
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
            unsigned long long* xa = acc.xa + size_t(r) * kXAWords;
            unsigned long long* xb = acc.xb + size_t(r) * kXBWords;
            const unsigned long long used = c.used[i];
            if (atomicAdd(&xa[XA_COUNT], 1ull) == 0) atomicAdd(&acc.node_live[node], 1u);
            atomicAdd(&xa[XA_SUM_USED], used);
            atomicMax(&xb[XB_MAX_USED], used);
            atomicMax(&xb[XB_LAST_TS], c.ts_us[i]);
            atomicAdd(&xa[XA_SUM_SM], (unsigned long long)c.sm_util[i]);
            atomicAdd(&xa[XA_SUM_MEM], (unsigned long long)c.mem_util[i]);
          } else {
            // GaugeVec.Set: the last write in sample order wins (main.go:147,150)
            atomicMax(&acc.r_seq[r], (uint64_t(seq) << 32) | uint64_t(i));
          }
        }
        if (!(e.val & kMore)) break;
      }
      pos = pos + 1 == nt.cap ? 0 : pos + 1;
    }
    if (hit) ++matched; else ++dropped;
  }
  if (PASS == 0) {
    // block-level reduction of the statistics
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


// an empty batch still has to report "0 matched, 0 dropped, no new series"
__global__ void k_ingest_nothing(unsigned long long* counters, unsigned long long* host_post) {
  if (threadIdx.x == 0) { host_post[C_MATCHED] = 0; host_post[C_DROPPED] = 0; host_post[C_NEW] = 0; counters[C_CHURN] = 0; }
}

// X mode: start a new window — every per-series accumulator and the per-node "has a value" flags to zero, one launch
__global__ void k_clear_window(DevAcc acc, uint32_t n_series, uint32_t n_nodes) {
  const uint32_t stride = gridDim.x * blockDim.x;
  ulonglong2* a2 = reinterpret_cast<ulonglong2*>(acc.xa);
  ulonglong2* b2 = reinterpret_cast<ulonglong2*>(acc.xb);
  for (size_t i = blockIdx.x * blockDim.x + threadIdx.x; i < size_t(n_series) * (kXAWords / 2); i += stride) a2[i] = make_ulonglong2(0, 0);
  for (size_t i = blockIdx.x * blockDim.x + threadIdx.x; i < size_t(n_series) * (kXBWords / 2); i += stride) b2[i] = make_ulonglong2(0, 0);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_nodes; i += stride) acc.node_live[i] = 0;
}

// R modes: after the join, a series whose last writer is in this batch takes
// that sample's values: float64(used) and the device's total (main.go:147,149).
__global__ void k_finalize_r(DevAcc acc, DevCols c, const unsigned long long* gpu_total, uint32_t G, uint32_t seq,
                             uint32_t n_series, uint32_t n_nodes, unsigned long long* counters, unsigned long long* host_post) {
  // runs right behind the join on the same stream: it also hands the join's statistics over (see k_ingest_epilogue),
  // which saves the R modes a launch per cycle
  if (counters && blockIdx.x == 0 && threadIdx.x == 0) {
    host_post[C_MATCHED] = counters[C_MATCHED]; host_post[C_DROPPED] = counters[C_DROPPED]; host_post[C_NEW] = counters[C_NEW];
    counters[C_CHURN] = counters[C_NEW];
    counters[C_MATCHED] = 0; counters[C_DROPPED] = 0; counters[C_NEW] = 0;
  }
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_series) return;
  const unsigned long long s = acc.r_seq[r];
  if (uint32_t(s >> 32) != seq) return;
  const uint64_t i = s & 0xffffffffull;
  uint32_t node = 0;
  if (c.node) node = c.node[i];
  else if (c.node_off) {
    uint32_t lo = 0, hi = n_nodes;
    while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (c.node_off[mid] <= i) lo = mid; else hi = mid; }
    node = lo;
  }
  acc.r_used[r] = c.used[i];
  acc.r_total[r] = gpu_total[node * G + c.gpu[i]];
}

// ---- series index rebuild (cold path, on churn) ----------------------------------
__global__ void k_merge_pending(uint2* bm, const uint32_t* pending, uint32_t n_words, uint32_t* popc_out) {
  const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= n_words) return;
  uint2 b = bm[w];
  b.x |= pending[w];
  bm[w] = b;
  popc_out[w] = __popc(b.x);
}

// generic 3-step exclusive scan (u32 in, u64 out), 1024 items per block
constexpr int kScanBlock = 256, kScanItems = 4;
__global__ void k_scan_reduce(const uint32_t* in, uint64_t n, unsigned long long* block_sums) {
  __shared__ unsigned long long sh[kScanBlock / 32];
  const uint64_t base = uint64_t(blockIdx.x) * kScanBlock * kScanItems;
  unsigned long long s = 0;
  for (int k = 0; k < kScanItems; ++k) {
    const uint64_t i = base + uint64_t(k) * kScanBlock + threadIdx.x;
    if (i < n) s += in[i];
  }
  for (int o = 16; o; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) { unsigned long long tot = 0; for (int k = 0; k < kScanBlock / 32; ++k) tot += sh[k]; block_sums[blockIdx.x] = tot; }
}
__global__ void k_scan_blocksums(unsigned long long* block_sums, uint32_t nb, unsigned long long* total) {
  // single block, serial over chunks of 1024
  __shared__ unsigned long long sh[1024];
  __shared__ unsigned long long carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < nb; base += 1024) {
    const uint32_t i = base + threadIdx.x;
    unsigned long long v = i < nb ? block_sums[i] : 0;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
      unsigned long long t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
      __syncthreads();
      sh[threadIdx.x] += t;
      __syncthreads();
    }
    if (i < nb) block_sums[i] = carry + sh[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 0) carry += sh[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}
// item layout inside a block: thread-major chunks of kScanItems consecutive items
__global__ void k_scan_apply(const uint32_t* in, uint64_t n, const unsigned long long* block_sums, unsigned long long* out64,
                             uint32_t* out32) {
  __shared__ unsigned long long sh[kScanBlock / 32];
  const uint64_t base = uint64_t(blockIdx.x) * kScanBlock * kScanItems + uint64_t(threadIdx.x) * kScanItems;
  uint32_t v[kScanItems];
  unsigned long long s = 0;
  for (int k = 0; k < kScanItems; ++k) { v[k] = base + k < n ? in[base + k] : 0; s += v[k]; }
  unsigned long long inc = s;
  for (int o = 1; o < 32; o <<= 1) { unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o); if ((threadIdx.x & 31) >= o) inc += t; }
  if ((threadIdx.x & 31) == 31) sh[threadIdx.x >> 5] = inc;
  __syncthreads();
  unsigned long long warp_off = 0;
  for (int k = 0; k < int(threadIdx.x >> 5); ++k) warp_off += sh[k];
  unsigned long long run = block_sums[blockIdx.x] + warp_off + inc - s;
  for (int k = 0; k < kScanItems; ++k) {
    if (base + k < n) { if (out64) out64[base + k] = run; if (out32) out32[base + k] = uint32_t(run); }
    run += v[k];
  }
}

__global__ void k_set_prefix(uint2* bm, const uint32_t* prefix, uint32_t n_words) {
  const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w < n_words) bm[w].y = prefix[w];
}

struct DevLabels {  // label pieces: text of slot = A[a] + G[gidx] + B[a]
  const uint8_t *A_blob, *B_blob, *G_blob;
  const unsigned long long *A_off, *B_off, *G_off;
  const Run* runs;
  uint32_t n_runs;
};

// one thread per bitmap word: emit the series of its set bits
__global__ void k_build_series(const uint2* bm, uint32_t n_words, DevLabels L, uint32_t* ser_slot, uint32_t* ser_a,
                               uint32_t* ser_gidx, uint32_t* ser_len) {
  const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= n_words) return;
  uint2 b = bm[w];
  uint32_t bits = b.x, r = b.y;
  while (bits) {
    const uint32_t slot = (w << 5) + (__ffs(bits) - 1);
    bits &= bits - 1;
    uint32_t lo = 0, hi = L.n_runs;  // last run with slot_start <= slot
    while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (L.runs[mid].slot_start <= slot) lo = mid; else hi = mid; }
    const Run run = L.runs[lo];
    const uint32_t a = run.a_first + (slot - run.slot_start);
    uint32_t len = uint32_t(L.A_off[a + 1] - L.A_off[a]) + uint32_t(L.B_off[a + 1] - L.B_off[a]);
    if (run.gidx != 0xFFFFFFFFu) len += uint32_t(L.G_off[run.gidx + 1] - L.G_off[run.gidx]);
    ser_slot[r] = slot; ser_a[r] = a; ser_gidx[r] = run.gidx; ser_len[r] = len;
    ++r;
  }
}

// one warp per series: concatenate its label pieces into the series label blob
__global__ void k_fill_labels(DevLabels L, const uint32_t* ser_a, const uint32_t* ser_gidx, const unsigned long long* lbl_off,
                              uint32_t n_series, uint8_t* blob) {
  const uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (r >= n_series) return;
  const uint32_t a = ser_a[r], g = ser_gidx[r];
  uint8_t* dst = blob + lbl_off[r];
  const uint32_t la = uint32_t(L.A_off[a + 1] - L.A_off[a]);
  for (uint32_t i = lane; i < la; i += 32) dst[i] = L.A_blob[L.A_off[a] + i];
  dst += la;
  if (g != 0xFFFFFFFFu) {
    const uint32_t lg = uint32_t(L.G_off[g + 1] - L.G_off[g]);
    for (uint32_t i = lane; i < lg; i += 32) dst[i] = L.G_blob[L.G_off[g] + i];
    dst += lg;
  }
  const uint32_t lb = uint32_t(L.B_off[a + 1] - L.B_off[a]);
  for (uint32_t i = lane; i < lb; i += 32) dst[i] = L.B_blob[L.B_off[a] + i];
}

__device__ __forceinline__ uint32_t rank_of(const uint2* bm, uint32_t slot) {
  const uint2 b = bm[slot >> 5];
  return b.y + __popc(b.x & ((1u << (slot & 31)) - 1));
}

// move per-series state from the old ranks to the new ones
__global__ void k_remap(const uint32_t* old_slot, uint32_t n_old, const uint2* bm, int n_arrays, int words,
                        const unsigned long long* const* old_arr, unsigned long long* const* new_arr) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_old) return;
  const uint32_t s = old_slot[r];
  if (!((bm[s >> 5].x >> (s & 31)) & 1u)) return;  // the series was expired (kgx_expire): its state is dropped
  const uint32_t nr = rank_of(bm, s);
  for (int k = 0; k < n_arrays; ++k)
    for (int w = 0; w < words; ++w) new_arr[k][size_t(nr) * words + w] = old_arr[k][size_t(r) * words + w];
}

// X mode, explicit expiry: take every series whose newest sample is older than `cutoff` out of the series index.
// out[0] counts the ones that were exported (had samples in the window), out[1] all of them.
__global__ void k_mark_expired(const uint32_t* ser_slot, uint32_t n_series, const unsigned long long* xa,
                               const unsigned long long* xb, unsigned long long cutoff, uint2* bm, unsigned long long* out) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_series || xb[size_t(r) * kXBWords + XB_LAST_TS] >= cutoff) return;
  const uint32_t s = ser_slot[r];
  atomicAnd(&bm[s >> 5].x, ~(1u << (s & 31)));
  if (xa[size_t(r) * kXAWords + XA_COUNT]) atomicAdd(&out[0], 1ull);
  atomicAdd(&out[1], 1ull);
}

__global__ void k_node_ranks(const uint2* bm, const uint32_t* node_slot_base, uint32_t n_nodes, uint32_t n_slots,
                             uint32_t n_series, uint32_t* node_ser_off) {
  const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n > n_nodes) return;
  const uint32_t s = node_slot_base[n];
  node_ser_off[n] = s >= n_slots ? n_series : rank_of(bm, s);
}

// R modes, table change: seed carried series (slot, used, total)
__global__ void k_seed_carried(const uint32_t* slots, const unsigned long long* used, const unsigned long long* total,
                               uint32_t n, const uint2* bm, DevAcc acc) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t r = rank_of(bm, slots[i]);
  acc.r_seq[r] = 1;  // older than any ingest (their seq starts at 1)
  acc.r_used[r] = used[i];
  acc.r_total[r] = total[i];
}
__global__ void k_set_bits(const uint32_t* slots, uint32_t n, uint32_t* pending) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) atomicOr(&pending[slots[i] >> 5], 1u << (slots[i] & 31));
}

// ---- render --------------------------------------------------------------------------
struct FamMeta {
  uint8_t name[64];
  uint8_t header[320];  // "# HELP <name> <help>\n# TYPE <name> gauge\n"
  uint32_t name_len, header_len;
  int32_t kind;
  int32_t pad;
};
constexpr int kMaxFamilies = 8;

struct __align__(16) Tile {
  uint32_t first;   // rank of the first series
  uint32_t node;
  uint16_t count;   // series in the tile
  uint8_t family;
  uint8_t flags;    // bit0: first tile of (node, family): carries the header
                    // bit1: first tile of the node: records the body offset
  uint32_t pad_;
  // byte range of the tile's label text in the series label blob (= lbl_off[first], lbl_off[first + count]): kept in
  // the descriptor so that the writer's "ticket -> descriptor -> label copy" chain is two memory round trips, not three
  unsigned long long lbl0, lbl1;
};
static_assert(sizeof(Tile) == 32, "Tile is two 16-byte loads");
__global__ void k_tile_labels(Tile* tiles, uint32_t n_tiles, const unsigned long long* lbl_off) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_tiles) return;
  tiles[t].lbl0 = lbl_off[tiles[t].first];
  tiles[t].lbl1 = lbl_off[tiles[t].first + tiles[t].count];
}

struct DevRender {
  const FamMeta* fam;  // [n_families], device global memory (one schema per ctx)
  uint32_t n_families;
  const Tile* tiles;
  uint32_t n_tiles;
  const unsigned long long* lbl_off;  // [n_series+1]
  const uint32_t* ser_len;            // [n_series] label bytes of each series
  const uint8_t* lbl_blob;
  const uint32_t* ser_gidx;
  const unsigned long long* gpu_total;  // X: [n_nodes*G]
  unsigned long long* tile_state;       // [n_tiles] look-back words
  unsigned long long* node_body_off;    // [n_nodes+1]
  uint8_t* out;
  uint32_t smem_bytes;                  // capacity of the tile buffer
  uint32_t epoch;                       // tag of this launch in the look-back words and body offsets (see kStEpoch)
  uint32_t check_churn;                 // 1: leave at once if counters[C_CHURN] != 0 (the host then rebuilds the index)
  unsigned long long* host_len;         // mapped host word that receives the body length (may be null)
};

// A look-back word: state (2 bits) | epoch (14 bits) | byte count (48 bits).  A word only counts as published if it
// carries the epoch of the running launch, so the array never has to be cleared between launches: what an earlier
// launch left behind reads as "not there yet".  The host bumps the epoch per launch and clears the array when it
// wraps (or whenever it clears anyway: epoch 0).  The per-node body offsets are tagged the same way.
constexpr unsigned long long kStAgg = 1ull << 62, kStPrefix = 2ull << 62, kStMask = 3ull << 62;
constexpr int kStEpochShift = 48;
constexpr uint32_t kStEpochMax = 0x3fff;
constexpr unsigned long long kStValue = (1ull << kStEpochShift) - 1;
__device__ __forceinline__ unsigned long long st_tag(uint32_t epoch) { return (unsigned long long)epoch << kStEpochShift; }
__device__ __forceinline__ bool st_published(unsigned long long s, uint32_t epoch) {
  return (s & kStMask) != 0 && uint32_t(s >> kStEpochShift & kStEpochMax) == epoch;
}

// Two tile tickets take turns, picked by the parity of the launch's epoch: a launch draws from one and its first
// thread zeroes the other, which nobody in this launch touches and the NEXT launch (epoch + 1) draws from.  So the
// ticket needs neither a memset per launch nor a "last CTA out" hand-shake (whose __threadfence per CTA costs more
// than it saves: measured on the join, 0.41 -> 0.69 ms).  The host clears both when it restarts the epoch.
__device__ __forceinline__ unsigned long long* ticket_of(unsigned long long* counters, uint32_t epoch) {
  if (blockIdx.x == 0 && threadIdx.x == 0) counters[C_TICKET + ((epoch + 1) & 1)] = 0;
  return &counters[C_TICKET + (epoch & 1)];
}

__device__ __forceinline__ double series_value(int kind, const DevAcc& acc, const DevRender& R, uint32_t r, bool& live) {
  switch (kind) {
    case V_R_USED: live = acc.r_seq[r] != 0; return __ull2double_rn(acc.r_used[r]);            // main.go:147
    case V_R_PERC: live = acc.r_seq[r] != 0;                                                    // main.go:149
      return __dmul_rn(__ddiv_rn(__ull2double_rn(acc.r_used[r]), __ull2double_rn(acc.r_total[r])), 100.0);
    default: {
      const ulonglong2* xa = reinterpret_cast<const ulonglong2*>(acc.xa + size_t(r) * kXAWords);
      const unsigned long long* xb = acc.xb + size_t(r) * kXBWords;
      switch (kind) {  // count sits beside sum_used: one 16-byte load serves both
        case V_X_SUM: { const ulonglong2 a = xa[0]; live = a.x != 0; return __ull2double_rn(a.y); }
        case V_X_PERC: { const ulonglong2 a = xa[0]; live = a.x != 0;
          return __dmul_rn(__ddiv_rn(__ull2double_rn(a.y), __ull2double_rn(R.gpu_total[R.ser_gidx[r]])), 100.0); }
        case V_X_SM: { const unsigned long long n = acc.xa[size_t(r) * kXAWords + XA_COUNT]; const ulonglong2 b = xa[1]; live = n != 0;
          return __ddiv_rn(__ull2double_rn(b.x), __ull2double_rn(n)); }
        case V_X_MEM: { const unsigned long long n = acc.xa[size_t(r) * kXAWords + XA_COUNT]; const ulonglong2 b = xa[1]; live = n != 0;
          return __ddiv_rn(__ull2double_rn(b.y), __ull2double_rn(n)); }
        case V_X_MAX: live = acc.xa[size_t(r) * kXAWords + XA_COUNT] != 0; return __ull2double_rn(xb[XB_MAX_USED]);
        default: live = acc.xa[size_t(r) * kXAWords + XA_COUNT] != 0; return __ddiv_rn(__ull2double_rn(xb[XB_LAST_TS]), 1e6);
      }
    }
  }
}

constexpr int kRenderThreads = 256;

// One CTA per tile (taken in ticket order so look-back never waits on a tile
// that has not started).  v1 writer: every thread composes its own line in the
// shared-memory tile buffer, then the CTA streams the buffer out coalesced.
__global__ void __launch_bounds__(kRenderThreads) k_render(DevRender R, DevAcc acc, unsigned long long* counters) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ uint32_t s_warp[kRenderThreads / 32];
  __shared__ uint32_t s_tile;
  __shared__ unsigned long long s_base;
  const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  unsigned long long* ticket = ticket_of(counters, R.epoch);
  if (R.check_churn && *(volatile unsigned long long*)&counters[C_CHURN] != 0) return;
  for (;;) {
    __syncthreads();
    if (tid == 0) s_tile = uint32_t(atomicAdd(ticket, 1ull));
    __syncthreads();
    const uint32_t t = s_tile;
    if (t >= R.n_tiles) return;
    const Tile tile = R.tiles[t];
    const FamMeta& fm = R.fam[tile.family];
    // ---- pass 1: exact length of every line
    uint8_t txt[kMaxValueText];
    int vlen = 0;
    uint32_t len = 0, lbl_len = 0;
    unsigned long long lbl_at = 0;
    if (tid < tile.count) {
      const uint32_t r = tile.first + tid;
      bool live;
      const double v = series_value(fm.kind, acc, R, r, live);
      if (live) {
        vlen = format_value(v, txt);
        lbl_at = R.lbl_off[r];
        lbl_len = uint32_t(R.lbl_off[r + 1] - lbl_at);
        len = fm.name_len + lbl_len + 1 + uint32_t(vlen) + 1;
      }
    }
    const uint32_t hdr = ((tile.flags & 1) && acc.node_live[tile.node] != 0) ? fm.header_len : 0;
    // ---- exclusive scan of the line lengths over the CTA
    uint32_t inc = len;
    for (int o = 1; o < 32; o <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    uint32_t woff = 0, total = hdr;
    for (int k = 0; k < kRenderThreads / 32; ++k) { if (k < int(wid)) woff += s_warp[k]; total += s_warp[k]; }
    const uint32_t off = hdr + woff + inc - len;
    // ---- inter-tile offset: decoupled look-back over the tile chain
    if (tid == 0) {
      unsigned long long base = 0;
      if (t > 0) {
        atomicExch(&R.tile_state[t], kStAgg | st_tag(R.epoch) | total);
        for (uint32_t p = t; p-- > 0;) {
          unsigned long long s;
          do { s = *(volatile unsigned long long*)&R.tile_state[p]; } while (!st_published(s, R.epoch));
          base += s & kStValue;
          if ((s & kStMask) == kStPrefix) break;
        }
      }
      atomicExch(&R.tile_state[t], kStPrefix | st_tag(R.epoch) | (base + total));
      s_base = base;
      if (tile.flags & 2) R.node_body_off[tile.node] = st_tag(R.epoch) | base;
      if (t == R.n_tiles - 1) { counters[C_TOTAL_LEN] = base + total; if (R.host_len) *R.host_len = base + total; }
    }
    __syncthreads();
    const unsigned long long base = s_base;
    const uint32_t shift = uint32_t(base & 15);  // keep smem and global congruent mod 16
    // ---- pass 2: compose the tile text in shared memory
    uint8_t* buf = smem + shift;
    if (hdr) for (uint32_t i = tid; i < hdr; i += kRenderThreads) buf[i] = fm.header[i];
    if (len) {
      uint8_t* p = buf + off;
      for (uint32_t i = 0; i < fm.name_len; ++i) p[i] = fm.name[i];
      p += fm.name_len;
      const uint8_t* src = R.lbl_blob + lbl_at;
      for (uint32_t i = 0; i < lbl_len; ++i) p[i] = src[i];
      p += lbl_len;
      *p++ = ' ';
      for (int i = 0; i < vlen; ++i) p[i] = txt[i];
      p[vlen] = '\n';
    }
    __syncthreads();
    // ---- stream the tile out: byte head/tail, 16-byte body
    uint8_t* g = R.out + base;
    const uint32_t head = min(total, (16u - shift) & 15u);
    const uint32_t body = (total - head) & ~15u;
    if (tid < head) g[tid] = buf[tid];
    const uint4* sb = reinterpret_cast<const uint4*>(buf + head);
    uint4* gb = reinterpret_cast<uint4*>(g + head);
    for (uint32_t i = tid; i < body / 16; i += kRenderThreads) gb[i] = sb[i];
    for (uint32_t i = head + body + tid; i < total; i += kRenderThreads) g[i] = buf[i];
  }
}

// ---- values only (Collector read-back) -----------------------------------------------
__global__ void k_values(DevRender R, DevAcc acc, int kind, uint32_t n_series, double* values, uint8_t* live_out) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_series) return;
  bool live;
  values[r] = series_value(kind, acc, R, r, live);
  live_out[r] = live;
}

// ---- NVML wire format -> columns (R modes) ------------------------------------------------
struct NvmlProcessInfo { uint32_t pid, pad_; unsigned long long used; uint32_t gi, ci; };  // nvml.h:294-303
// one thread per row: which device (node*G+gpu) owns it, then scatter into the columns
__global__ void k_unpack_nvml(const NvmlProcessInfo* infos, const unsigned long long* dev_off, uint32_t n_dev, uint32_t G,
                              uint64_t n, uint32_t* pid, uint32_t* gpu, uint32_t* node, unsigned long long* used) {
  const uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  uint32_t lo = 0, hi = n_dev;  // last device with dev_off <= i
  while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (dev_off[mid] <= i) lo = mid; else hi = mid; }
  const NvmlProcessInfo p = infos[i];
  pid[i] = p.pid; used[i] = p.used; gpu[i] = lo % G; node[i] = lo / G;
}
__global__ void k_node_off_from_dev_off(const unsigned long long* dev_off, uint32_t n_nodes, uint32_t G, unsigned long long* node_off) {
  const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n <= n_nodes) node_off[n] = dev_off[size_t(n) * G];
}

// X mode: a sample row = a process-info row + the utilisation sample NVML reports for the same pid on the same
// device (nvmlDeviceGetProcessUtilization; the last one in array order when a pid occurs twice; none -> sm = mem =
// ts = 0).  One CTA per device at a time: the device's samples go into a shared-memory hash, its rows probe it.
struct NvmlProcessUtilizationSample { uint32_t pid, pad_; unsigned long long ts; uint32_t sm, mem, enc, dec; };  // nvml.h:1277-1285
constexpr uint32_t kUtilCells = 2048;
constexpr int kUtilThreads = 256;
__global__ void __launch_bounds__(kUtilThreads) k_join_nvml_util(const unsigned long long* dev_off, const NvmlProcessUtilizationSample* utils,
                                                                  const unsigned long long* util_off, uint32_t n_dev, const uint32_t* pid_col,
                                                                  uint32_t* sm, uint32_t* mem, unsigned long long* ts) {
  __shared__ unsigned long long cells[kUtilCells];  // pid << 32 | (index + 1); 0 = empty
  const uint32_t tid = threadIdx.x;
  for (uint32_t d = blockIdx.x; d < n_dev; d += gridDim.x) {
    const unsigned long long r0 = dev_off[d], r1 = dev_off[d + 1], u0 = util_off[d], m = util_off[d + 1] - u0;
    if (r1 == r0) continue;  // uniform over the CTA
    const bool hashed = m * 2 <= kUtilCells;
    if (hashed) {
      for (uint32_t i = tid; i < kUtilCells; i += kUtilThreads) cells[i] = 0;
      __syncthreads();
      for (uint32_t j = tid; j < m; j += kUtilThreads) {
        const uint32_t pid = utils[u0 + j].pid;
        const unsigned long long mine = ((unsigned long long)pid << 32) | (j + 1);
        uint32_t pos = dev_hash_pid(pid) & (kUtilCells - 1);
        for (;;) {
          const unsigned long long old = atomicCAS(&cells[pos], 0ull, mine);
          if (old == 0) break;
          if (uint32_t(old >> 32) == pid) { atomicMax(&cells[pos], mine); break; }
          pos = (pos + 1) & (kUtilCells - 1);
        }
      }
      __syncthreads();
    }
    for (unsigned long long i = r0 + tid; i < r1; i += kUtilThreads) {
      const uint32_t pid = pid_col[i];
      long long found = -1;
      if (hashed) {
        uint32_t pos = dev_hash_pid(pid) & (kUtilCells - 1);
        for (;;) {
          const unsigned long long cell = cells[pos];
          if (cell == 0) break;
          if (uint32_t(cell >> 32) == pid) { found = (long long)(cell & 0xffffffffull) - 1; break; }
          pos = (pos + 1) & (kUtilCells - 1);
        }
      } else {  // an unusually crowded device: scan its samples
        for (unsigned long long j = 0; j < m; ++j)
          if (utils[u0 + j].pid == pid) found = (long long)j;
      }
      if (found >= 0) {
        const NvmlProcessUtilizationSample u = utils[u0 + found];
        sm[i] = u.sm; mem[i] = u.mem; ts[i] = u.ts;
      } else {
        sm[i] = 0; mem[i] = 0; ts[i] = 0;
      }
    }
    __syncthreads();  // the hash is cleared for the next device
  }
}

// A few result words -> mapped host memory, by plain stores (system scope after the kernel ends): small read-backs
// must not queue on the device->host copy engine behind a body that is still being sent.
__global__ void k_post_words(const unsigned long long* src, unsigned long long* mapped_host, int n) {
  if (int(threadIdx.x) < n) mapped_host[threadIdx.x] = src[threadIdx.x];
}

// ---- float printer entry (tests) ---------------------------------------------------------
__global__ void k_format_values(const double* v, uint64_t n, uint8_t* out32) {
  const uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  uint8_t txt[kMaxValueText];
  const int len = format_value(v[i], txt);
  uint8_t* rec = out32 + 32 * i;
  for (int k = 0; k < 31; ++k) rec[k] = k < len ? txt[k] : 0;
  rec[31] = uint8_t(len);
}

}  // namespace kgx
