// kgx_api.cu — the C ABI of include/kgx.h: context, device memory, and the
// launch sequences of the join -> reduce -> render path.  No CPU fallback: every
// entry point that computes needs the CUDA device the ctx was created on.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/kgx.h"
#include "kgx_kernels.cuh"
#include "kgx_writer.cuh"
#include "kgx_writer5.cuh"
#include "kgx_ingest2.cuh"
#include "kgx_table.hpp"
#include "kgx_build.cuh"

using namespace kgx;

namespace {

struct DBuf {  // growable device buffer
  void* p = nullptr;
  size_t cap = 0;
};

struct TimedRegion { int kind; cudaEvent_t a, b; };

}  // namespace

struct kgx_ctx {
  int device = 0, mode = 0;
  bool profile = false, own_stream = false;
  cudaStream_t stream = nullptr;
  cudaStream_t copy_stream = nullptr;  // kgx_render_begin: body D2H overlaps the next ingest
  cudaEvent_t ev_rendered = nullptr;
  // Bodies leave through two device buffers in turn: while one is still being copied to the host, the next scrape
  // is rendered into the other (kgx_render_begin).  *_busy: a copy started by kgx_render_begin still reads that buffer.
  cudaEvent_t ev_copy_out = nullptr, ev_copy_alt = nullptr;
  bool out_busy = false, alt_busy = false;
  // mu serialises everything that touches the series state (one writer, any number of readers: a reader sees the last
  // completed ingest).  mu_in only guards the staging of a kgx_ingest batch (host->device copies into d_cols on
  // in_stream): that part runs WITHOUT mu, so a scrape never waits for a batch that is still being copied in, only
  // for a join that is actually running.  Lock order: mu_in, then mu.
  std::mutex mu, mu_in;
  cudaStream_t in_stream = nullptr;
  uint64_t table_gen = 0;  // bumped by kgx_set_table
  int sm_count = 148;
  std::atomic<size_t> device_bytes{0};

  bool have_table = false;
  Plan plan;
  // table
  DBuf d_node_tab, d_tab, d_pk_slot0, d_pk_stride, d_grank, d_node_slot_base, d_node_a_base;
  uint32_t ingest_smem = 0, ingest_cells = 0, ingest_pks = 0, ingest_words = 0;  // staging buffer of the node-grouped join
  DBuf d_A_blob, d_B_blob, d_G_blob, d_A_off, d_B_off, d_G_off, d_runs;
  uint32_t n_runs = 0;
  // table build on the device (kgx_build.cuh): scratch chunks, and — R modes — who every slot and pod of the
  // current generation is, so that series with a value can be carried into the next table without leaving the GPU
  bool device_build = true, build_dirty = false;
  std::vector<DBuf> arena;
  size_t arena_i = 0, arena_off = 0;
  DBuf d_slot_pid[2], d_slot_pod[2], d_pods_node[2], d_pods_blob[2], d_pods_off[2];
  int gen = 0;
  uint32_t n_pods = 0, max_words = 0;
  // index
  DBuf d_bm, d_pending, d_popc, d_prefix, d_block_sums;
  uint32_t n_words = 0;
  // series
  uint32_t n_series = 0;
  DBuf d_ser_slot, d_ser_a, d_ser_gidx, d_ser_len, d_lbl_off, d_lbl_blob, d_node_ser_off;
  uint64_t lbl_bytes = 0;
  std::vector<uint32_t> node_ser_off;
  // per-series state
  DBuf d_acc[6];  // X: [0] = xa, [1] = xb records per series (kgx_kernels.cuh) ; R: r_seq,r_used,r_total
  DBuf d_acc_spare[6];  // target of the next index rebuild
  DBuf d_node_live, d_gpu_total;
  bool have_total = false;
  // render
  DBuf d_fam, d_tiles, d_tile_state, d_node_body_off, d_out, d_out_alt;
  uint32_t n_tiles = 0, tile_lines = 0, render_smem = 0;
  bool render_v2 = false;  // second-generation writer (needs metric names of 15..64 bytes)
  RenderLayout lay{};
  int render_grid = 0, render_threads = 128;  // measured best on B200: direct stores, 128-line tiles
  size_t last_len = 0;
  bool rendered = false;
  // Small configurations (one node, a few hundred processes) are bound by calls into the driver, not by bytes: a batch
  // whose columns fit kSmallBytes crosses in ONE copy from a pinned staging buffer, and a body whose upper bound fits
  // leaves in one copy queued right behind the writer (no second synchronisation).
  uint64_t out_bound = 0;          // upper bound of the body size (what d_out holds)
  uint8_t *h_small_in = nullptr, *h_small_out = nullptr;  // pinned, kSmallBytes each, allocated on first use
  DBuf d_small_in;
  std::vector<uint64_t> body_off;  // host copy of the last body's per-node offsets (kgx_body_node), fetched on first use
  bool body_off_valid = false;
  std::vector<FamMeta> fam;
  // misc
  DBuf d_counters;
  unsigned long long* h_counters = nullptr;  // pinned and mapped: kernels post results here (k_post_words)
  unsigned long long* h_counters_dev = nullptr;  // its device alias
  DBuf d_cols[9];  // staging of a kgx_ingest batch (guarded by mu_in): the seven columns, node_off, gpu_total
  DBuf d_tmp[6];
  uint32_t seq = 0;
  // kgx_ingest_device returns once its kernels are queued: the statistics and the churn check of that join are
  // settled by the next call that needs them (settle_ingest), or ride on the writer's own synchronisation
  bool reset_pending = false;  // X: kgx_reset_window was called; the next node-grouped join clears as it goes (else flush_reset)
  bool ingest_pending = false;
  kgx::DevCols pending_cols{};
  bool pending_grouped = false;
  uint32_t render_epoch = 0, body_epoch = 0;  // tag of the look-back words / body offsets of the latest launch
  kgx_stats stats{};
  kgx_profile prof{};
  std::vector<TimedRegion> regions;
};

namespace {

constexpr size_t kSmallBytes = 256 << 10;
constexpr int kCounterWords = 16;  // C_COUNT words the kernels know + two scratch words for kgx_expire
constexpr int kExpireWords = 8;

const char* const kErrNames[] = {"ok", "invalid argument", "out of memory", "CUDA error", "output buffer too small",
                                 "call out of order", "size beyond a documented maximum"};

// Detail of the last error, per calling thread: the poll loop and the scrape handlers call into one ctx concurrently
// and must not see (or tear) each other's message.
thread_local std::string tl_err;

#define KGX_CUDA(c, call)                                                                      \
  do {                                                                                         \
    cudaError_t e_ = (call);                                                                   \
    if (e_ != cudaSuccess) {                                                                   \
      (void)(c);                                                                               \
      tl_err = std::string(#call) + ": " + cudaGetErrorString(e_);                             \
      return e_ == cudaErrorMemoryAllocation ? KGX_E_NOMEM : KGX_E_CUDA;                       \
    }                                                                                          \
  } while (0)

#define KGX_TRY(expr)         \
  do {                        \
    int rc_ = (expr);         \
    if (rc_ != KGX_OK) return rc_; \
  } while (0)

int fail(kgx_ctx*, int code, const char* msg) { tl_err = msg; return code; }

int dev_reserve(kgx_ctx* c, DBuf& b, size_t bytes) {
  if (bytes <= b.cap && b.p) return KGX_OK;
  if (b.p) { cudaFree(b.p); c->device_bytes -= b.cap; b.p = nullptr; b.cap = 0; }
  size_t want = std::max<size_t>(bytes + bytes / 8, 256);  // headroom: series and tables grow a little every cycle
  KGX_CUDA(c, cudaMalloc(&b.p, want));
  b.cap = want;
  c->device_bytes += want;
  return KGX_OK;
}
void dev_free(kgx_ctx* c, DBuf& b) {
  if (b.p) { cudaFree(b.p); c->device_bytes -= b.cap; }
  b.p = nullptr; b.cap = 0;
}
template <class T>
int upload(kgx_ctx* c, DBuf& b, const T* src, size_t n) {
  KGX_TRY(dev_reserve(c, b, n * sizeof(T)));
  if (n) KGX_CUDA(c, cudaMemcpyAsync(b.p, src, n * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  return KGX_OK;
}
template <class T> T* ptr(const DBuf& b) { return static_cast<T*>(b.p); }

struct Timed {  // CUDA-event bracket around a launch sequence when profiling
  kgx_ctx* c; int kind; cudaEvent_t a = nullptr, b = nullptr;
  Timed(kgx_ctx* c_, int kind_) : c(c_), kind(kind_) {
    if (!c->profile) return;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a, c->stream);
  }
  void launches(int n) { c->stats.kernel_launches += n; c->prof.launches[kind] += n; }
  ~Timed() {
    if (!c->profile) return;
    cudaEventRecord(b, c->stream);
    c->regions.push_back(TimedRegion{kind, a, b});
  }
};

// body copies started by kgx_render_begin still read d_out / d_out_alt: finish them before those are rewritten or freed
void wait_copy(kgx_ctx* c) {
  if (c->out_busy || c->alt_busy) { cudaStreamSynchronize(c->copy_stream); c->out_busy = c->alt_busy = false; }
}

int launch_check(kgx_ctx* c, const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { (void)c; tl_err = std::string(what) + ": " + cudaGetErrorString(e); return KGX_E_CUDA; }
  return KGX_OK;
}

inline uint32_t cdiv(uint64_t a, uint64_t b) { return uint32_t((a + b - 1) / b); }

// exclusive scan of n u32 items -> out64 and/or out32; total -> *total_host (sync)
int device_scan(kgx_ctx* c, const uint32_t* in, uint64_t n, unsigned long long* out64, uint32_t* out32,
                unsigned long long* total_host) {
  const uint32_t nb = std::max<uint32_t>(1, cdiv(n, kScanBlock * kScanItems));
  KGX_TRY(dev_reserve(c, c->d_block_sums, (size_t(nb) + 1) * 8));
  unsigned long long* bs = ptr<unsigned long long>(c->d_block_sums);
  k_scan_reduce<<<nb, kScanBlock, 0, c->stream>>>(in, n, bs);
  k_scan_blocksums<<<1, 1024, 0, c->stream>>>(bs, nb, bs + nb);
  k_scan_apply<<<nb, kScanBlock, 0, c->stream>>>(in, n, bs, out64, out32);
  KGX_TRY(launch_check(c, "scan"));
  KGX_CUDA(c, cudaMemcpyAsync(total_host, bs + nb, 8, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  return KGX_OK;
}

DevAcc acc_view(kgx_ctx* c) {
  DevAcc a{};
  if (c->mode == KGX_MODE_X) {
    a.xa = ptr<unsigned long long>(c->d_acc[0]); a.xb = ptr<unsigned long long>(c->d_acc[1]);
  } else {
    a.r_seq = ptr<unsigned long long>(c->d_acc[0]); a.r_used = ptr<unsigned long long>(c->d_acc[1]);
    a.r_total = ptr<unsigned long long>(c->d_acc[2]);
  }
  a.node_live = ptr<uint32_t>(c->d_node_live);
  return a;
}
int n_acc_arrays(const kgx_ctx* c) { return c->mode == KGX_MODE_X ? 2 : 3; }  // X: xa, xb (kgx_kernels.cuh)
int acc_words(const kgx_ctx* c, int k) { return c->mode == KGX_MODE_X ? (k == 0 ? int(kXAWords) : int(kXBWords)) : 1; }  // words per series

// the writer is compiled for four CTA sizes (= lines per tile); 128 measured best on B200
const void* writer_func(const kgx_ctx* c) {
  switch (c->render_threads) {
    case 32: return reinterpret_cast<const void*>(&k_render5<32>);
    case 64: return reinterpret_cast<const void*>(&k_render5<64>);
    case 256: return reinterpret_cast<const void*>(&k_render5<256>);
    default: return reinterpret_cast<const void*>(&k_render5<128>);
  }
}
void launch_writer(kgx_ctx* c, int grid, const DevRender& R, unsigned long long* counters) {
  switch (c->render_threads) {
    case 32: k_render5<32><<<grid, 32, c->render_smem, c->stream>>>(R, c->lay, acc_view(c), counters); break;
    case 64: k_render5<64><<<grid, 64, c->render_smem, c->stream>>>(R, c->lay, acc_view(c), counters); break;
    case 256: k_render5<256><<<grid, 256, c->render_smem, c->stream>>>(R, c->lay, acc_view(c), counters); break;
    default: k_render5<128><<<grid, 128, c->render_smem, c->stream>>>(R, c->lay, acc_view(c), counters); break;
  }
}

DevLabels labels_view(kgx_ctx* c) {
  DevLabels L{};
  L.A_blob = ptr<uint8_t>(c->d_A_blob); L.B_blob = ptr<uint8_t>(c->d_B_blob); L.G_blob = ptr<uint8_t>(c->d_G_blob);
  L.A_off = ptr<unsigned long long>(c->d_A_off); L.B_off = ptr<unsigned long long>(c->d_B_off);
  L.G_off = ptr<unsigned long long>(c->d_G_off);
  L.runs = ptr<Run>(c->d_runs); L.n_runs = c->n_runs;
  return L;
}

void build_family_meta(kgx_ctx* c) {
  c->fam.clear();
  for (const FamilyDef& f : families(c->mode)) {
    FamMeta m{};
    m.name_len = uint32_t(strlen(f.name));
    memcpy(m.name, f.name, m.name_len);
    std::string h = std::string("# HELP ") + f.name + " ";
    for (const char* p = f.help; *p; ++p) {  // expfmt escapes \ and newline in help (G2)
      if (*p == '\\') h += "\\\\"; else if (*p == '\n') h += "\\n"; else h += *p;
    }
    h += std::string("\n# TYPE ") + f.name + " gauge\n";
    m.header_len = uint32_t(h.size());
    memcpy(m.header, h.data(), h.size());
    m.kind = f.value_kind;
    c->fam.push_back(m);
  }
}

// Merge the pending slots into the series index, re-rank every series, move
// the per-series state to the new ranks, re-materialise label text and tiles.
int rebuild_index(kgx_ctx* c) {
  wait_copy(c);
  Timed tm(c, KGX_K_INDEX);
  const uint32_t nw = c->n_words;
  const uint32_t old_n = c->n_series;
  KGX_TRY(dev_reserve(c, c->d_popc, size_t(nw) * 4));
  KGX_TRY(dev_reserve(c, c->d_prefix, size_t(nw) * 4));
  k_merge_pending<<<cdiv(nw, 256), 256, 0, c->stream>>>(ptr<uint2>(c->d_bm), ptr<uint32_t>(c->d_pending), nw,
                                                        ptr<uint32_t>(c->d_popc));
  unsigned long long total = 0;
  KGX_TRY(device_scan(c, ptr<uint32_t>(c->d_popc), nw, nullptr, ptr<uint32_t>(c->d_prefix), &total));
  k_set_prefix<<<cdiv(nw, 256), 256, 0, c->stream>>>(ptr<uint2>(c->d_bm), ptr<uint32_t>(c->d_prefix), nw);
  tm.launches(5);
  const uint32_t S = uint32_t(total);
  // new per-series state, zeroed, then the old state moved over
  const int na = n_acc_arrays(c);
  DBuf* new_acc = c->d_acc_spare;  // the two sets swap roles at every rebuild: no allocation in steady churn
  for (int k = 0; k < na; ++k) {
    KGX_TRY(dev_reserve(c, new_acc[k], size_t(S) * 8 * acc_words(c, k)));
    KGX_CUDA(c, cudaMemsetAsync(new_acc[k].p, 0, size_t(S) * 8 * acc_words(c, k), c->stream));
  }
  if (old_n) {
    const unsigned long long* h_old[6]; unsigned long long* h_new[6];
    for (int k = 0; k < na; ++k) { h_old[k] = ptr<unsigned long long>(c->d_acc[k]); h_new[k] = ptr<unsigned long long>(new_acc[k]); }
    KGX_TRY(dev_reserve(c, c->d_tmp[0], 2 * 6 * sizeof(void*)));
    KGX_CUDA(c, cudaMemcpyAsync(c->d_tmp[0].p, h_old, sizeof h_old, cudaMemcpyHostToDevice, c->stream));
    KGX_CUDA(c, cudaMemcpyAsync(static_cast<char*>(c->d_tmp[0].p) + sizeof h_old, h_new, sizeof h_new, cudaMemcpyHostToDevice, c->stream));
    for (int k = 0; k < na; ++k)  // one launch per array: their records differ in size
      k_remap<<<cdiv(old_n, 256), 256, 0, c->stream>>>(ptr<uint32_t>(c->d_ser_slot), old_n, ptr<uint2>(c->d_bm), 1, acc_words(c, k),
                                                       reinterpret_cast<const unsigned long long* const*>(c->d_tmp[0].p) + k,
                                                       reinterpret_cast<unsigned long long* const*>(static_cast<char*>(c->d_tmp[0].p) + sizeof h_old) + k);
    tm.launches(1);
    KGX_CUDA(c, cudaStreamSynchronize(c->stream));  // h_old/h_new are stack arrays
  }
  for (int k = 0; k < na; ++k) std::swap(c->d_acc[k], c->d_acc_spare[k]);
  // series arrays
  KGX_TRY(dev_reserve(c, c->d_ser_slot, size_t(S) * 4));
  KGX_TRY(dev_reserve(c, c->d_ser_a, size_t(S) * 4));
  KGX_TRY(dev_reserve(c, c->d_ser_gidx, size_t(S) * 4));
  KGX_TRY(dev_reserve(c, c->d_ser_len, size_t(S) * 4 + 4));
  KGX_TRY(dev_reserve(c, c->d_lbl_off, (size_t(S) + 1) * 8));
  DevLabels L = labels_view(c);
  k_build_series<<<cdiv(nw, 128), 128, 0, c->stream>>>(ptr<uint2>(c->d_bm), nw, L, ptr<uint32_t>(c->d_ser_slot),
                                                       ptr<uint32_t>(c->d_ser_a), ptr<uint32_t>(c->d_ser_gidx),
                                                       ptr<uint32_t>(c->d_ser_len));
  unsigned long long lbl_total = 0;
  KGX_TRY(device_scan(c, ptr<uint32_t>(c->d_ser_len), S, ptr<unsigned long long>(c->d_lbl_off), nullptr, &lbl_total));
  KGX_CUDA(c, cudaMemcpyAsync(ptr<unsigned long long>(c->d_lbl_off) + S, &lbl_total, 8, cudaMemcpyHostToDevice, c->stream));
  KGX_TRY(dev_reserve(c, c->d_lbl_blob, size_t(lbl_total) + 64));
  if (S) k_fill_labels<<<cdiv(uint64_t(S) * 32, 256), 256, 0, c->stream>>>(L, ptr<uint32_t>(c->d_ser_a), ptr<uint32_t>(c->d_ser_gidx),
                                                                           ptr<unsigned long long>(c->d_lbl_off), S,
                                                                           ptr<uint8_t>(c->d_lbl_blob));
  // per-node series ranges -> tiles
  const uint32_t NN = c->plan.n_nodes;
  KGX_TRY(dev_reserve(c, c->d_node_ser_off, (size_t(NN) + 1) * 4));
  k_node_ranks<<<cdiv(NN + 1, 256), 256, 0, c->stream>>>(ptr<uint2>(c->d_bm), ptr<uint32_t>(c->d_node_slot_base), NN,
                                                         uint32_t(c->plan.n_slots), S, ptr<uint32_t>(c->d_node_ser_off));
  tm.launches(6);
  KGX_TRY(launch_check(c, "index rebuild"));
  c->node_ser_off.assign(size_t(NN) + 1, 0);
  KGX_CUDA(c, cudaMemcpyAsync(c->node_ser_off.data(), c->d_node_ser_off.p, (size_t(NN) + 1) * 4, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  c->n_series = S;
  c->lbl_bytes = lbl_total;
  // tiles: (node, family, chunk of series) in output order
  std::vector<Tile> tiles;
  const uint32_t TL = c->tile_lines;
  if (TL == 0) return fail(c, KGX_E_STATE, "render tile geometry not set");
  const int nf = int(c->fam.size());
  for (uint32_t n = 0; n < NN; ++n) {
    const uint32_t lo = c->node_ser_off[n], hi = c->node_ser_off[n + 1];
    if (hi == lo) continue;
    for (int f = 0; f < nf; ++f)
      for (uint32_t k = lo; k < hi; k += TL) {
        Tile t{};
        t.first = k; t.node = n; t.count = uint16_t(std::min(TL, hi - k)); t.family = uint8_t(f);
        t.flags = uint8_t((k == lo ? 1 : 0) | ((k == lo && f == 0) ? 2 : 0));
        tiles.push_back(t);
      }
  }
  c->n_tiles = uint32_t(tiles.size());
  KGX_TRY(upload(c, c->d_tiles, tiles.data(), tiles.size()));
  if (c->n_tiles) {
    k_tile_labels<<<cdiv(c->n_tiles, 256), 256, 0, c->stream>>>(ptr<Tile>(c->d_tiles), c->n_tiles, ptr<unsigned long long>(c->d_lbl_off));
    tm.launches(1);
  }
  KGX_TRY(dev_reserve(c, c->d_tile_state, size_t(c->n_tiles) * 8 + 8));
  KGX_CUDA(c, cudaMemsetAsync(c->d_tile_state.p, 0, c->d_tile_state.cap, c->stream));  // (re)allocated: no stale epoch tags
  KGX_CUDA(c, cudaMemsetAsync(c->d_node_body_off.p, 0xff, c->d_node_body_off.cap, c->stream));
  KGX_CUDA(c, cudaMemsetAsync(ptr<unsigned long long>(c->d_counters) + C_TICKET, 0, 16, c->stream));  // both tickets
  c->render_epoch = 0;
  // R modes: every series has a value, so a node is live iff it has series
  if (c->mode != KGX_MODE_X) {
    std::vector<uint32_t> live(NN);
    for (uint32_t n = 0; n < NN; ++n) live[n] = c->node_ser_off[n + 1] - c->node_ser_off[n];
    KGX_CUDA(c, cudaMemcpyAsync(c->d_node_live.p, live.data(), size_t(NN) * 4, cudaMemcpyHostToDevice, c->stream));
  }
  // output capacity: an upper bound of the body size
  uint64_t cap = 64;
  for (const FamMeta& m : c->fam)
    cap += uint64_t(S) * (m.name_len + 2 + kMaxValueText) + lbl_total + uint64_t(NN) * m.header_len;
  KGX_TRY(dev_reserve(c, c->d_out, size_t(cap)));
  c->out_bound = cap;
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  c->stats.rebuilds += 1;
  return KGX_OK;
}

void free_table_state(kgx_ctx* c) {
  wait_copy(c);
  DBuf* all[] = {&c->d_node_a_base, &c->d_node_tab, &c->d_tab, &c->d_pk_slot0, &c->d_pk_stride, &c->d_grank, &c->d_node_slot_base,
                 &c->d_A_blob, &c->d_B_blob, &c->d_G_blob, &c->d_A_off, &c->d_B_off, &c->d_G_off, &c->d_runs,
                 &c->d_bm, &c->d_pending, &c->d_popc, &c->d_prefix, &c->d_ser_slot, &c->d_ser_a, &c->d_ser_gidx,
                 &c->d_ser_len, &c->d_lbl_off, &c->d_lbl_blob, &c->d_node_ser_off, &c->d_node_live, &c->d_gpu_total,
                 &c->d_tiles, &c->d_tile_state, &c->d_node_body_off, &c->d_out, &c->d_out_alt};
  for (DBuf* b : all) dev_free(c, *b);
  for (auto& b : c->d_acc) dev_free(c, b);
  for (auto& b : c->d_acc_spare) dev_free(c, b);
  for (int g = 0; g < 2; ++g) {
    dev_free(c, c->d_slot_pid[g]); dev_free(c, c->d_slot_pod[g]); dev_free(c, c->d_pods_node[g]);
    dev_free(c, c->d_pods_blob[g]); dev_free(c, c->d_pods_off[g]);
  }
  for (auto& b : c->arena) dev_free(c, b);
  c->arena.clear(); c->arena_i = 0; c->arena_off = 0; c->n_pods = 0; c->n_runs = 0;
  c->n_series = 0; c->n_tiles = 0; c->n_words = 0; c->lbl_bytes = 0; c->have_total = false; c->rendered = false;
  c->last_len = 0;
}

// A new table replaces the old one in the same buffers (they only ever grow): under churn the
// pod->PID map changes every cycle and must not cost thirty cudaFree/cudaMalloc pairs each time.
void reset_table_state(kgx_ctx* c) {
  wait_copy(c);
  c->n_series = 0; c->n_tiles = 0; c->n_words = 0; c->lbl_bytes = 0; c->have_total = false; c->rendered = false;
  c->last_len = 0;
}

int collect_carried(kgx_ctx* c, std::vector<Carried>* out) {
  out->clear();
  if (c->mode == KGX_MODE_X || !c->have_table || c->n_series == 0) return KGX_OK;
  const uint32_t S = c->n_series;
  std::vector<uint32_t> slot(S);
  std::vector<unsigned long long> seq(S), used(S), total(S);
  KGX_CUDA(c, cudaMemcpyAsync(slot.data(), c->d_ser_slot.p, size_t(S) * 4, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaMemcpyAsync(seq.data(), c->d_acc[0].p, size_t(S) * 8, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaMemcpyAsync(used.data(), c->d_acc[1].p, size_t(S) * 8, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaMemcpyAsync(total.data(), c->d_acc[2].p, size_t(S) * 8, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  out->reserve(S);
  for (uint32_t r = 0; r < S; ++r) {
    if (!seq[r]) continue;
    const uint32_t s = slot[r], pod = c->plan.slot_pod[s];
    out->push_back(Carried{c->plan.pod_node[pod], c->plan.slot_pid[s], c->plan.pod_names[pod], used[r], total[r]});
  }
  return KGX_OK;
}

int upload_plan(kgx_ctx* c) {
  Plan& p = c->plan;
  KGX_TRY(upload(c, c->d_node_tab, p.node_tab.data(), p.node_tab.size()));
  KGX_TRY(upload(c, c->d_tab, p.tab.data(), p.tab.size()));
  KGX_TRY(upload(c, c->d_pk_slot0, p.pk_slot0.data(), p.pk_slot0.size()));
  KGX_TRY(upload(c, c->d_pk_stride, p.pk_stride.data(), p.pk_stride.size()));
  KGX_TRY(upload(c, c->d_grank, p.grank.data(), p.grank.size()));
  KGX_TRY(upload(c, c->d_node_a_base, p.node_a_base.data(), p.node_a_base.size()));
  KGX_TRY(upload(c, c->d_node_slot_base, p.node_slot_base.data(), p.node_slot_base.size()));
  KGX_TRY(upload(c, c->d_A_blob, p.A_blob.data(), p.A_blob.size()));
  KGX_TRY(upload(c, c->d_B_blob, p.B_blob.data(), p.B_blob.size()));
  KGX_TRY(upload(c, c->d_G_blob, p.G_blob.data(), p.G_blob.size()));
  KGX_TRY(upload(c, c->d_A_off, p.A_off.data(), p.A_off.size()));
  KGX_TRY(upload(c, c->d_B_off, p.B_off.data(), p.B_off.size()));
  KGX_TRY(upload(c, c->d_G_off, p.G_off.data(), p.G_off.size()));
  KGX_TRY(upload(c, c->d_runs, p.runs.data(), p.runs.size()));
  c->n_runs = uint32_t(p.runs.size());
  c->max_words = 0;  // series-index words one node spans
  for (uint32_t n = 0; n < p.n_nodes; ++n)
    c->max_words = std::max(c->max_words, ((p.node_slot_base[n + 1] + 31) >> 5) - (p.node_slot_base[n] >> 5));
  return KGX_OK;
}

// The dynamic shared-memory limit is an attribute of the FUNCTION (per device), shared by every ctx in the process: it is
// only ever raised, or a ctx with a small table would pull it below what another ctx's launches need.
int raise_smem_limit(kgx_ctx* c, const void* fn, uint32_t bytes) {
  static std::mutex mu;
  static std::vector<std::pair<std::pair<int, const void*>, uint32_t>> seen;
  std::lock_guard<std::mutex> g(mu);
  for (auto& e : seen)
    if (e.first.first == c->device && e.first.second == fn) {
      if (bytes <= e.second) return KGX_OK;
      KGX_CUDA(c, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes)));
      e.second = bytes;
      return KGX_OK;
    }
  KGX_CUDA(c, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes)));
  seen.push_back({{c->device, fn}, bytes});
  return KGX_OK;
}

// Everything that depends on the table only through its sizes (plan.n_slots, n_nodes, G, max_*; ctx.max_words):
// the join's staging buffer, the series index, the render tile geometry.  Shared by the host and the device build.
int finish_plan(kgx_ctx* c) {
  Plan& p = c->plan;
  {  // node-grouped join: stage a node's table slice (+ its podkey slot tables) in shared memory when it fits
    const uint64_t budget = 100 * 1024;  // two CTAs per SM
    const uint32_t max_words = c->max_words;
    const uint64_t need = uint64_t(p.max_node_cells) * sizeof(Entry) + uint64_t(p.max_node_pks) * 8 + uint64_t(max_words) * 8;
    if (need <= budget) { c->ingest_cells = p.max_node_cells; c->ingest_pks = p.max_node_pks; c->ingest_words = max_words; }
    else { c->ingest_cells = 0; c->ingest_pks = 0; c->ingest_words = 0; }  // too large: probe in global memory, still node by node
    c->ingest_smem = uint32_t(uint64_t(c->ingest_cells) * sizeof(Entry) + uint64_t(c->ingest_pks) * 8 + uint64_t(c->ingest_words) * 8 + 16);
    const void* fns[] = {(const void*)&k_ingest_nodes<true, 0>, (const void*)&k_ingest_nodes<true, 1>,
                         (const void*)&k_ingest_nodes<false, 0>, (const void*)&k_ingest_nodes<false, 1>};
    for (const void* f : fns) KGX_TRY(raise_smem_limit(c, f, c->ingest_smem));
  }
  c->n_words = cdiv(p.n_slots, 32) + 1;
  KGX_TRY(dev_reserve(c, c->d_bm, size_t(c->n_words) * 8));
  KGX_TRY(dev_reserve(c, c->d_pending, size_t(c->n_words) * 4));
  KGX_CUDA(c, cudaMemsetAsync(c->d_bm.p, 0, size_t(c->n_words) * 8, c->stream));
  KGX_CUDA(c, cudaMemsetAsync(c->d_pending.p, 0, size_t(c->n_words) * 4, c->stream));
  KGX_TRY(dev_reserve(c, c->d_node_live, size_t(p.n_nodes) * 4));
  KGX_CUDA(c, cudaMemsetAsync(c->d_node_live.p, 0, size_t(p.n_nodes) * 4, c->stream));
  KGX_TRY(dev_reserve(c, c->d_gpu_total, size_t(p.n_nodes) * p.G * 8));
  KGX_CUDA(c, cudaMemsetAsync(c->d_gpu_total.p, 0, size_t(p.n_nodes) * p.G * 8, c->stream));
  KGX_TRY(dev_reserve(c, c->d_node_body_off, (size_t(p.n_nodes) + 1) * 8));
  KGX_TRY(upload(c, c->d_fam, c->fam.data(), c->fam.size()));
  // tile geometry: lines per tile so that a whole tile fits the shared-memory buffers
  uint32_t max_name = 0, min_name = ~0u, max_header = 0;
  for (const FamMeta& m : c->fam) {
    max_name = std::max(max_name, m.name_len); min_name = std::min(min_name, m.name_len);
    max_header = std::max(max_header, m.header_len);
  }
  const uint64_t max_line = uint64_t(max_name) + p.max_label_len + 2 + kMaxValueText;
  c->render_v2 = min_name >= 15 && max_name <= 64 && c->fam.size() <= size_t(kMaxFamilies) && !getenv("KGX_RENDER_V1");
  int occ = 1;
  if (c->render_v2) {
    // per line: value slot + staged label.  Small CTAs: many independent tiles per SM hide each other's latencies
    // lines per tile = threads per CTA; measured best on B200: 128
    c->render_threads = 128;
    if (const char* e = getenv("KGX_RENDER_THREADS")) c->render_threads = atoi(e);
    if (c->render_threads != 32 && c->render_threads != 64 && c->render_threads != 256) c->render_threads = 128;
    const uint64_t slot_bufs = 2;  // the tile being emitted and the one being measured
    const uint64_t fixed = uint64_t(kMaxFamilies) * kNameStride + slot_bufs * c->render_threads * kSlotBytes + 32 + (32 + 128 + 48) + 32;
    uint64_t budget = 200 * 1024;
    if (fixed + p.max_label_len > budget) return fail(c, KGX_E_LIMIT, "label text too long for one render tile");
    c->tile_lines = uint32_t(std::min<uint64_t>(uint64_t(c->render_threads), (budget - fixed) / std::max<uint64_t>(1, p.max_label_len)));
    auto up16 = [](uint64_t v) { return uint32_t((v + 15) & ~15ull); };
    c->lay.names = 0;
    c->lay.slots = up16(uint64_t(kMaxFamilies) * kNameStride);
    c->lay.stage = up16(c->lay.slots + slot_bufs * c->render_threads * kSlotBytes + 32);
    c->lay.total = up16(c->lay.stage + 32 + uint64_t(c->tile_lines) * p.max_label_len + 128 + 48);
    c->render_smem = c->lay.total;
    KGX_TRY(raise_smem_limit(c, writer_func(c), c->render_smem));
    KGX_CUDA(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, writer_func(c), c->render_threads, c->render_smem));
  } else {
    const uint64_t budget = 96 * 1024;
    if (max_line + max_header + 32 > budget) return fail(c, KGX_E_LIMIT, "label text too long for one render tile");
    c->tile_lines = uint32_t(std::min<uint64_t>(kRenderThreads, (budget - max_header - 32) / max_line));
    c->render_smem = uint32_t((max_header + uint64_t(c->tile_lines) * max_line + 32 + 15) & ~15ull);
    KGX_TRY(raise_smem_limit(c, reinterpret_cast<const void*>(&k_render), c->render_smem));
    KGX_CUDA(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_render, kRenderThreads, c->render_smem));
  }
  c->render_grid = std::max(1, occ) * c->sm_count;
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  return KGX_OK;
}

}  // namespace
#include "kgx_build_host.cuh"
namespace {

// kgx_set_table with the plan built on the GPU.  `*carried` receives what must be re-seeded after the index exists.
int set_table_device(kgx_ctx* c, const kgx_table* t, CarriedDev* carried) {
  const bool x = c->mode == KGX_MODE_X;
  KGX_TRY(check_table_args(c, t, x));
  wait_copy(c);
  Arena ar(c);
  c->build_dirty = false;
  KGX_TAKE(uint32_t, small, 16);
  KGX_CUDA(c, cudaMemsetAsync(small, 0, 64, c->stream));
  const BuildSmall sm{small, small + 1, small + 1 + BS_COUNT};
  TableIn in;
  KGX_TRY(upload_table(c, ar, t, x, &in));
  int rc;
  if (x) {
    rc = build_table_device_x(c, ar, in, sm);
  } else {
    const int g = c->gen;
    const OldGen og{ptr<uint32_t>(c->d_slot_pid[g]), ptr<uint32_t>(c->d_slot_pod[g]), ptr<uint32_t>(c->d_pods_node[g]),
                    ptr<uint8_t>(c->d_pods_blob[g]), ptr<unsigned long long>(c->d_pods_off[g]), c->have_table ? c->n_pods : 0u};
    rc = collect_carried_device(c, ar, og, carried);
    if (rc == KGX_OK) rc = build_table_device_r(c, ar, in, sm, og, *carried);
  }
  if (rc != KGX_OK) {
    if (c->build_dirty) { c->have_table = false; reset_table_state(c); }  // the old table's buffers were already being rewritten
    return rc;
  }
  uint32_t st[BS_COUNT] = {};
  KGX_TRY(read_back(c, st, sm.stats, BS_COUNT));
  Plan& p = c->plan;
  const uint64_t n_slots = p.n_slots, n_a = p.n_a, n_entries = p.n_table_entries;
  p = Plan();  // the host-side arrays of a host-built plan are stale now
  p.mode = c->mode; p.n_nodes = t->n_nodes; p.G = t->gpus_per_node;
  p.n_slots = n_slots; p.n_a = n_a; p.n_table_entries = n_entries;
  p.max_label_len = st[BS_MAX_AB] + st[BS_MAX_G];
  p.max_node_cells = st[BS_MAX_NODE_CELLS]; p.max_node_pks = st[BS_MAX_NODE_PKS];
  c->max_words = st[BS_MAX_WORDS];
  return KGX_OK;
}

// A pending window reset, carried out on its own (the next call is not a node-grouped join that could fold it in).
int flush_reset(kgx_ctx* c) {
  if (!c->reset_pending) return KGX_OK;
  c->reset_pending = false;
  if (!c->have_table || c->mode != KGX_MODE_X) return KGX_OK;
  Timed tm(c, KGX_K_OTHER);
  const uint32_t items = std::max<uint32_t>(c->n_series, c->plan.n_nodes);
  const uint32_t grid = std::max(1u, std::min<uint32_t>(cdiv(items, 256), uint32_t(c->sm_count) * 8));
  k_clear_window<<<grid, 256, 0, c->stream>>>(acc_view(c), c->n_series, c->plan.n_nodes);
  tm.launches(1);
  return launch_check(c, "k_clear_window");
}

// CTAs that share one node's rows in the node-grouped join: 1 for a fleet; a lone busy node is split (about 2 K rows each)
uint32_t ingest_parts(const kgx_ctx* c, uint64_t n_rows) {
  const uint32_t ctas = uint32_t(c->sm_count) * 2;
  if (c->plan.n_nodes >= ctas) return 1;
  const uint64_t by_rows = std::max<uint64_t>(1, n_rows / std::max<uint32_t>(1, c->plan.n_nodes) / 2048);
  return uint32_t(std::min<uint64_t>(std::min<uint64_t>(by_rows, ctas / c->plan.n_nodes), 64));
}

// Launch one pass of the join over the batch `dc` (PASS 0: steady state; PASS 1: after an index rebuild, folds exactly
// the matches pass 0 had to skip).  `reset`: a window reset is folded into this pass (node-grouped X-mode batches only).
void launch_ingest_pass(kgx_ctx* c, const DevCols& dc, bool grouped, int pass, bool reset = false) {
  const Plan& p = c->plan;
  const bool x = c->mode == KGX_MODE_X;
  unsigned long long* ctr = ptr<unsigned long long>(c->d_counters);
  DevTable t{ptr<NodeTab>(c->d_node_tab), ptr<Entry>(c->d_tab), ptr<uint32_t>(c->d_pk_slot0), ptr<uint32_t>(c->d_pk_stride),
             ptr<uint32_t>(c->d_grank), p.n_nodes, p.G};
  DevIndex ix{ptr<uint2>(c->d_bm), ptr<uint32_t>(c->d_pending), c->n_words};
  const int grid = int(std::min<uint64_t>(uint64_t(c->sm_count) * 8, std::max<uint64_t>(1, cdiv(dc.n, 256))));
  const DevNodeMeta nm{ptr<uint32_t>(c->d_node_a_base), c->ingest_cells, c->ingest_pks, ptr<uint32_t>(c->d_node_slot_base), c->ingest_words,
                       ptr<uint32_t>(c->d_node_ser_off), reset ? 1u : 0u, ingest_parts(c, dc.n)};
  const int ngrid = int(std::min<uint32_t>(p.n_nodes * nm.parts, uint32_t(c->sm_count) * 2));
  if (grouped) {  // rows grouped by node: one CTA per node, table slice staged in shared memory
    if (x && pass == 0) k_ingest_nodes<true, 0><<<ngrid, kIngestThreads, c->ingest_smem, c->stream>>>(t, nm, ix, acc_view(c), dc, c->seq, ctr);
    else if (x) k_ingest_nodes<true, 1><<<ngrid, kIngestThreads, c->ingest_smem, c->stream>>>(t, nm, ix, acc_view(c), dc, c->seq, ctr);
    else if (pass == 0) k_ingest_nodes<false, 0><<<ngrid, kIngestThreads, c->ingest_smem, c->stream>>>(t, nm, ix, acc_view(c), dc, c->seq, ctr);
    else k_ingest_nodes<false, 1><<<ngrid, kIngestThreads, c->ingest_smem, c->stream>>>(t, nm, ix, acc_view(c), dc, c->seq, ctr);
  } else {
    if (x && pass == 0) k_ingest<true, 0><<<grid, 256, 0, c->stream>>>(t, ix, acc_view(c), dc, c->seq, ctr);
    else if (x) k_ingest<true, 1><<<grid, 256, 0, c->stream>>>(t, ix, acc_view(c), dc, c->seq, ctr);
    else if (pass == 0) k_ingest<false, 0><<<grid, 256, 0, c->stream>>>(t, ix, acc_view(c), dc, c->seq, ctr);
    else k_ingest<false, 1><<<grid, 256, 0, c->stream>>>(t, ix, acc_view(c), dc, c->seq, ctr);
  }
}

// `with_epilogue`: the launch also hands the join's statistics over; returns false in *did if nothing was launched
int launch_finalize_r(kgx_ctx* c, const DevCols& dc, bool grouped, bool with_epilogue = false, bool* did = nullptr) {
  if (did) *did = false;
  if (c->mode == KGX_MODE_X || !c->n_series || !dc.n) return KGX_OK;
  if (did) *did = true;
  Timed tm(c, KGX_K_FINALIZE);
  DevCols fc = dc;
  if (grouped) fc.node = nullptr;  // the node of a row is derived the way the join derived it: from node_off
  k_finalize_r<<<cdiv(c->n_series, 256), 256, 0, c->stream>>>(acc_view(c), fc, ptr<unsigned long long>(c->d_gpu_total), c->plan.G,
                                                              c->seq, c->n_series, c->plan.n_nodes,
                                                              with_epilogue ? ptr<unsigned long long>(c->d_counters) : nullptr, c->h_counters_dev);
  tm.launches(1);
  return launch_check(c, "k_finalize_r");
}

// The join of the pending batch saw series the index does not hold: merge them in, fold what pass 0 skipped.
int handle_churn(kgx_ctx* c) {
  KGX_TRY(rebuild_index(c));
  {
    Timed tm(c, KGX_K_INGEST);
    launch_ingest_pass(c, c->pending_cols, c->pending_grouped, 1);
    tm.launches(1);
    KGX_TRY(launch_check(c, "k_ingest churn pass"));
  }
  KGX_CUDA(c, cudaMemsetAsync(c->d_pending.p, 0, size_t(c->n_words) * 4, c->stream));
  KGX_CUDA(c, cudaMemsetAsync(ptr<unsigned long long>(c->d_counters) + C_CHURN, 0, 8, c->stream));
  return launch_finalize_r(c, c->pending_cols, c->pending_grouped);
}

// Statistics and churn check of the latest join (queued by ingest_device_locked): wait for it, read what the hand-over
// kernel behind it posted to mapped host memory, and run the churn path if it saw new series.
int settle_ingest(kgx_ctx* c) {
  if (!c->ingest_pending) return KGX_OK;
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  c->ingest_pending = false;
  c->stats.n_matched = c->h_counters[C_MATCHED];
  c->stats.n_dropped = c->h_counters[C_DROPPED];
  if (c->h_counters[C_NEW]) {
    KGX_TRY(handle_churn(c));
    KGX_CUDA(c, cudaStreamSynchronize(c->stream));  // the caller's device columns may go away after this call
  }
  return KGX_OK;
}

// Queue the join + reduce of one batch whose columns are device memory.  Returns with the work in flight
// (c->ingest_pending); settle_ingest() or the next render completes it.
int ingest_device_locked(kgx_ctx* c, const kgx_cols* s) {
  if (!c->have_table) return fail(c, KGX_E_STATE, "kgx_ingest before kgx_set_table");
  if (!s->gpu_total) return fail(c, KGX_E_INVAL, "cols: gpu_total missing");
  if (s->n && (!s->pid || !s->gpu || !s->used)) return fail(c, KGX_E_INVAL, "cols: pid/gpu/used missing");
  const bool x = c->mode == KGX_MODE_X;
  if (x && s->n && (!s->sm_util || !s->mem_util || !s->ts_us)) return fail(c, KGX_E_INVAL, "cols: X mode needs sm_util, mem_util, ts_us");
  if (s->n >= 0xffffffffull) return fail(c, KGX_E_LIMIT, "more than 2^32-1 samples in one ingest");
  KGX_TRY(settle_ingest(c));  // one batch in flight at a time: its churn pass needs its columns
  const Plan& p = c->plan;
  KGX_CUDA(c, cudaMemcpyAsync(c->d_gpu_total.p, s->gpu_total, size_t(p.n_nodes) * p.G * 8, cudaMemcpyDefault, c->stream));
  c->have_total = true;
  c->seq += 1;
  c->rendered = false;
  const DevCols dc{s->pid, s->gpu, s->node, s->sm_util, s->mem_util, reinterpret_cast<const unsigned long long*>(s->used),
                   reinterpret_cast<const unsigned long long*>(s->ts_us), reinterpret_cast<const unsigned long long*>(s->node_off), s->n};
  const bool grouped = s->node_off != nullptr && !getenv("KGX_INGEST_V1");
  {
    Timed tm(c, KGX_K_INGEST);
    // a pending window reset: the join clears each node as it takes it — if one CTA owns the node
    const bool fold_reset = c->reset_pending && x && grouped && s->n && c->n_series && ingest_parts(c, s->n) == 1;
    if (!fold_reset) KGX_TRY(flush_reset(c));
    c->reset_pending = false;
    if (s->n) launch_ingest_pass(c, dc, grouped, 0, fold_reset);
    else k_ingest_nothing<<<1, 32, 0, c->stream>>>(ptr<unsigned long long>(c->d_counters), c->h_counters_dev);
    tm.launches(1);
    KGX_TRY(launch_check(c, "k_ingest"));
  }
  // The statistics reach mapped host memory by a store of a kernel queued behind the join, not by a D2H copy: a copy
  // would queue on the device->host copy engine behind a body that kgx_render_begin may still be sending (tens of ms).
  bool handed_over = false;  // R modes: k_finalize_r posts them too
  KGX_TRY(launch_finalize_r(c, dc, grouped, true, &handed_over));
  if (s->n && !handed_over) {
    Timed tm(c, KGX_K_INGEST);
    k_ingest_epilogue<<<1, 32, 0, c->stream>>>(ptr<unsigned long long>(c->d_counters), c->h_counters_dev);
    tm.launches(1);
    KGX_TRY(launch_check(c, "k_ingest_epilogue"));
  }
  c->pending_cols = dc;
  c->pending_grouped = grouped;
  c->ingest_pending = true;
  return KGX_OK;
}

int launch_render(kgx_ctx* c, uint8_t* out, bool guard) {
  unsigned long long* ctr = ptr<unsigned long long>(c->d_counters);
  if (c->render_epoch >= kStEpochMax) {  // the 14-bit epoch tag wraps: clear the tagged arrays once, start over
    KGX_CUDA(c, cudaMemsetAsync(c->d_tile_state.p, 0, c->d_tile_state.cap, c->stream));
    KGX_CUDA(c, cudaMemsetAsync(c->d_node_body_off.p, 0xff, c->d_node_body_off.cap, c->stream));
    KGX_CUDA(c, cudaMemsetAsync(ptr<unsigned long long>(c->d_counters) + C_TICKET, 0, 16, c->stream));
    c->render_epoch = 0;
  }
  c->render_epoch += 1;
  c->body_epoch = c->render_epoch;
  DevRender R{};
  R.fam = ptr<FamMeta>(c->d_fam); R.n_families = uint32_t(c->fam.size());
  R.tiles = ptr<Tile>(c->d_tiles); R.n_tiles = c->n_tiles;
  R.lbl_off = ptr<unsigned long long>(c->d_lbl_off); R.lbl_blob = ptr<uint8_t>(c->d_lbl_blob);
  R.ser_len = ptr<uint32_t>(c->d_ser_len);
  R.ser_gidx = ptr<uint32_t>(c->d_ser_gidx); R.gpu_total = ptr<unsigned long long>(c->d_gpu_total);
  R.tile_state = ptr<unsigned long long>(c->d_tile_state); R.node_body_off = ptr<unsigned long long>(c->d_node_body_off);
  R.out = out; R.smem_bytes = c->render_smem;
  R.epoch = c->render_epoch; R.check_churn = guard ? 1u : 0u; R.host_len = c->h_counters_dev + C_TOTAL_LEN;
  Timed tm(c, KGX_K_RENDER);
  const int grid = int(std::min<uint32_t>(c->n_tiles, uint32_t(c->render_grid)));
  if (c->render_v2) launch_writer(c, grid, R, ctr);
  else k_render<<<grid, kRenderThreads, c->render_smem, c->stream>>>(R, acc_view(c), ctr);
  tm.launches(1);
  return launch_check(c, "k_render");
}

// `h_copy` (pinned, `h_cap` bytes): also queue the copy of the body's upper bound right behind the writer, so that a
// small body needs no second synchronisation; *copied = bytes that copy covered (0: none).
int render_device_locked(kgx_ctx* c, size_t* len, bool keep_copies = false, uint8_t* h_copy = nullptr, size_t h_cap = 0,
                         size_t* copied = nullptr) {
  if (copied) *copied = 0;
  auto queue_copy = [&]() -> int {
    if (!h_copy || c->out_bound == 0 || c->out_bound > h_cap) { if (copied) *copied = 0; return KGX_OK; }
    KGX_CUDA(c, cudaMemcpyAsync(h_copy, c->d_out.p, size_t(c->out_bound), cudaMemcpyDeviceToHost, c->stream));
    if (copied) *copied = size_t(c->out_bound);
    return KGX_OK;
  };
  if (!keep_copies) wait_copy(c);
  c->body_off_valid = false;
  // a scrape may arrive before the first poll cycle (the reference starts its HTTP server first, main.go:68-74):
  // an empty registry gathers nothing, so the body is empty
  if (!c->have_table) { c->last_len = 0; c->rendered = true; if (len) *len = 0; return KGX_OK; }
  KGX_TRY(flush_reset(c));  // (no-op unless kgx_reset_window was the last call)
  if (c->n_tiles == 0) KGX_TRY(settle_ingest(c));  // a first batch may bring the first series
  if (c->n_tiles == 0) { c->last_len = 0; c->rendered = true; if (len) *len = 0; c->stats.rendered_bytes = 0; return KGX_OK; }
  // A join may still be in flight (kgx_ingest_device).  The writer is queued behind it without a host round trip and
  // checks on the device that the join met no new series; if it did, the writer leaves at once and the churn path
  // runs here before a second launch.  Steady state: two kernels, one synchronisation.
  const bool guard = c->ingest_pending;
  KGX_TRY(launch_render(c, ptr<uint8_t>(c->d_out), guard));
  KGX_TRY(queue_copy());
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  if (guard) {
    c->ingest_pending = false;
    c->stats.n_matched = c->h_counters[C_MATCHED];
    c->stats.n_dropped = c->h_counters[C_DROPPED];
    if (c->h_counters[C_NEW]) {
      KGX_TRY(handle_churn(c));
      KGX_TRY(launch_render(c, ptr<uint8_t>(c->d_out), false));
      KGX_TRY(queue_copy());
      KGX_CUDA(c, cudaStreamSynchronize(c->stream));
    }
  }
  c->last_len = size_t(c->h_counters[C_TOTAL_LEN]);
  c->rendered = true;
  c->stats.rendered_bytes = c->last_len;
  if (len) *len = c->last_len;
  return KGX_OK;
}

// per-node offsets of the last rendered body, on the host (cached until the next render)
int fetch_body_offsets(kgx_ctx* c) {
  if (c->body_off_valid) return KGX_OK;
  const size_t NN = c->plan.n_nodes;
  c->body_off.assign(NN + 1, 0);
  if (c->have_table && c->n_tiles) {
    KGX_CUDA(c, cudaMemcpyAsync(c->body_off.data(), c->d_node_body_off.p, (NN + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
    KGX_CUDA(c, cudaStreamSynchronize(c->stream));
    c->body_off[NN] = c->last_len;
    for (size_t k = NN; k-- > 0;) {  // an entry counts if the latest launch wrote it; nodes without series have an empty body
      const uint64_t w = c->body_off[k];
      c->body_off[k] = (w >> kStEpochShift) == c->body_epoch ? (w & kStValue) : c->body_off[k + 1];
    }
  }
  c->body_off_valid = true;
  return KGX_OK;
}

}  // namespace

namespace {
// defined with kgx_render_noalloc at the end of this file: the device address of its in-library scratch array
int noalloc_scratch(kgx_ctx* c, unsigned long long** out);
}  // namespace

// =============================== C ABI ========================================
extern "C" {

int kgx_abi_version(void) { return KGX_ABI_VERSION; }

const char* kgx_strerror(int code) {
  int i = -code;
  return (i >= 0 && i < int(sizeof kErrNames / sizeof kErrNames[0])) ? kErrNames[i] : "unknown error";
}

const char* kgx_last_error(kgx_ctx* c) { return c ? tl_err.c_str() : "null ctx"; }

int kgx_create(int device, const kgx_opts* opts, kgx_ctx** out) {
  if (!out || !opts) return KGX_E_INVAL;
  *out = nullptr;
  if (opts->abi_version != KGX_ABI_VERSION || opts->mode < 0 || opts->mode > 2) return KGX_E_INVAL;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return KGX_E_CUDA;  // no CPU fallback
  if (cudaSetDevice(device) != cudaSuccess) return KGX_E_CUDA;
  kgx_ctx* c = new (std::nothrow) kgx_ctx();
  if (!c) return KGX_E_NOMEM;
  c->device = device; c->mode = opts->mode; c->profile = opts->profile != 0;
  c->device_build = getenv("KGX_HOST_TABLE_BUILD") == nullptr;
  if (opts->stream) c->stream = static_cast<cudaStream_t>(opts->stream);
  else { if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return KGX_E_CUDA; } c->own_stream = true; }
  cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device);
  build_family_meta(c);
  if (cudaHostAlloc(reinterpret_cast<void**>(&c->h_counters), C_COUNT * 8, cudaHostAllocMapped) != cudaSuccess ||
      cudaHostGetDevicePointer(reinterpret_cast<void**>(&c->h_counters_dev), c->h_counters, 0) != cudaSuccess ||
      dev_reserve(c, c->d_counters, kCounterWords * 8) != KGX_OK) { kgx_destroy(c); return KGX_E_CUDA; }
  cudaMemsetAsync(c->d_counters.p, 0, kCounterWords * 8, c->stream);
  memset(c->h_counters, 0, C_COUNT * 8);
  // set-up for kgx_render_noalloc: resolve the scratch array's address on this GPU now (this device is current), so
  // that the scrape-time call makes no symbol look-up; if it fails here the call itself tries again and reports it
  { unsigned long long* unused = nullptr; (void)noalloc_scratch(c, &unused); tl_err.clear(); }
  *out = c;
  return KGX_OK;
}

void kgx_destroy(kgx_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  wait_copy(c);
  free_table_state(c);
  dev_free(c, c->d_fam); dev_free(c, c->d_counters); dev_free(c, c->d_block_sums);
  for (auto& b : c->d_cols) dev_free(c, b);
  for (auto& b : c->d_tmp) dev_free(c, b);
  for (auto& r : c->regions) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  if (c->h_counters) cudaFreeHost(c->h_counters);
  if (c->h_small_in) cudaFreeHost(c->h_small_in);
  if (c->h_small_out) cudaFreeHost(c->h_small_out);
  dev_free(c, c->d_small_in);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  if (c->in_stream) cudaStreamDestroy(c->in_stream);
  if (c->ev_rendered) cudaEventDestroy(c->ev_rendered);
  if (c->ev_copy_out) cudaEventDestroy(c->ev_copy_out);
  if (c->ev_copy_alt) cudaEventDestroy(c->ev_copy_alt);
  if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

int kgx_set_table(kgx_ctx* c, const kgx_table* table) {
  if (!c) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (!table) return fail(c, KGX_E_INVAL, "null table");
  KGX_TRY(settle_ingest(c));
  c->reset_pending = false;  // X mode: a new table clears the window anyway
  if (c->have_table && c->mode != KGX_MODE_X &&
      (table->n_nodes != c->plan.n_nodes || table->gpus_per_node != c->plan.G))
    return fail(c, KGX_E_INVAL, "R modes: n_nodes/gpus_per_node cannot change while series persist");
  // KGX_TRACE=1: where the cold path spends its time (stderr)
  const bool trace = getenv("KGX_TRACE") != nullptr;
  auto now = [] { return std::chrono::steady_clock::now(); };
  auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
    return std::chrono::duration<double, std::milli>(b - a).count();
  };
  const auto t0 = now();
  if (c->device_build) {  // the plan is built on the GPU (kgx_build.cuh); KGX_HOST_TABLE_BUILD=1 at kgx_create selects the host builder
    CarriedDev cd;
    KGX_TRY(set_table_device(c, table, &cd));
    const auto t1 = now();
    reset_table_state(c);
    c->have_table = false;  // the old table's buffers are gone; the new one counts only once its geometry is set
    KGX_TRY(finish_plan(c));
    c->have_table = true;
    c->table_gen += 1;
    c->stats.n_slots = c->plan.n_slots;
    if (cd.n) {  // R modes: series that already had a value keep it (main.go never deletes a GaugeVec child)
      k_set_bits<<<cdiv(cd.n, 256), 256, 0, c->stream>>>(cd.slot, cd.n, ptr<uint32_t>(c->d_pending));
      KGX_TRY(rebuild_index(c));
      k_seed_carried<<<cdiv(cd.n, 256), 256, 0, c->stream>>>(cd.slot, cd.used, cd.total, cd.n, ptr<uint2>(c->d_bm), acc_view(c));
      c->stats.kernel_launches += 2;
      KGX_CUDA(c, cudaMemsetAsync(c->d_pending.p, 0, size_t(c->n_words) * 4, c->stream));
      KGX_TRY(launch_check(c, "carry-over"));
      KGX_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    {  // scratch of a small table is kept for the next cycle (churn); a fleet-sized one is given back
      size_t held = 0;
      for (const DBuf& b : c->arena) held += b.cap;
      if (held > (size_t(256) << 20)) {
        KGX_CUDA(c, cudaStreamSynchronize(c->stream));
        for (DBuf& b : c->arena) dev_free(c, b);
        c->arena.clear();
      }
    }
    if (trace)
      fprintf(stderr, "kgx_set_table (device build): plan %.2f ms (%u carried series), index + re-seed %.2f ms\n", ms(t0, t1), cd.n,
              ms(t1, now()));
    return KGX_OK;
  }
  std::vector<Carried> carried;
  KGX_TRY(collect_carried(c, &carried));
  const auto t1 = now();
  Plan np;
  std::string err;
  int rc = build_plan(c->mode, table, carried, &np, &err);
  if (rc != KGX_OK) { tl_err = err; return rc; }
  const auto t2 = now();
  reset_table_state(c);
  c->plan = std::move(np);
  c->have_table = false;  // see the device-build branch
  KGX_TRY(upload_plan(c));
  KGX_TRY(finish_plan(c));
  c->have_table = true;
  c->table_gen += 1;
  const auto t3 = now();
  c->stats.n_slots = c->plan.n_slots;
  if (!carried.empty()) {
    const uint32_t n = uint32_t(carried.size());
    std::vector<unsigned long long> used(n), total(n);
    for (uint32_t i = 0; i < n; ++i) { used[i] = carried[i].used; total[i] = carried[i].total; }
    KGX_TRY(upload(c, c->d_tmp[1], c->plan.carried_slot.data(), n));
    KGX_TRY(upload(c, c->d_tmp[2], used.data(), n));
    KGX_TRY(upload(c, c->d_tmp[3], total.data(), n));
    k_set_bits<<<cdiv(n, 256), 256, 0, c->stream>>>(ptr<uint32_t>(c->d_tmp[1]), n, ptr<uint32_t>(c->d_pending));
    KGX_TRY(rebuild_index(c));
    k_seed_carried<<<cdiv(n, 256), 256, 0, c->stream>>>(ptr<uint32_t>(c->d_tmp[1]), ptr<unsigned long long>(c->d_tmp[2]),
                                                        ptr<unsigned long long>(c->d_tmp[3]), n, ptr<uint2>(c->d_bm), acc_view(c));
    c->stats.kernel_launches += 2;
    KGX_CUDA(c, cudaMemsetAsync(c->d_pending.p, 0, size_t(c->n_words) * 4, c->stream));
    KGX_TRY(launch_check(c, "carry-over"));
    KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  if (trace)
    fprintf(stderr, "kgx_set_table: read back %zu series %.2f ms, host plan %.2f ms, upload %.2f ms, re-seed %.2f ms\n",
            carried.size(), ms(t0, t1), ms(t1, t2), ms(t2, t3), ms(t3, now()));
  return KGX_OK;
}

int kgx_ingest_device(kgx_ctx* c, const kgx_cols* cols) {
  if (!c) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (!cols) return fail(c, KGX_E_INVAL, "null cols");
  return ingest_device_locked(c, cols);  // in flight on return: settled by the next call (see kgx.h)
}

int kgx_ingest(kgx_ctx* c, const kgx_cols* cols) {
  if (!c) return KGX_E_INVAL;
  std::lock_guard<std::mutex> gin(c->mu_in);  // one batch is staged at a time
  cudaSetDevice(c->device);
  if (!cols) return fail(c, KGX_E_INVAL, "null cols");
  uint32_t NN = 0, G = 0;
  uint64_t gen = 0;
  {
    std::lock_guard<std::mutex> g(c->mu);
    if (!c->have_table) return fail(c, KGX_E_STATE, "kgx_ingest before kgx_set_table");
    KGX_TRY(settle_ingest(c));  // a kgx_ingest_device batch still in flight may be reading d_cols
    NN = c->plan.n_nodes; G = c->plan.G; gen = c->table_gen;
  }
  if (!cols->gpu_total) return fail(c, KGX_E_INVAL, "cols: gpu_total missing");
  kgx_cols d = *cols;
  const size_t n = size_t(cols->n);
  if (cols->node_off) {  // the grouping promise is checked here, where the offsets are still host memory
    if (cols->node_off[0] != 0 || cols->node_off[NN] != n) return fail(c, KGX_E_INVAL, "cols: node_off must start at 0 and end at n");
    for (uint32_t k = 0; k < NN; ++k)
      if (cols->node_off[k + 1] < cols->node_off[k]) return fail(c, KGX_E_INVAL, "cols: node_off not ascending");
  }
  // Stage the batch on the copy-in stream, WITHOUT the state mutex: scrapes go on rendering the last completed
  // reduction while hundreds of megabytes cross PCIe (the reference's HTTP goroutines likewise never wait for the poll
  // loop's I/O, main.go:68-74).
  if (!c->in_stream) KGX_CUDA(c, cudaStreamCreateWithFlags(&c->in_stream, cudaStreamNonBlocking));
  const void* src[9] = {cols->pid, cols->gpu, cols->used, cols->node, cols->sm_util, cols->mem_util, cols->ts_us, cols->node_off, cols->gpu_total};
  const size_t bytes[9] = {n * 4, n * 4, n * 8, n * 4, n * 4, n * 4, n * 8, (size_t(NN) + 1) * 8, size_t(NN) * G * 8};
  const void* dst[9];
  size_t packed = 0, at[9];
  for (int k = 0; k < 9; ++k) { at[k] = packed; if (src[k]) packed += (bytes[k] + 15) & ~size_t(15); }
  if (packed <= kSmallBytes && !c->h_small_in &&
      cudaHostAlloc(reinterpret_cast<void**>(&c->h_small_in), kSmallBytes, cudaHostAllocDefault) != cudaSuccess) {
    c->h_small_in = nullptr;
    cudaGetLastError();
  }
  if (packed <= kSmallBytes && c->h_small_in) {  // a small batch: pack the columns, one copy instead of nine
    KGX_TRY(dev_reserve(c, c->d_small_in, kSmallBytes));
    for (int k = 0; k < 9; ++k) {
      dst[k] = nullptr;
      if (!src[k]) continue;
      if (bytes[k]) memcpy(c->h_small_in + at[k], src[k], bytes[k]);
      dst[k] = static_cast<uint8_t*>(c->d_small_in.p) + at[k];
    }
    if (packed) KGX_CUDA(c, cudaMemcpyAsync(c->d_small_in.p, c->h_small_in, packed, cudaMemcpyHostToDevice, c->in_stream));
  } else {
    for (int k = 0; k < 9; ++k) {
      dst[k] = nullptr;
      if (!src[k]) continue;
      KGX_TRY(dev_reserve(c, c->d_cols[k], bytes[k]));
      if (bytes[k]) KGX_CUDA(c, cudaMemcpyAsync(c->d_cols[k].p, src[k], bytes[k], cudaMemcpyHostToDevice, c->in_stream));
      dst[k] = c->d_cols[k].p;
    }
  }
  KGX_CUDA(c, cudaStreamSynchronize(c->in_stream));  // the caller's buffers are free from here on
  d.pid = static_cast<const uint32_t*>(dst[0]); d.gpu = static_cast<const uint32_t*>(dst[1]);
  d.used = static_cast<const uint64_t*>(dst[2]); d.node = static_cast<const uint32_t*>(dst[3]);
  d.sm_util = static_cast<const uint32_t*>(dst[4]); d.mem_util = static_cast<const uint32_t*>(dst[5]);
  d.ts_us = static_cast<const uint64_t*>(dst[6]); d.node_off = static_cast<const uint64_t*>(dst[7]);
  d.gpu_total = static_cast<const uint64_t*>(dst[8]);
  std::lock_guard<std::mutex> g(c->mu);
  if (c->table_gen != gen || !c->have_table) return fail(c, KGX_E_STATE, "kgx_set_table ran while this batch was being copied in");
  KGX_TRY(ingest_device_locked(c, &d));
  return settle_ingest(c);  // the statistics are final when this returns
}

int kgx_ingest_nvml(kgx_ctx* c, const kgx_nvml_process_info* infos, const uint64_t* dev_off, const uint64_t* gpu_total) {
  if (!c) return KGX_E_INVAL;
  std::lock_guard<std::mutex> gin(c->mu_in);  // the staging buffers are shared with kgx_ingest
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (c->mode == KGX_MODE_X) return fail(c, KGX_E_STATE, "kgx_ingest_nvml: nvmlProcessInfo_t carries no utilisation or timestamp; R modes only");
  if (!c->have_table) return fail(c, KGX_E_STATE, "kgx_ingest_nvml before kgx_set_table");
  if (!dev_off || !gpu_total) return fail(c, KGX_E_INVAL, "kgx_ingest_nvml: dev_off / gpu_total missing");
  static_assert(sizeof(kgx_nvml_process_info) == 24 && sizeof(NvmlProcessInfo) == 24, "nvmlProcessInfo_t is 24 bytes");
  KGX_TRY(settle_ingest(c));
  const uint32_t n_dev = c->plan.n_nodes * c->plan.G;
  const size_t n = size_t(dev_off[n_dev]);
  if (dev_off[0] != 0) return fail(c, KGX_E_INVAL, "kgx_ingest_nvml: dev_off[0] must be 0");
  for (uint32_t d = 0; d < n_dev; ++d)
    if (dev_off[d + 1] < dev_off[d]) return fail(c, KGX_E_INVAL, "kgx_ingest_nvml: dev_off not ascending");
  if (n && !infos) return fail(c, KGX_E_INVAL, "kgx_ingest_nvml: infos missing");
  KGX_TRY(dev_reserve(c, c->d_tmp[1], n * 24));
  KGX_TRY(dev_reserve(c, c->d_tmp[2], (size_t(n_dev) + 1) * 8));
  const size_t bytes[8] = {n * 4, n * 4, n * 8, n * 4, 0, 0, 0, (size_t(c->plan.n_nodes) + 1) * 8};
  for (int k : {0, 1, 2, 3, 7}) KGX_TRY(dev_reserve(c, c->d_cols[k], bytes[k]));
  if (n) KGX_CUDA(c, cudaMemcpyAsync(c->d_tmp[1].p, infos, n * 24, cudaMemcpyHostToDevice, c->stream));
  KGX_CUDA(c, cudaMemcpyAsync(c->d_tmp[2].p, dev_off, (size_t(n_dev) + 1) * 8, cudaMemcpyHostToDevice, c->stream));
  {
    Timed tm(c, KGX_K_OTHER);
    if (n) k_unpack_nvml<<<cdiv(n, 256), 256, 0, c->stream>>>(ptr<NvmlProcessInfo>(c->d_tmp[1]), ptr<unsigned long long>(c->d_tmp[2]), n_dev,
                                                           c->plan.G, n, ptr<uint32_t>(c->d_cols[0]), ptr<uint32_t>(c->d_cols[1]),
                                                           ptr<uint32_t>(c->d_cols[3]), ptr<unsigned long long>(c->d_cols[2]));
    k_node_off_from_dev_off<<<cdiv(c->plan.n_nodes + 1, 256), 256, 0, c->stream>>>(ptr<unsigned long long>(c->d_tmp[2]), c->plan.n_nodes,
                                                                                  c->plan.G, ptr<unsigned long long>(c->d_cols[7]));
    tm.launches(n ? 2 : 1);
    KGX_TRY(launch_check(c, "k_unpack_nvml"));
  }
  kgx_cols d{};
  d.n = n;
  d.pid = ptr<uint32_t>(c->d_cols[0]); d.gpu = ptr<uint32_t>(c->d_cols[1]); d.used = ptr<uint64_t>(c->d_cols[2]);
  d.node = ptr<uint32_t>(c->d_cols[3]); d.node_off = ptr<uint64_t>(c->d_cols[7]); d.gpu_total = gpu_total;
  KGX_TRY(ingest_device_locked(c, &d));
  return settle_ingest(c);  // the caller's buffers are free and the statistics final when this returns
}

int kgx_ingest_nvml_x(kgx_ctx* c, const kgx_nvml_process_info* infos, const uint64_t* dev_off,
                      const kgx_nvml_process_utilization_sample* utils, const uint64_t* util_off, const uint64_t* gpu_total) {
  if (!c) return KGX_E_INVAL;
  std::lock_guard<std::mutex> gin(c->mu_in);  // the staging buffers are shared with kgx_ingest
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (c->mode != KGX_MODE_X) return fail(c, KGX_E_STATE, "kgx_ingest_nvml_x: X mode only (R modes take kgx_ingest_nvml)");
  if (!c->have_table) return fail(c, KGX_E_STATE, "kgx_ingest_nvml_x before kgx_set_table");
  if (!dev_off || !util_off || !gpu_total) return fail(c, KGX_E_INVAL, "kgx_ingest_nvml_x: dev_off / util_off / gpu_total missing");
  static_assert(sizeof(kgx_nvml_process_utilization_sample) == 32 && sizeof(NvmlProcessUtilizationSample) == 32,
                "nvmlProcessUtilizationSample_t is 32 bytes");
  KGX_TRY(settle_ingest(c));
  const uint32_t n_dev = c->plan.n_nodes * c->plan.G;
  if (dev_off[0] != 0 || util_off[0] != 0) return fail(c, KGX_E_INVAL, "kgx_ingest_nvml_x: dev_off[0] and util_off[0] must be 0");
  for (uint32_t d = 0; d < n_dev; ++d)
    if (dev_off[d + 1] < dev_off[d] || util_off[d + 1] < util_off[d]) return fail(c, KGX_E_INVAL, "kgx_ingest_nvml_x: offsets not ascending");
  const size_t n = size_t(dev_off[n_dev]), m = size_t(util_off[n_dev]);
  if ((n && !infos) || (m && !utils)) return fail(c, KGX_E_INVAL, "kgx_ingest_nvml_x: infos / utils missing");
  KGX_TRY(dev_reserve(c, c->d_tmp[1], n * 24));
  KGX_TRY(dev_reserve(c, c->d_tmp[2], (size_t(n_dev) + 1) * 8));
  KGX_TRY(dev_reserve(c, c->d_tmp[4], m * 32));
  KGX_TRY(dev_reserve(c, c->d_tmp[5], (size_t(n_dev) + 1) * 8));
  const size_t bytes[8] = {n * 4, n * 4, n * 8, n * 4, n * 4, n * 4, n * 8, (size_t(c->plan.n_nodes) + 1) * 8};
  for (int k = 0; k < 8; ++k) KGX_TRY(dev_reserve(c, c->d_cols[k], bytes[k]));
  if (n) KGX_CUDA(c, cudaMemcpyAsync(c->d_tmp[1].p, infos, n * 24, cudaMemcpyHostToDevice, c->stream));
  if (m) KGX_CUDA(c, cudaMemcpyAsync(c->d_tmp[4].p, utils, m * 32, cudaMemcpyHostToDevice, c->stream));
  KGX_CUDA(c, cudaMemcpyAsync(c->d_tmp[2].p, dev_off, (size_t(n_dev) + 1) * 8, cudaMemcpyHostToDevice, c->stream));
  KGX_CUDA(c, cudaMemcpyAsync(c->d_tmp[5].p, util_off, (size_t(n_dev) + 1) * 8, cudaMemcpyHostToDevice, c->stream));
  {
    Timed tm(c, KGX_K_OTHER);
    if (n) {
      k_unpack_nvml<<<cdiv(n, 256), 256, 0, c->stream>>>(ptr<NvmlProcessInfo>(c->d_tmp[1]), ptr<unsigned long long>(c->d_tmp[2]), n_dev, c->plan.G,
                                                         n, ptr<uint32_t>(c->d_cols[0]), ptr<uint32_t>(c->d_cols[1]), ptr<uint32_t>(c->d_cols[3]),
                                                         ptr<unsigned long long>(c->d_cols[2]));
      const int grid = int(std::min<uint32_t>(n_dev, uint32_t(c->sm_count) * 8));
      k_join_nvml_util<<<grid, kUtilThreads, 0, c->stream>>>(ptr<unsigned long long>(c->d_tmp[2]), ptr<NvmlProcessUtilizationSample>(c->d_tmp[4]),
                                                             ptr<unsigned long long>(c->d_tmp[5]), n_dev, ptr<uint32_t>(c->d_cols[0]),
                                                             ptr<uint32_t>(c->d_cols[4]), ptr<uint32_t>(c->d_cols[5]),
                                                             ptr<unsigned long long>(c->d_cols[6]));
    }
    k_node_off_from_dev_off<<<cdiv(c->plan.n_nodes + 1, 256), 256, 0, c->stream>>>(ptr<unsigned long long>(c->d_tmp[2]), c->plan.n_nodes,
                                                                                  c->plan.G, ptr<unsigned long long>(c->d_cols[7]));
    tm.launches(n ? 3 : 1);
    KGX_TRY(launch_check(c, "k_join_nvml_util"));
  }
  kgx_cols d{};
  d.n = n;
  d.pid = ptr<uint32_t>(c->d_cols[0]); d.gpu = ptr<uint32_t>(c->d_cols[1]); d.used = ptr<uint64_t>(c->d_cols[2]);
  d.node = ptr<uint32_t>(c->d_cols[3]); d.sm_util = ptr<uint32_t>(c->d_cols[4]); d.mem_util = ptr<uint32_t>(c->d_cols[5]);
  d.ts_us = ptr<uint64_t>(c->d_cols[6]); d.node_off = ptr<uint64_t>(c->d_cols[7]); d.gpu_total = gpu_total;
  KGX_TRY(ingest_device_locked(c, &d));
  return settle_ingest(c);  // the caller's buffers are free and the statistics final when this returns
}

int kgx_reset_window(kgx_ctx* c) {
  if (!c) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (c->mode != KGX_MODE_X) return fail(c, KGX_E_STATE, "R modes never reset: GaugeVec children persist");
  if (!c->have_table) return fail(c, KGX_E_STATE, "kgx_reset_window before kgx_set_table");
  KGX_TRY(settle_ingest(c));
  // carried out by the next node-grouped join as it goes (each CTA clears the node it owns), or by whatever call
  // comes first and needs the state (flush_reset)
  c->reset_pending = true;
  c->rendered = false;
  return KGX_OK;
}

int kgx_expire(kgx_ctx* c, uint64_t older_than_ts_us, uint64_t* n_expired) {
  if (!c) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (n_expired) *n_expired = 0;
  if (c->mode != KGX_MODE_X) return fail(c, KGX_E_STATE, "R modes never expire a series: main.go has no Delete/Reset");
  if (!c->have_table) return fail(c, KGX_E_STATE, "kgx_expire before kgx_set_table");
  KGX_TRY(settle_ingest(c));
  KGX_TRY(flush_reset(c));
  if (c->n_series == 0) return KGX_OK;
  unsigned long long* ctr = ptr<unsigned long long>(c->d_counters) + kExpireWords;  // two spare counter words
  KGX_CUDA(c, cudaMemsetAsync(ctr, 0, 16, c->stream));
  {
    Timed tm(c, KGX_K_INDEX);
    const DevAcc a = acc_view(c);
    k_mark_expired<<<cdiv(c->n_series, 256), 256, 0, c->stream>>>(ptr<uint32_t>(c->d_ser_slot), c->n_series, a.xa, a.xb,
                                                                 older_than_ts_us, ptr<uint2>(c->d_bm), ctr);
    tm.launches(1);
    KGX_TRY(launch_check(c, "k_mark_expired"));
  }
  unsigned long long out[2] = {0, 0};
  KGX_CUDA(c, cudaMemcpyAsync(out, ctr, 16, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  if (n_expired) *n_expired = out[0];
  if (out[1] == 0) return KGX_OK;
  c->rendered = false;
  return rebuild_index(c);  // re-rank the survivors, move their state, re-materialise labels and tiles
}

int kgx_render_device(kgx_ctx* c, size_t* len) {
  if (!c) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  return render_device_locked(c, len);
}

int kgx_render(kgx_ctx* c, uint8_t* out, size_t cap, size_t* len) {
  if (!c || !len) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (!c->h_small_out && cudaHostAlloc(reinterpret_cast<void**>(&c->h_small_out), kSmallBytes, cudaHostAllocDefault) != cudaSuccess) {
    c->h_small_out = nullptr;
    cudaGetLastError();  // no pinned memory: the two-step path below serves every size
  }
  size_t copied = 0;
  KGX_TRY(render_device_locked(c, len, false, c->h_small_out, kSmallBytes, &copied));
  if (*len > cap) return fail(c, KGX_E_NOSPACE, "output buffer too small");
  if (*len) {
    if (!out) return fail(c, KGX_E_INVAL, "null output buffer");
    if (*len <= copied) {  // a small body: it arrived with the writer's own synchronisation
      memcpy(out, c->h_small_out, *len);
    } else {
      KGX_CUDA(c, cudaMemcpyAsync(out, c->d_out.p, *len, cudaMemcpyDeviceToHost, c->stream));
      KGX_CUDA(c, cudaStreamSynchronize(c->stream));
    }
  }
  return KGX_OK;
}

int kgx_render_begin(kgx_ctx* c, uint8_t* out, size_t cap, size_t* len) {
  if (!c || !len) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (!c->copy_stream) {
    KGX_CUDA(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    KGX_CUDA(c, cudaEventCreateWithFlags(&c->ev_rendered, cudaEventDisableTiming));
    KGX_CUDA(c, cudaEventCreateWithFlags(&c->ev_copy_out, cudaEventDisableTiming));
    KGX_CUDA(c, cudaEventCreateWithFlags(&c->ev_copy_alt, cudaEventDisableTiming));
  }
  if (c->out_busy) {  // the previous body is still leaving from d_out: this one is rendered into the other buffer
    if (c->alt_busy) { KGX_CUDA(c, cudaEventSynchronize(c->ev_copy_alt)); c->alt_busy = false; }  // the one before that
    if (c->d_out_alt.cap < c->d_out.cap) {
      dev_free(c, c->d_out_alt);
      KGX_CUDA(c, cudaMalloc(&c->d_out_alt.p, c->d_out.cap));
      c->d_out_alt.cap = c->d_out.cap;
      c->device_bytes += c->d_out.cap;
    }
    std::swap(c->d_out, c->d_out_alt);
    std::swap(c->out_busy, c->alt_busy);
    std::swap(c->ev_copy_out, c->ev_copy_alt);
  }
  KGX_TRY(render_device_locked(c, len, true));
  if (*len > cap) return fail(c, KGX_E_NOSPACE, "output buffer too small");
  if (*len) {
    if (!out) return fail(c, KGX_E_INVAL, "null output buffer");
    KGX_CUDA(c, cudaEventRecord(c->ev_rendered, c->stream));
    KGX_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_rendered, 0));
    KGX_CUDA(c, cudaMemcpyAsync(out, c->d_out.p, *len, cudaMemcpyDeviceToHost, c->copy_stream));
    KGX_CUDA(c, cudaEventRecord(c->ev_copy_out, c->copy_stream));
    c->out_busy = true;
  }
  return KGX_OK;
}

int kgx_render_end(kgx_ctx* c) {
  if (!c) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (c->alt_busy) {  // the older of two bodies in flight
    KGX_CUDA(c, cudaEventSynchronize(c->ev_copy_alt));
    c->alt_busy = false;
  } else if (c->out_busy) {
    KGX_CUDA(c, cudaEventSynchronize(c->ev_copy_out));
    c->out_busy = false;
  }
  return KGX_OK;
}

int kgx_body_device(kgx_ctx* c, const uint8_t** dptr, size_t* len) {
  if (!c || !dptr || !len) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  if (!c->rendered) return fail(c, KGX_E_STATE, "no rendered body");
  *dptr = ptr<uint8_t>(c->d_out); *len = c->last_len;
  return KGX_OK;
}

int kgx_body_offsets(kgx_ctx* c, uint64_t* off, size_t n) {
  if (!c || !off) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (!c->rendered) return fail(c, KGX_E_STATE, "no rendered body");
  if (n != size_t(c->plan.n_nodes) + 1) return fail(c, KGX_E_INVAL, "offset array must have n_nodes+1 entries");
  KGX_TRY(fetch_body_offsets(c));
  std::copy(c->body_off.begin(), c->body_off.end(), off);
  return KGX_OK;
}

int kgx_body_node(kgx_ctx* c, uint32_t node, uint8_t* out, size_t cap, size_t* len) {
  if (!c || !len) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (!c->rendered) return fail(c, KGX_E_STATE, "no rendered body");
  if (!c->have_table) { *len = 0; return KGX_OK; }  // nothing registered yet: every node's body is empty
  if (node >= c->plan.n_nodes) return fail(c, KGX_E_INVAL, "node out of range");
  KGX_TRY(fetch_body_offsets(c));
  const uint64_t lo = c->body_off[node], hi = c->body_off[size_t(node) + 1];
  *len = size_t(hi - lo);
  if (*len > cap) return fail(c, KGX_E_NOSPACE, "output buffer too small");
  if (*len) {
    if (!out) return fail(c, KGX_E_INVAL, "null output buffer");
    KGX_CUDA(c, cudaMemcpyAsync(out, ptr<uint8_t>(c->d_out) + lo, *len, cudaMemcpyDeviceToHost, c->stream));
    KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  return KGX_OK;
}

int kgx_body_read(kgx_ctx* c, uint64_t offset, uint8_t* out, size_t cap, size_t* n) {
  if (!c || !n) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  *n = 0;
  if (!c->rendered) return fail(c, KGX_E_STATE, "no rendered body");
  if (offset > c->last_len) return fail(c, KGX_E_INVAL, "offset beyond the end of the body");
  const size_t take = size_t(std::min<uint64_t>(cap, c->last_len - offset));
  if (take) {
    if (!out) return fail(c, KGX_E_INVAL, "null output buffer");
    KGX_CUDA(c, cudaMemcpyAsync(out, ptr<uint8_t>(c->d_out) + offset, take, cudaMemcpyDeviceToHost, c->stream));
    KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  *n = take;
  return KGX_OK;
}

int kgx_family_count(kgx_ctx* c) { return c ? int(c->fam.size()) : KGX_E_INVAL; }

int kgx_family_info(kgx_ctx* c, int family, const char** name, const char** help, const char* const** names, int* n_labels) {
  if (!c || family < 0 || family >= int(c->fam.size())) return KGX_E_INVAL;
  const FamilyDef& f = families(c->mode)[size_t(family)];
  if (name) *name = f.name;
  if (help) *help = f.help;
  int nl = 0;
  const char* const* ln = label_names(c->mode, &nl);
  if (names) *names = ln;
  if (n_labels) *n_labels = nl;
  return KGX_OK;
}

int kgx_read_values(kgx_ctx* c, int family, double* values, uint32_t* slots, size_t cap, size_t* n_out) {
  if (!c || !n_out) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (!c->have_table) { *n_out = 0; return KGX_OK; }  // nothing registered yet: nothing to collect
  if (family < 0 || family >= int(c->fam.size())) return fail(c, KGX_E_INVAL, "family index out of range");
  KGX_TRY(settle_ingest(c));
  KGX_TRY(flush_reset(c));
  const uint32_t S = c->n_series;
  *n_out = 0;
  if (!S) return KGX_OK;
  KGX_TRY(dev_reserve(c, c->d_tmp[1], size_t(S) * 8));
  KGX_TRY(dev_reserve(c, c->d_tmp[2], size_t(S)));
  DevRender R{};
  R.ser_gidx = ptr<uint32_t>(c->d_ser_gidx); R.gpu_total = ptr<unsigned long long>(c->d_gpu_total);
  k_values<<<cdiv(S, 256), 256, 0, c->stream>>>(R, acc_view(c), c->fam[size_t(family)].kind, S, ptr<double>(c->d_tmp[1]),
                                                ptr<uint8_t>(c->d_tmp[2]));
  c->stats.kernel_launches += 1;
  KGX_TRY(launch_check(c, "k_values"));
  std::vector<double> v(S);
  std::vector<uint8_t> live(S);
  std::vector<uint32_t> slot(S);
  KGX_CUDA(c, cudaMemcpyAsync(v.data(), c->d_tmp[1].p, size_t(S) * 8, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaMemcpyAsync(live.data(), c->d_tmp[2].p, S, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaMemcpyAsync(slot.data(), c->d_ser_slot.p, size_t(S) * 4, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  size_t n = 0;
  for (uint32_t r = 0; r < S; ++r) {
    if (!live[r]) continue;
    if (n < cap) { if (values) values[n] = v[r]; if (slots) slots[n] = slot[r]; }
    ++n;
  }
  *n_out = n;
  return n <= cap ? KGX_OK : fail(c, KGX_E_NOSPACE, "value buffer too small");
}

int kgx_read_labels(kgx_ctx* c, char* blob, size_t blob_cap, uint64_t* off, size_t off_cap, size_t* n_out, size_t* blob_len) {
  if (!c || !n_out || !blob_len) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  *n_out = 0; *blob_len = 0;
  KGX_TRY(settle_ingest(c));
  KGX_TRY(flush_reset(c));
  if (!c->have_table || !c->n_series) { if (off && off_cap) off[0] = 0; return KGX_OK; }
  const uint32_t S = c->n_series;
  // which series carry a value (the same rule as kgx_read_values; it does not depend on the family)
  KGX_TRY(dev_reserve(c, c->d_tmp[1], size_t(S) * 8));
  KGX_TRY(dev_reserve(c, c->d_tmp[2], size_t(S)));
  DevRender R{};
  R.ser_gidx = ptr<uint32_t>(c->d_ser_gidx); R.gpu_total = ptr<unsigned long long>(c->d_gpu_total);
  k_values<<<cdiv(S, 256), 256, 0, c->stream>>>(R, acc_view(c), c->fam[0].kind, S, ptr<double>(c->d_tmp[1]), ptr<uint8_t>(c->d_tmp[2]));
  c->stats.kernel_launches += 1;
  KGX_TRY(launch_check(c, "k_values"));
  std::vector<uint8_t> live(S);
  std::vector<uint64_t> all_off(size_t(S) + 1);
  KGX_CUDA(c, cudaMemcpyAsync(live.data(), c->d_tmp[2].p, S, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaMemcpyAsync(all_off.data(), c->d_lbl_off.p, (size_t(S) + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  size_t n = 0, bytes = 0;
  for (uint32_t r = 0; r < S; ++r)
    if (live[r]) { ++n; bytes += size_t(all_off[r + 1] - all_off[r]); }
  *n_out = n; *blob_len = bytes;
  if (!blob && !off) return KGX_OK;  // size query
  if (!blob || !off || blob_cap < bytes || off_cap < n + 1) return fail(c, KGX_E_NOSPACE, "label buffers too small (sizes returned)");
  if (n == S) {  // every series has a value: the device blob is already the answer, in exposition order
    std::copy(all_off.begin(), all_off.end(), off);
    if (bytes) KGX_CUDA(c, cudaMemcpyAsync(blob, c->d_lbl_blob.p, bytes, cudaMemcpyDeviceToHost, c->stream));
    KGX_CUDA(c, cudaStreamSynchronize(c->stream));
    return KGX_OK;
  }
  // some series of the index carry no value in this window: copy maximal runs of live series
  size_t k = 0, at = 0;
  for (uint32_t r = 0; r < S;) {
    if (!live[r]) { ++r; continue; }
    uint32_t e = r;
    while (e < S && live[e]) ++e;
    const size_t run = size_t(all_off[e] - all_off[r]);
    if (run) KGX_CUDA(c, cudaMemcpyAsync(blob + at, ptr<uint8_t>(c->d_lbl_blob) + all_off[r], run, cudaMemcpyDeviceToHost, c->stream));
    for (uint32_t q = r; q < e; ++q) off[k++] = at + (all_off[q] - all_off[r]);
    at += run;
    r = e;
  }
  off[k] = at;
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  return KGX_OK;
}

int kgx_slot_labels(kgx_ctx* c, uint32_t slot, char* out, size_t cap, size_t* len) {
  if (!c || !len) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  if (!c->have_table || slot >= c->plan.n_slots) return fail(c, KGX_E_INVAL, "slot out of range");
  if (c->device_build) {  // the label pieces only exist on the GPU: fetch the three of this slot
    cudaSetDevice(c->device);
    if (c->plan.runs.size() != c->n_runs) {
      c->plan.runs.resize(c->n_runs);
      KGX_CUDA(c, cudaMemcpyAsync(c->plan.runs.data(), c->d_runs.p, size_t(c->n_runs) * sizeof(Run), cudaMemcpyDeviceToHost, c->stream));
      KGX_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    const std::vector<Run>& runs = c->plan.runs;
    size_t lo = 0, hi = runs.size();
    while (hi - lo > 1) { size_t mid = (lo + hi) / 2; if (runs[mid].slot_start <= slot) lo = mid; else hi = mid; }
    const Run& r = runs[lo];
    const uint32_t a = r.a_first + (slot - r.slot_start);
    std::string s;
    auto piece = [&](const DBuf& off, const DBuf& blob, uint64_t i) -> int {
      unsigned long long o[2];
      KGX_CUDA(c, cudaMemcpyAsync(o, ptr<unsigned long long>(off) + i, 16, cudaMemcpyDeviceToHost, c->stream));
      KGX_CUDA(c, cudaStreamSynchronize(c->stream));
      const size_t at = s.size();
      s.resize(at + size_t(o[1] - o[0]));
      if (o[1] > o[0]) {
        KGX_CUDA(c, cudaMemcpyAsync(&s[at], ptr<uint8_t>(blob) + o[0], size_t(o[1] - o[0]), cudaMemcpyDeviceToHost, c->stream));
        KGX_CUDA(c, cudaStreamSynchronize(c->stream));
      }
      return KGX_OK;
    };
    KGX_TRY(piece(c->d_A_off, c->d_A_blob, a));
    if (r.gidx != 0xFFFFFFFFu) KGX_TRY(piece(c->d_G_off, c->d_G_blob, r.gidx));
    KGX_TRY(piece(c->d_B_off, c->d_B_blob, a));
    *len = s.size();
    if (out && cap) memcpy(out, s.data(), std::min(cap, s.size()));
    return KGX_OK;
  }
  const Plan& p = c->plan;
  size_t lo = 0, hi = p.runs.size();
  while (hi - lo > 1) { size_t mid = (lo + hi) / 2; if (p.runs[mid].slot_start <= slot) lo = mid; else hi = mid; }
  const Run& r = p.runs[lo];
  const uint32_t a = r.a_first + (slot - r.slot_start);
  std::string s(p.A_blob, size_t(p.A_off[a]), size_t(p.A_off[a + 1] - p.A_off[a]));
  if (r.gidx != 0xFFFFFFFFu) s.append(p.G_blob, size_t(p.G_off[r.gidx]), size_t(p.G_off[r.gidx + 1] - p.G_off[r.gidx]));
  s.append(p.B_blob, size_t(p.B_off[a]), size_t(p.B_off[a + 1] - p.B_off[a]));
  *len = s.size();
  if (out && cap) memcpy(out, s.data(), std::min(cap, s.size()));
  return KGX_OK;
}

int kgx_format_values(kgx_ctx* c, const double* values, size_t n, uint8_t* out32) {
  if (!c || (n && (!values || !out32))) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (!n) return KGX_OK;
  KGX_TRY(dev_reserve(c, c->d_tmp[1], n * 8));
  KGX_TRY(dev_reserve(c, c->d_tmp[2], n * 32));
  KGX_CUDA(c, cudaMemcpyAsync(c->d_tmp[1].p, values, n * 8, cudaMemcpyHostToDevice, c->stream));
  k_format_values<<<cdiv(n, 256), 256, 0, c->stream>>>(ptr<double>(c->d_tmp[1]), n, ptr<uint8_t>(c->d_tmp[2]));
  c->stats.kernel_launches += 1;
  KGX_TRY(launch_check(c, "k_format_values"));
  KGX_CUDA(c, cudaMemcpyAsync(out32, c->d_tmp[2].p, n * 32, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  return KGX_OK;
}

void* kgx_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) return nullptr;
  return p;
}
void kgx_host_free(void* p) { if (p) cudaFreeHost(p); }

int kgx_get_stats(kgx_ctx* c, kgx_stats* out) {
  if (!c || !out) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  KGX_TRY(settle_ingest(c));
  c->stats.n_series = c->n_series;
  c->stats.device_bytes = c->device_bytes;
  *out = c->stats;
  return KGX_OK;
}

int kgx_sync(kgx_ctx* c) {
  if (!c) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  KGX_TRY(settle_ingest(c));
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  return KGX_OK;
}

int kgx_get_profile(kgx_ctx* c, kgx_profile* out) {
  if (!c || !out) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  for (auto& r : c->regions) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) c->prof.ms[r.kind] += ms;
    cudaEventDestroy(r.a); cudaEventDestroy(r.b);
  }
  c->regions.clear();
  *out = c->prof;
  return KGX_OK;
}

int kgx_set_profile(kgx_ctx* c, int on) {
  if (!c) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  c->profile = on != 0;
  return KGX_OK;
}

int kgx_profile_reset(kgx_ctx* c) {
  if (!c) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  for (auto& r : c->regions) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  c->regions.clear();
  c->prof = kgx_profile{};
  return KGX_OK;
}

#ifdef KGX_PHASE_CLOCKS
// experiment builds only (tools/phase_clocks.py): cycles thread 0 of every writer CTA spent in each phase of a tile
int kgx_debug_phase_clocks(unsigned long long* out16, int reset) {
  if (out16 && cudaMemcpyFromSymbol(out16, kgx::g_phase_clk, 16 * 8) != cudaSuccess) return KGX_E_CUDA;
  if (reset) { unsigned long long z[16] = {}; if (cudaMemcpyToSymbol(kgx::g_phase_clk, z, sizeof z) != cudaSuccess) return KGX_E_CUDA; }
  return KGX_OK;
}
#endif

}  // extern "C"

// =================== render without any allocation ==========================
// kgx_render_noalloc: the exposition writer for callers that must not touch the
// CUDA allocator at scrape time.  Two kinds of embedder are in that position:
// a /metrics handler that shares its GPU with a training job which has already
// claimed all device memory (a cudaMalloc there fails, or worse, succeeds and
// starves the job), and a handler with a latency budget that cannot absorb the
// device-wide synchronisation cudaMalloc/cudaFree imply.  For them the writer's
// scratch — the look-back word of every tile, the ticket and length counters,
// and the per-node body offsets — lives in one fixed array that is part of the
// library image, and the text goes straight into a device buffer the caller
// owns.  The array costs KGX_NOALLOC_SCRATCH_BYTES on every GPU that loads the
// library, until the process exits.
namespace {

constexpr uint32_t kNoallocMaxTiles = KGX_NOALLOC_MAX_TILES;
constexpr uint32_t kNoallocMaxNodes = KGX_NOALLOC_MAX_NODES;
constexpr size_t kNoallocWords = size_t(C_COUNT) + kNoallocMaxTiles + kNoallocMaxNodes + 1;
static_assert(kNoallocWords * 8 == KGX_NOALLOC_SCRATCH_BYTES, "header and scratch array disagree");

// [0, C_COUNT)            counters (ticket, total length)
// [C_COUNT, +max tiles)   look-back words
// [.., +max nodes + 1)    body offset of every node
static __device__ __align__(256) unsigned long long g_noalloc_scratch[kNoallocWords];

constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
unsigned long long* g_noalloc_ptr[kMaxDevices] = {};  // device address of the array, per GPU

// The array's name is only a host-side handle here; what a kernel needs is the
// device address, which differs per GPU.  Looked up once per device, with that
// device current.
int noalloc_scratch(kgx_ctx* c, unsigned long long** out) {
  if (c->device < 0 || c->device >= kMaxDevices) return fail(c, KGX_E_LIMIT, "device index beyond 63");
  std::lock_guard<std::mutex> g(g_noalloc_mu);
  if (!g_noalloc_ptr[c->device]) {
    void* p = nullptr;
    KGX_CUDA(c, cudaGetSymbolAddress(&p, g_noalloc_scratch));
    g_noalloc_ptr[c->device] = static_cast<unsigned long long*>(p);
  }
  *out = g_noalloc_ptr[c->device];
  return KGX_OK;
}

}  // namespace

extern "C" int kgx_render_noalloc(kgx_ctx* c, uint8_t* d_out, size_t cap, size_t* len) {
  if (!c || !len) return KGX_E_INVAL;
  std::lock_guard<std::mutex> g(c->mu);
  cudaSetDevice(c->device);
  if (!c->have_table) return fail(c, KGX_E_STATE, "kgx_render_noalloc before kgx_set_table");
  KGX_TRY(settle_ingest(c));  // only a kgx_ingest_device can still be in flight here (kgx.h: kgx_sync it first)
  KGX_TRY(flush_reset(c));
  if (c->n_tiles == 0) { *len = 0; return KGX_OK; }
  if (c->n_tiles > kNoallocMaxTiles)
    return fail(c, KGX_E_LIMIT, "kgx_render_noalloc: more render tiles than KGX_NOALLOC_MAX_TILES; use kgx_render");
  if (c->plan.n_nodes > kNoallocMaxNodes)
    return fail(c, KGX_E_LIMIT, "kgx_render_noalloc: more nodes than KGX_NOALLOC_MAX_NODES; use kgx_render");
  // the writer does not bounds-check: the buffer must hold the worst case
  uint64_t bound = 64;
  for (const FamMeta& m : c->fam)
    bound += uint64_t(c->n_series) * (m.name_len + 2 + kMaxValueText) + c->lbl_bytes + uint64_t(c->plan.n_nodes) * m.header_len;
  if (!d_out || cap < bound) { *len = size_t(bound); return fail(c, KGX_E_NOSPACE, "kgx_render_noalloc: output buffer below the worst-case size (returned in *len)"); }
  if (reinterpret_cast<uintptr_t>(d_out) & 15) return fail(c, KGX_E_INVAL, "kgx_render_noalloc: output buffer must be 16-byte aligned");
  unsigned long long* scratch = nullptr;
  KGX_TRY(noalloc_scratch(c, &scratch));
  // the array still holds the previous call's words: everything the kernel
  // expects to start from zero (or from "unset") is cleared here
  KGX_CUDA(c, cudaMemsetAsync(scratch, 0, (size_t(C_COUNT) + c->n_tiles) * 8, c->stream));
  KGX_CUDA(c, cudaMemsetAsync(scratch + C_COUNT + kNoallocMaxTiles, 0xff, (size_t(c->plan.n_nodes) + 1) * 8, c->stream));
  DevRender R{};
  R.fam = ptr<FamMeta>(c->d_fam); R.n_families = uint32_t(c->fam.size());
  R.tiles = ptr<Tile>(c->d_tiles); R.n_tiles = c->n_tiles;
  R.lbl_off = ptr<unsigned long long>(c->d_lbl_off); R.lbl_blob = ptr<uint8_t>(c->d_lbl_blob);
  R.ser_len = ptr<uint32_t>(c->d_ser_len);
  R.ser_gidx = ptr<uint32_t>(c->d_ser_gidx); R.gpu_total = ptr<unsigned long long>(c->d_gpu_total);
  R.tile_state = scratch + C_COUNT;
  R.node_body_off = scratch + C_COUNT + kNoallocMaxTiles;
  R.out = d_out; R.smem_bytes = c->render_smem;
  R.epoch = 0; R.check_churn = 0; R.host_len = nullptr;  // the array was cleared above: untagged words, no hand-over
  const int grid = int(std::min<uint32_t>(c->n_tiles, uint32_t(c->render_grid)));
  if (c->render_v2) launch_writer(c, grid, R, scratch);
  else k_render<<<grid, kRenderThreads, c->render_smem, c->stream>>>(R, acc_view(c), scratch);
  c->stats.kernel_launches += 1;
  KGX_TRY(launch_check(c, "k_render (noalloc)"));
  k_post_words<<<1, 32, 0, c->stream>>>(scratch + C_TOTAL_LEN, c->h_counters_dev + C_TICKET + 1, 1);
  c->stats.kernel_launches += 1;
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  *len = size_t(c->h_counters[C_TICKET + 1]);
  return KGX_OK;
}
