// kgx_writer.cuh — the exposition writer (sm_100a): series state + label text in,
// Prometheus text exposition out, byte-identical to what the reference serves
// through promhttp (main.go:69-71 -> Registry.Gather + expfmt, rules O1-O3, G1-G2,
// F1-F5 of SURVEY.md 8c).  One pass, fused per tile of <= THREADS series of one
// family of one node (exact lengths -> scan -> scatter, the north star's three
// stages, without an intermediate):
//
//   * the label text of a tile is ONE contiguous byte range of the series label
//     blob: a single TMA bulk copy (cp.async.bulk global->shared, completion on
//     an mbarrier) stages it.  Warp 0 issues it for the NEXT tile — together
//     with that tile's ticket and descriptor — before the CTA starts on it, so
//     the ticket, the descriptor and the labels are not waited for;
//   * each thread owns one line and formats its value once, straight into a
//     shared-memory slot (" <value>\n");
//   * line lengths -> CTA scan; the tile's size is published at once, and every
//     warp resolves the tile's global byte offset itself by decoupled look-back
//     over 32 predecessors per step.  Tickets are handed out in output order,
//     so a tile never waits on one that has not started;
//   * each thread then streams its line to the output as 16-byte chunks aligned
//     to the GLOBAL address: a chunk belongs to the line that holds its first
//     byte, so no two lanes ever write the same chunk and every store is a full
//     streaming STG.128.  The four sources of a line (metric name, staged
//     labels, value slot, head of the next line's name) are appended one after
//     the other, all lanes of a warp in the same source at the same time; word
//     loads + funnel shifts realign, computed byte masks merge the one chunk
//     that straddles two sources.  Only the head of a tile's first line and the
//     tail of its last one are partial chunks: those few bytes are byte stores.
//
// Needs metric names of 15..64 bytes (the tail of a chunk is completed with
// the next line's name); other schemas take the generic k_render.
#pragma once
#include "kgx_async.cuh"
#include "kgx_kernels.cuh"

namespace kgx {

struct RenderLayout {  // byte offsets inside dynamic shared memory (all 16-aligned)
  uint32_t names;      // kMaxFamilies x 96: [16 pad][name (<=64)][16 pad]
  uint32_t slots;      // THREADS x 28: " <value>\n"
  uint32_t stage;      // staged label text (+32 slack before, +128+48 after)
  uint32_t total;
};
constexpr uint32_t kSlotBytes = 28;  // 7 words: odd word stride -> conflict-free byte stores
constexpr uint32_t kNameStride = 96;
#ifndef KGX_LOOKBACK
#define KGX_LOOKBACK 1
#endif
constexpr int kLookBack = KGX_LOOKBACK;  // predecessors per lane and look-back step

__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void stg128_cs(uint8_t* p, uint4 v) {
  asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// bytes [0,e) from lo, bytes [e,16) from hi (0 < e < 16).  The masks are computed:
// the writer is bound by the L1/LSU data pipe, not by the ALU.
__device__ __forceinline__ uint4 merge16(uint4 lo, uint4 hi, uint32_t e) {
  auto mk = [e](int w) -> uint32_t {  // the bytes of word w that are < e
    const int n = min(max(int(e) - 4 * w, 0), 4);
    return __funnelshift_rc(0xffffffffu, 0u, uint32_t(32 - 8 * n));
  };
  const uint4 m = make_uint4(mk(0), mk(1), mk(2), mk(3));
  return make_uint4((lo.x & m.x) | (hi.x & ~m.x), (lo.y & m.y) | (hi.y & ~m.y), (lo.z & m.z) | (hi.z & ~m.z),
                    (lo.w & m.w) | (hi.w & ~m.w));
}

#ifdef KGX_PHASE_CLOCKS  // experiment build only (tools/phase_clocks.py): where a tile's time goes, thread 0's view
__device__ unsigned long long g_phase_clk[16];
#define KGX_PC_DECL unsigned long long pc_acc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0}; long long pc_t = clock64(); unsigned long long pc_tiles = 0
#define KGX_PC(i) do { if (tid == 0) { const long long n_ = clock64(); pc_acc[i] += (unsigned long long)(n_ - pc_t); pc_t = n_; } } while (0)
#define KGX_PC_FLUSH do { if (tid == 0) { for (int i_ = 0; i_ < 10; ++i_) atomicAdd(&g_phase_clk[i_], pc_acc[i_]); atomicAdd(&g_phase_clk[10], pc_tiles); } } while (0)
#else
#define KGX_PC_DECL
#define KGX_PC(i)
#define KGX_PC_FLUSH
#endif

template <int THREADS>
__global__ void __launch_bounds__(THREADS) k_render3(DevRender R, RenderLayout lay, DevAcc acc, unsigned long long* counters) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ unsigned long long s_scan[THREADS / 32];
  __shared__ uint4 s_tail;
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ uint32_t s_next;  // ticket of the tile to process next
  __shared__ Tile s_desc;      // its descriptor
  __shared__ uint32_t s_l0;    // where its first label starts inside the staged range
  const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  // the join that ran just before saw series the index does not hold yet: this body would miss them.  Every CTA leaves;
  // the host rebuilds the index, folds what was skipped and launches again (churn only).
  unsigned long long* ticket = ticket_of(counters, R.epoch);
  if (R.check_churn && *(volatile unsigned long long*)&counters[C_CHURN] != 0) return;
  const uint32_t sm0 = uint32_t(__cvta_generic_to_shared(smem));
  const uint32_t bar = uint32_t(__cvta_generic_to_shared(&s_bar));

  // one-time: metric names of every family, with slack on both sides
  for (uint32_t i = tid; i < kMaxFamilies * kNameStride; i += THREADS) {
    const uint32_t f = i / kNameStride, o = i % kNameStride;
    smem[lay.names + i] = (o >= 16 && o < 16 + 64 && f < R.n_families) ? R.fam[f].name[o - 16] : uint8_t('#');
  }
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // The ticket counter is one address for the whole grid and its reply takes ~0.5 us under load.  Thread 0 draws
  // the next ticket as soon as its own line is out, just before the end-of-tile barrier, so part of that latency
  // passes while the rest of the CTA finishes.  (Drawing it any earlier — a tile ahead — lets tiles start out of
  // ticket order and costs 3.7x; fetching the descriptor before the barrier as well measured slower, 1.601 vs
  // 1.586 ms: profiles/README.md.)
  unsigned long long tk_next = 0;
  if (tid == 0) tk_next = atomicAdd(ticket, 1ull);
  auto fetch_next = [&]() {  // warp 0: the drawn ticket, its descriptor, and the bulk copy of its label text
    const uint32_t nt = __shfl_sync(0xffffffffu, uint32_t(tk_next), 0);
    if (nt < R.n_tiles && lane == 0) {
      const Tile d = R.tiles[nt];
      const unsigned long long L0 = d.lbl0, L1 = d.lbl1;
      const unsigned long long g0 = L0 & ~127ull;  // whole 128-byte lines
      const uint32_t bytes = uint32_t(((L1 + 15ull) & ~15ull) - g0);
      if (bytes) {
        mbar_expect_tx(bar, bytes);
        bulk_g2s(sm0 + lay.stage + 32, R.lbl_blob + g0, bytes, bar);
      } else {
        mbar_arrive(bar);
      }
      s_desc = d;
      s_l0 = uint32_t(L0 - g0);
    }
    if (lane == 0) s_next = nt;
  };
  if (wid == 0) fetch_next();
  __syncthreads();

  uint32_t parity = 0;
  KGX_PC_DECL;
  for (;;) {
    const uint32_t t = s_next;
    if (t >= R.n_tiles) {
      KGX_PC_FLUSH;
      return;
    }
    const Tile tile = s_desc;
    const uint32_t l0 = s_l0;
    const FamMeta& fm = R.fam[tile.family];
    const uint32_t name_len = fm.name_len;
    // ---- pass 1: value text and exact length of my line
    const uint32_t slot = sm0 + lay.slots + tid * kSlotBytes;
    uint32_t len = 0, lbl_len = 0, vlen = 0;
    bool live = false;
    if (tid < tile.count) {
      const uint32_t r = tile.first + tid;
      lbl_len = R.ser_len[r];  // 4 bytes per series; offsets inside the tile come from the CTA scan below
      double v = series_value(fm.kind, acc, R, r, live);
#ifdef KGX_PHASE_CLOCKS
      asm volatile("mov.b64 %0, %0;" : "+d"(v));
      KGX_PC(0);  // state loads
#endif
      if (live) {
        uint8_t* sp = smem + lay.slots + tid * kSlotBytes;
        sp[0] = ' ';
        vlen = uint32_t(format_value(v, sp + 1));
        sp[1 + vlen] = '\n';
        len = name_len + lbl_len + vlen + 2;
      }
    }
    KGX_PC(1);  // format
    const uint32_t hdr = ((tile.flags & 1) && acc.node_live[tile.node] != 0) ? fm.header_len : 0;
    // ---- exclusive scan over the CTA, three fields at once: output bytes (24 bits), label bytes of
    // every series, printed or not (24 bits), lines printed (16 bits)
    const unsigned long long mine = (uint64_t(live) << 48) | (uint64_t(lbl_len) << 24) | len;
    unsigned long long inc = mine;
    for (int o = 1; o < 32; o <<= 1) { unsigned long long u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
    if (lane == 31) s_scan[wid] = inc;
    __syncthreads();  // (1) also: everyone has read s_next / s_desc / s_l0
    KGX_PC(2);  // scan barrier
    unsigned long long woff = 0, all = 0;
    for (int k = 0; k < THREADS / 32; ++k) { if (k < int(wid)) woff += s_scan[k]; all += s_scan[k]; }
    const unsigned long long exc = woff + inc - mine;
    const uint32_t off = hdr + (uint32_t(exc) & 0xffffffu);
    const uint32_t total = hdr + (uint32_t(all) & 0xffffffu);
    const uint32_t lbl_src = sm0 + lay.stage + 32 + l0 + (uint32_t(exc >> 24) & 0xffffffu);  // my label in the staged range
    const bool first_live = live && uint32_t(exc >> 48) == 0;
    const bool last_live = live && uint32_t(exc >> 48) + 1 == uint32_t(all >> 48);
    // ---- global byte offset of the tile: decoupled look-back, 32 predecessors per step.
    // Every warp resolves it for itself (same loads, same answer): no barrier, no broadcast.
    if (tid == 0 && t > 0) atomicExch(&R.tile_state[t], kStAgg | st_tag(R.epoch) | total);
    unsigned long long base = 0;
    if (t > 0) {
      // One step of the walk covers 32 * kLookBack predecessors: kLookBack coalesced loads of 32 consecutive look-back
      // words each (lane 0 of load 0 is the direct predecessor), all in flight together.  With thousands of tiles in
      // flight the nearest tile that already knows its prefix is hundreds of tickets back, so a 32-tile step meant
      // ~20 dependent L2 round trips per tile (profiles/README.md, round 2).
      long long p = (long long)t - 1 - lane;
      for (;;) {
        unsigned long long st[kLookBack];
#pragma unroll
        for (int k = 0; k < kLookBack; ++k) {
          const long long q = p - 32 * k;
          st[k] = q >= 0 ? (unsigned long long)*(volatile unsigned long long*)&R.tile_state[q] : (kStPrefix | st_tag(R.epoch));  // before tile 0: prefix 0
        }
        unsigned long long v = 0;
        bool found = false;
#pragma unroll
        for (int k = 0; k < kLookBack; ++k) {
          const long long q = p - 32 * k;
          while (!st_published(st[k], R.epoch)) st[k] = *(volatile unsigned long long*)&R.tile_state[q];  // not published yet (or an older launch's word)
          const uint32_t pfx = __ballot_sync(0xffffffffu, (st[k] & kStMask) == kStPrefix);
          if (!found) {  // warp-uniform
            const uint32_t first = pfx ? uint32_t(__ffs(pfx) - 1) : 32u;
            if (lane <= first) v += st[k] & kStValue;
            found = pfx != 0;
          }
        }
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        base += v;
        if (found) break;
        p -= 32 * kLookBack;
      }
    }
    KGX_PC(3);  // look-back
    if (tid == 0) {
      atomicExch(&R.tile_state[t], kStPrefix | st_tag(R.epoch) | (base + total));
      if (tile.flags & 2) R.node_body_off[tile.node] = st_tag(R.epoch) | base;
      if (t == R.n_tiles - 1) { counters[C_TOTAL_LEN] = base + total; if (R.host_len) *R.host_len = base + total; }
    }
    uint8_t* gtile = R.out + base;            // first byte of the tile in the output
    uint8_t* galn = R.out + (base & ~15ull);  // 16-aligned origin: position v lives at galn + v
    const uint32_t sh = uint32_t(base & 15);
    if (hdr) for (uint32_t i = tid; i < hdr; i += THREADS) gtile[i] = fm.header[i];
    mbar_wait(bar, parity);  // staged labels have landed (issued a whole tile ago)
    parity ^= 1;
    KGX_PC(4);  // header + label wait
    // ---- pass 2: stream my line to the output as 16-byte chunks aligned to the global address
    if (live) {
      const uint32_t g = sh + off;                                              // first byte of my line (position)
      const uint32_t b0 = g + name_len, b1 = b0 + lbl_len, b2 = b1 + vlen + 2;  // source boundaries; b2 = line end
      const uint32_t start = (g + 15u) & ~15u;
      const uint32_t stop = last_live ? (b2 & ~15u) : ((b2 + 15u) & ~15u);      // my full chunks: [start, stop)
      const uint32_t name_sm = sm0 + lay.names + tile.family * kNameStride + 16;
      const uint32_t d0 = name_sm - g, d1 = lbl_src - b0, d2 = slot - b1, d3 = name_sm - b2;  // source = position + delta
      if (first_live) {  // the bytes before my first aligned chunk: nobody else writes them
        for (uint32_t q = g; q < start && q < b2; ++q) galn[q] = smem[lay.names + tile.family * kNameStride + 16 + (q - g)];
      }
      // `a4` holds the chunk at p, of which the first `e` bytes are already final.
      uint32_t p = start, e = 0;
      uint4 a4 = make_uint4(0, 0, 0, 0);
      auto append = [&](uint32_t d, uint32_t b) {  // source = position + d, up to position b (exclusive)
        // the source is walked in 16-byte steps: four new words per step, the fifth carried over
        uint32_t wa = (p + d) & ~3u;
        const uint32_t r = ((p + d) & 3u) * 8u;
        uint32_t w0 = lds32(wa), w1 = lds32(wa + 4), w2 = lds32(wa + 8), w3 = lds32(wa + 12), w4 = lds32(wa + 16);
        auto chunk = [&]() {
          return make_uint4(__funnelshift_r(w0, w1, r), __funnelshift_r(w1, w2, r), __funnelshift_r(w2, w3, r),
                            __funnelshift_r(w3, w4, r));
        };
        auto step = [&]() { wa += 16; w0 = w4; w1 = lds32(wa + 4); w2 = lds32(wa + 8); w3 = lds32(wa + 12); w4 = lds32(wa + 16); };
        const uint4 nw = chunk();
        a4 = e ? merge16(a4, nw, e) : nw;
        if (b >= p + 16) {
          stg128_cs(galn + p, a4);
          p += 16;
          step();
          while (p + 16 <= b) {
            stg128_cs(galn + p, chunk());
            p += 16;
            step();
          }
          a4 = chunk();
        }
        e = b - p;
      };
      append(d0, b0);  // metric name
      append(d1, b1);  // labels
      append(d2, b2);  // " <value>\n"
      if (!last_live) {
        append(d3, stop);  // the head of the next line's name completes my last chunk
      } else if (e) {      // the tile ends inside a chunk: its bytes belong to me, the rest to the next tile
        s_tail = a4;       // one thread per tile gets here: park the chunk, then byte-copy
        const uint8_t* sp = reinterpret_cast<const uint8_t*>(&s_tail);
        for (uint32_t q = 0; q < e; ++q) galn[p + q] = sp[q];
      }
    }
    KGX_PC(5);  // stream my line
    if (tid == 0) tk_next = atomicAdd(ticket, 1ull);  // in flight while the rest of the CTA finishes its lines
    __syncthreads();             // (2) label stage and value slots are free
    KGX_PC(6);  // end-of-tile barrier
    if (wid == 0) fetch_next();  // next ticket, descriptor, label copy
    KGX_PC(7);  // ticket + descriptor
    __syncthreads();             // (3) s_next / s_desc / s_l0 visible
    KGX_PC(8);
#ifdef KGX_PHASE_CLOCKS
    ++pc_tiles;
#endif
  }
}

}  // namespace kgx
