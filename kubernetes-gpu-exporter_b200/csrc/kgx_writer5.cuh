// kgx_writer5.cuh — the exposition writer, software-pipelined over tiles (sm_100a).
//
// The building blocks are described in kgx_writer.cuh (TMA-staged label text, one line per thread, 16-byte chunks
// aligned to the global address, decoupled look-back over tiles handed out in output order).  What this kernel adds is
// WHEN a tile's size is published.  A writer that formats a tile's values, publishes its size and at once needs the
// sizes of its 32 predecessors finds other CTAs formatting those at that very moment: it waits for the slowest of them,
// 41-45 % of a tile's time (round 1's k_render3; profiles/README.md, round 2: phase clocks).  Here a CTA keeps two tiles
// in flight:
//
//   iteration i:   MEASURE tile t(i+1)   state loads, value text into the other slot buffer, CTA scan,
//                                        size published                     <- a whole tile period early
//                  EMIT    tile t(i)     look-back (the predecessors' sizes have been out for a period), ticket of
//                                        t(i+2) drawn, header, lines streamed; thread 0 then loads t(i+2)'s descriptor
//                  one CTA barrier       then labels of t(i+1) -> stage (TMA)
//
// Tickets are still drawn in output order and sizes are still published in ticket order, so the look-back never waits
// on a tile that has not started; it just no longer waits on one that is still formatting.
#pragma once
#include "kgx_writer.cuh"

namespace kgx {

#ifndef KGX_LOOKBACK5
#define KGX_LOOKBACK5 1
#endif
constexpr int kLookBack5 = KGX_LOOKBACK5;  // look-back words per lane and step (32 * kLookBack5 predecessors per step)

struct LineState {  // what MEASURE leaves for EMIT, per thread
  uint32_t len, lbl_len, vlen;  // bytes of my line, of its label text, of its value text
  uint32_t vstart;              // where " <value>\n" starts inside my value slot (format_slot)
  uint32_t off;                 // my line's offset inside the tile (header included)
  uint32_t lbl_at;              // my label's offset inside the tile's label range
  uint32_t total, hdr;          // tile bytes, header bytes (CTA-uniform)
  bool live, first_live, last_live;
};

// Shared memory allows 10 CTAs of 128 lines per SM, but pinning the register budget to that (48 per thread) costs a
// spill and measured slower than the 56 registers the kernel wants at 9 CTAs: 1.420 vs 1.382 ms (64 registers at 8 CTAs:
// 1.423 ms).  So: 1 152 resident threads.
template <int THREADS>
__global__ void __launch_bounds__(THREADS, (1152 / THREADS < 32 ? 1152 / THREADS : 32)) k_render5(DevRender R, RenderLayout lay, DevAcc acc, unsigned long long* counters) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ unsigned long long s_scan[THREADS / 32];
  __shared__ uint4 s_tail;
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ uint32_t s_tk[3];   // tickets of the tile being emitted, the one being measured and the one after
  __shared__ Tile s_desc[3];     // their descriptors
  const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  unsigned long long* ticket = ticket_of(counters, R.epoch);
  if (R.check_churn && *(volatile unsigned long long*)&counters[C_CHURN] != 0) return;
  const uint32_t sm0 = uint32_t(__cvta_generic_to_shared(smem));
  const uint32_t bar = uint32_t(__cvta_generic_to_shared(&s_bar));

  for (uint32_t i = tid; i < kMaxFamilies * kNameStride; i += THREADS) {
    const uint32_t f = i / kNameStride, o = i % kNameStride;
    smem[lay.names + i] = (o >= 16 && o < 16 + 64 && f < R.n_families) ? R.fam[f].name[o - 16] : uint8_t('#');
  }
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  unsigned long long tk_next = 0;
  // thread 0: the ticket it drew goes into descriptor slot k, with its descriptor
  auto fetch_desc = [&](uint32_t k) {
    const uint32_t nt = uint32_t(tk_next);
    if (nt < R.n_tiles) s_desc[k] = R.tiles[nt];
    s_tk[k] = nt;
  };
  // warp 0: bulk copy of the label text of the tile in slot k into the (single) stage
  auto issue_labels = [&](uint32_t k) {
    if (lane == 0) {
      const unsigned long long L0 = s_desc[k].lbl0, L1 = s_desc[k].lbl1;
      const unsigned long long g0 = L0 & ~127ull;  // whole 128-byte lines
      const uint32_t bytes = uint32_t(((L1 + 15ull) & ~15ull) - g0);
      if (bytes) {
        mbar_expect_tx(bar, bytes);
        bulk_g2s(sm0 + lay.stage + 32, R.lbl_blob + g0, bytes, bar);
      } else {
        mbar_arrive(bar);
      }
    }
  };

  // ---- MEASURE the tile in pipeline slot k: value text into slot buffer k, exact line lengths, CTA scan, and the
  // tile's size published for the look-back of its successors.  Contains one CTA barrier.
  auto measure = [&](uint32_t k, uint32_t sb, LineState& L) {
    const Tile tile = s_desc[k];
    const uint32_t t = s_tk[k];
    const FamMeta& fm = R.fam[tile.family];
    L.len = 0; L.lbl_len = 0; L.vlen = 0; L.vstart = 0; L.live = false;
    if (tid < tile.count) {
      const uint32_t r = tile.first + tid;
      L.lbl_len = R.ser_len[r];
      bool live;
      const double v = series_value(fm.kind, acc, R, r, live);
      L.live = live;
      if (live) {
        uint32_t vs;
        L.vlen = format_slot(v, smem + lay.slots + (sb * THREADS + tid) * kSlotBytes, vs);
        L.vstart = vs - 1;  // the space before the value
        L.len = fm.name_len + L.lbl_len + L.vlen + 2;
      }
    }
    L.hdr = ((tile.flags & 1) && acc.node_live[tile.node] != 0) ? fm.header_len : 0;
    // exclusive scan over the CTA, three fields at once: output bytes (24 bits), label bytes of every series,
    // printed or not (24 bits), lines printed (16 bits)
    const unsigned long long mine = (uint64_t(L.live) << 48) | (uint64_t(L.lbl_len) << 24) | L.len;
    unsigned long long inc = mine;
    for (int o = 1; o < 32; o <<= 1) { unsigned long long u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
    if (lane == 31) s_scan[wid] = inc;
    __syncthreads();
    unsigned long long woff = 0, all = 0;
    for (int w = 0; w < THREADS / 32; ++w) { if (w < int(wid)) woff += s_scan[w]; all += s_scan[w]; }
    const unsigned long long exc = woff + inc - mine;
    L.off = L.hdr + (uint32_t(exc) & 0xffffffu);
    L.total = L.hdr + (uint32_t(all) & 0xffffffu);
    L.lbl_at = uint32_t(exc >> 24) & 0xffffffu;
    L.first_live = L.live && uint32_t(exc >> 48) == 0;
    L.last_live = L.live && uint32_t(exc >> 48) + 1 == uint32_t(all >> 48);
    if (tid == 0 && t > 0) atomicExch(&R.tile_state[t], kStAgg | st_tag(R.epoch) | L.total);
  };

  uint32_t parity = 0;
  KGX_PC_DECL;
  // ---- EMIT the tile in pipeline slot k (measured one iteration ago): global offset by look-back, header, lines
  auto emit = [&](uint32_t k, uint32_t sb, const LineState& L, bool more) {
    const Tile tile = s_desc[k];
    const uint32_t t = s_tk[k];
    const FamMeta& fm = R.fam[tile.family];
    const uint32_t name_len = fm.name_len;
    const uint32_t slot = sm0 + lay.slots + (sb * THREADS + tid) * kSlotBytes;
    // decoupled look-back, 32 predecessors per step; every warp resolves it for itself (no barrier, no broadcast)
    unsigned long long base = 0;
    if (t > 0) {
      // One step covers 32 * kLookBack5 predecessors: kLookBack5 coalesced loads of 32 consecutive look-back words each
      // (lane 0 of load 0 is the direct predecessor), all in flight together.  The predecessors' sizes were published a
      // tile period ago, so the walk rarely spins; steps wider than 32 words measured slower (1.51 / 1.53 / 1.67 ms at
      // 1 / 4 / 8 words per lane), so kLookBack5 is 1.
      long long p = (long long)t - 1 - lane;
      for (;;) {
        unsigned long long st[kLookBack5];
#pragma unroll
        for (int q = 0; q < kLookBack5; ++q) {
          const long long at = p - 32 * q;
          st[q] = at >= 0 ? (unsigned long long)*(volatile unsigned long long*)&R.tile_state[at] : (kStPrefix | st_tag(R.epoch));  // before tile 0: prefix 0
        }
        unsigned long long v = 0;
        bool found = false;
#pragma unroll
        for (int q = 0; q < kLookBack5; ++q) {
          const long long at = p - 32 * q;
          while (!st_published(st[q], R.epoch)) { st[q] = *(volatile unsigned long long*)&R.tile_state[at]; KGX_PC_COUNT(9); }
          const uint32_t pfx = __ballot_sync(0xffffffffu, (st[q] & kStMask) == kStPrefix);
          if (!found) {  // warp-uniform
            const uint32_t first = pfx ? uint32_t(__ffs(pfx) - 1) : 32u;
            if (lane <= first) v += st[q] & kStValue;
            found = pfx != 0;
          }
        }
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        base += v;
        KGX_PC_COUNT(8);
        if (found) break;
        p -= 32 * kLookBack5;
      }
    }
    if (tid == 0) {
      atomicExch(&R.tile_state[t], kStPrefix | st_tag(R.epoch) | (base + L.total));
      if (tile.flags & 2) R.node_body_off[tile.node] = st_tag(R.epoch) | base;
      if (t == R.n_tiles - 1) { counters[C_TOTAL_LEN] = base + L.total; if (R.host_len) *R.host_len = base + L.total; }
    }
    uint8_t* gtile = R.out + base;            // first byte of the tile in the output
    // 32-byte aligned origin: position v lives at galn + v.  Chunks are 16 bytes; two that fill a 32-byte sector
    // leave in one 256-bit store.
    const unsigned long long gabs = reinterpret_cast<unsigned long long>(R.out) + base;
    uint8_t* galn = reinterpret_cast<uint8_t*>(gabs & ~31ull);
    const uint32_t sh = uint32_t(gabs & 31);
    if (L.hdr) for (uint32_t i = tid; i < L.hdr; i += THREADS) gtile[i] = fm.header[i];
    // The ticket counter's reply takes ~0.5 us under load.  Drawn here — after the look-back, before the lines — it is
    // back by the time thread 0's line is out, and thread 0 fetches the descriptor before the end-of-tile barrier
    // (1.381 -> 1.351 ms).  Drawn earlier (before the look-back: 1.374; a whole iteration ahead: 1.59) tiles drift out of
    // ticket order and every look-back gets longer.
    if (tid == 0 && more) tk_next = atomicAdd(ticket, 1ull);
    KGX_PC(3);  // look-back
    mbar_wait(bar, parity);  // the tile's label text has landed (issued right after the previous end-of-tile barrier)
    parity ^= 1;
    KGX_PC(4);  // header + label wait
    if (L.live) {
      const uint32_t lbl_src = sm0 + lay.stage + 32 + uint32_t(tile.lbl0 & 127ull) + L.lbl_at;
      const uint32_t g = sh + L.off;                                                  // first byte of my line (position)
      const uint32_t b0 = g + name_len, b1 = b0 + L.lbl_len, b2 = b1 + L.vlen + 2;    // source boundaries; b2 = line end
      const uint32_t start = (g + 15u) & ~15u;
      const uint32_t stop = L.last_live ? (b2 & ~15u) : ((b2 + 15u) & ~15u);          // my full chunks: [start, stop)
      const uint32_t name_sm = sm0 + lay.names + tile.family * kNameStride + 16;
      const uint32_t d0 = name_sm - g, d1 = lbl_src - b0, d2 = slot + L.vstart - b1, d3 = name_sm - b2;  // source = position + delta
      if (L.first_live) {  // the bytes before my first aligned chunk: nobody else writes them
        for (uint32_t q = g; q < start && q < b2; ++q) galn[q] = smem[lay.names + tile.family * kNameStride + 16 + (q - g)];
      }
      uint32_t p = start, e = 0;  // `a4` holds the chunk at p, of which the first `e` bytes are already final
      uint4 a4 = make_uint4(0, 0, 0, 0);
      // A 16-byte store is half a 32-byte sector on its way to L2, and the stream of partial sectors is what the writer
      // waits for.  So a finished chunk in the lower half of a sector waits in `lo4` for its upper neighbour and the
      // two leave in one 256-bit store.  Straight-line code: the stores are predicated, the pending half is a select.
      uint4 lo4 = make_uint4(0, 0, 0, 0);
      auto put = [&](const uint4& c) {  // the finished chunk at position p
        const uint32_t odd = p & 16u;
        st_pair_or_single(galn + p, lo4, c, odd != 0 && p > start, odd != 0 && p == start);
        lo4.x = odd ? lo4.x : c.x; lo4.y = odd ? lo4.y : c.y; lo4.z = odd ? lo4.z : c.z; lo4.w = odd ? lo4.w : c.w;
        p += 16;
      };
      auto append = [&](uint32_t d, uint32_t b) {  // source = position + d, up to position b (exclusive)
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
          put(a4);
          step();
          while (p + 16 <= b) {
            put(chunk());
            step();
          }
          a4 = chunk();
        }
        e = b - p;
      };
      append(d0, b0);  // metric name
      append(d1, b1);  // labels
      append(d2, b2);  // " <value>\n"
      if (!L.last_live) append(d3, stop);  // the head of the next line's name completes my last chunk
      if ((p & 16u) && p > start) stg128_cs(galn + p - 16, lo4);  // a lower half whose neighbour is not mine
      if (L.last_live && e) {      // the tile ends inside a chunk: its bytes belong to me, the rest to the next tile
        s_tail = a4;
        const uint8_t* sp = reinterpret_cast<const uint8_t*>(&s_tail);
        for (uint32_t q = 0; q < e; ++q) galn[p + q] = sp[q];
      }
    }
    if (tid == 0 && more) fetch_desc((k + 2) % 3);  // nobody reads that slot before the end-of-tile barrier
  };

  // ---- prologue: first tile measured, second ticket drawn
  if (tid == 0) { tk_next = atomicAdd(ticket, 1ull); fetch_desc(0); }
  __syncthreads();
  if (s_tk[0] >= R.n_tiles) return;
  if (wid == 0) issue_labels(0);
  LineState cur, nxt;
  measure(0, 0, cur);
  if (tid == 0) { tk_next = atomicAdd(ticket, 1ull); fetch_desc(1); }
  __syncthreads();

  uint32_t d = 0, sb = 0;  // descriptor slot and value-slot buffer of the tile being emitted
  for (;;) {
    const uint32_t d1 = d == 2 ? 0 : d + 1;
    const bool more = s_tk[d1] < R.n_tiles;  // CTA-uniform
    if (more) measure(d1, sb ^ 1, nxt);
    KGX_PC(1);  // measure the next tile (state loads + format + scan)
    emit(d, sb, cur, more);
    KGX_PC(5);  // stream my line (+ thread 0: descriptor of the tile after next)
    __syncthreads();  // label stage and slot buffer sb are free; the new descriptor is visible
    KGX_PC(6);  // end-of-tile barrier
    if (!more) { KGX_PC_FLUSH; return; }
    if (wid == 0) issue_labels(d1);  // labels of the measured tile
    KGX_PC(7);
#ifdef KGX_PHASE_CLOCKS
    ++pc_tiles;
#endif
    cur = nxt;
    d = d1;
    sb ^= 1;
  }
}

}  // namespace kgx
