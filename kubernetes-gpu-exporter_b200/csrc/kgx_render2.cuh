// kgx_render2.cuh — exposition writer, second generation (sm_100a).
//
// Same contract as k_render (byte-identical output; replaces what main.go:69-71
// serves through Registry.Gather + expfmt), rebuilt around how the bytes move
// and around what has to wait for what:
//
//   * the label text of a tile is ONE contiguous byte range of the series label
//     blob: a single TMA bulk copy (cp.async.bulk global->shared, completion on
//     an mbarrier) stages it.  Warp 0 issues it for the NEXT tile — together
//     with that tile's ticket and descriptor — while the CTA is still flushing
//     the current one, so neither the ticket, nor the descriptor, nor the
//     labels are waited for;
//   * each thread owns one line and formats its value once, straight into a
//     shared-memory slot (" <value>\n");
//   * line lengths -> CTA scan; the tile's size is published to the look-back
//     chain at once, so successors never wait for more than this;
//   * each thread streams its line into the shared output buffer as aligned
//     16-byte chunks: a chunk belongs to the line that holds its first byte, so
//     every store is a full STS.128 and no two threads touch the same chunk.
//     The four sources of a line (metric name, staged labels, value slot, the
//     next line's name prefix) are appended one after the other, all lanes in
//     lock-step; word loads + funnel shifts realign, a mask table merges the
//     one chunk that straddles two sources;
//   * only then does warp 0 resolve the tile's global byte offset (decoupled
//     look-back, 32 predecessors per step): by now they have all published;
//   * the tile leaves through the LSU: every lane re-aligns 16 bytes from the
//     tile buffer to the global address (2 x LDS.128 + funnel shift) and issues
//     one streaming STG.128; the <16 bytes at either end go as byte stores.
#pragma once
#include "kgx_kernels.cuh"

namespace kgx {

struct RenderLayout {  // byte offsets inside dynamic shared memory (all 16-aligned)
  uint32_t names;      // kMaxFamilies x 96: [16 pad][name (<=64)][16 pad]
  uint32_t slots;      // THREADS x 28: " <value>\n"
  uint32_t stage;      // staged label text (+32 slack before, +128+48 after)
  uint32_t out;        // the tile's output text
  uint32_t total;
};
constexpr uint32_t kSlotBytes = 28;  // 7 words: odd word stride -> conflict-free byte stores
constexpr uint32_t kNameStride = 96;

__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// 16 bytes starting at shared byte address `a` (any alignment)
__device__ __forceinline__ uint4 gather16(uint32_t a) {
  const uint32_t b = a & ~3u, r = (a & 3u) * 8u;
  const uint32_t w0 = lds32(b), w1 = lds32(b + 4), w2 = lds32(b + 8), w3 = lds32(b + 12), w4 = lds32(b + 16);
  return make_uint4(__funnelshift_r(w0, w1, r), __funnelshift_r(w1, w2, r), __funnelshift_r(w2, w3, r),
                    __funnelshift_r(w3, w4, r));
}
// bytes [0,e) from lo, bytes [e,16) from hi (0 < e < 16); mask_sm: shared table of
// 16 byte masks, entry e = the first e bytes set
__device__ __forceinline__ uint4 merge16(uint4 lo, uint4 hi, uint32_t e, uint32_t mask_sm) {
  (void)mask_sm;  // masks are computed: the shared-memory pipe is the writer's bottleneck, the ALU is not
  auto mk = [e](int w) -> uint32_t {  // the bytes of word w that are < e
    const int n = min(max(int(e) - 4 * w, 0), 4);
    return __funnelshift_rc(0xffffffffu, 0u, uint32_t(32 - 8 * n));
  };
  const uint4 m = make_uint4(mk(0), mk(1), mk(2), mk(3));
  return make_uint4((lo.x & m.x) | (hi.x & ~m.x), (lo.y & m.y) | (hi.y & ~m.y), (lo.z & m.z) | (hi.z & ~m.z),
                    (lo.w & m.w) | (hi.w & ~m.w));
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
// TMA bulk copy global -> shared, completion counted on the mbarrier
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS) k_render2(DevRender R, RenderLayout lay, DevAcc acc, unsigned long long* counters) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ unsigned long long s_scan[THREADS / 32];
  __shared__ unsigned long long s_base;
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ __align__(16) uint32_t s_mask[16][4];
  __shared__ uint32_t s_next;          // ticket of the tile to process next
  __shared__ Tile s_desc;              // its descriptor
  __shared__ unsigned long long s_g0;  // start of its staged label range in the blob (128-aligned)
  const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint32_t sm0 = uint32_t(__cvta_generic_to_shared(smem));
  const uint32_t bar = uint32_t(__cvta_generic_to_shared(&s_bar));
  // one-time: metric names of every family, with slack on both sides
  for (uint32_t i = tid; i < kMaxFamilies * kNameStride; i += THREADS) {
    const uint32_t f = i / kNameStride, o = i % kNameStride;
    smem[lay.names + i] = (o >= 16 && o < 16 + 64 && f < R.n_families) ? R.fam[f].name[o - 16] : uint8_t('#');
  }
  if (tid < 64) {  // byte-mask table for merge16: entry e has the first e bytes set
    const uint32_t e = tid >> 2, w = tid & 3;
    const int n = int(e) - int(4 * w);
    s_mask[e][w] = n <= 0 ? 0u : (n >= 4 ? 0xffffffffu : ((1u << (8 * n)) - 1u));
  }
  const uint32_t mask_sm = uint32_t(__cvta_generic_to_shared(&s_mask[0][0]));
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // Warp 0: take the next ticket (tiles are handed out in output order, so the
  // look-back never waits on a tile that has not started), read its descriptor
  // and start the bulk copy of its label text.
  auto fetch_next = [&]() {
    uint32_t nt = 0;
    if (lane == 0) nt = uint32_t(atomicAdd(&counters[C_TICKET], 1ull));
    nt = __shfl_sync(0xffffffffu, nt, 0);
    if (nt < R.n_tiles && lane == 0) {
      const Tile d = R.tiles[nt];
      const unsigned long long L0 = R.lbl_off[d.first], L1 = R.lbl_off[d.first + d.count];
      const unsigned long long g0 = L0 & ~127ull;  // whole 128-byte lines
      const uint32_t bytes = uint32_t(((L1 + 15ull) & ~15ull) - g0);
      if (bytes) {
        mbar_expect_tx(bar, bytes);
        bulk_g2s(sm0 + lay.stage + 32, R.lbl_blob + g0, bytes, bar);
      } else {
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
      }
      s_desc = d;
      s_g0 = g0;
    }
    if (lane == 0) s_next = nt;
  };
  if (wid == 0) fetch_next();
  __syncthreads();

  uint32_t parity = 0;
  const uint32_t outb = sm0 + lay.out;  // byte p of the tile lives at outb + p
  for (;;) {
    const uint32_t t = s_next;
    if (t >= R.n_tiles) return;
    const Tile tile = s_desc;
    const unsigned long long g0 = s_g0;
    const FamMeta& fm = R.fam[tile.family];
    const uint32_t name_len = fm.name_len;
    // ---- pass 1: value text and exact length of my line
    const uint32_t slot = sm0 + lay.slots + tid * kSlotBytes;
    uint32_t len = 0, lbl_len = 0, vlen = 0, lbl_src = 0;
    bool live = false;
    if (tid < tile.count) {
      const uint32_t r = tile.first + tid;
      const double v = series_value(fm.kind, acc, R, r, live);
      if (live) {
        uint8_t* sp = smem + lay.slots + tid * kSlotBytes;
        sp[0] = ' ';
        vlen = uint32_t(format_value(v, sp + 1));
        sp[1 + vlen] = '\n';
        const unsigned long long at = R.lbl_off[r];
        lbl_len = uint32_t(R.lbl_off[r + 1] - at);
        lbl_src = sm0 + lay.stage + 32 + uint32_t(at - g0);
        len = name_len + lbl_len + vlen + 2;
      }
    }
    const uint32_t hdr = ((tile.flags & 1) && acc.node_live[tile.node] != 0) ? fm.header_len : 0;
    // ---- exclusive scan over the CTA: low word = bytes, high word = live lines
    const unsigned long long mine = (uint64_t(live) << 32) | len;
    unsigned long long inc = mine;
    for (int o = 1; o < 32; o <<= 1) { unsigned long long u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
    if (lane == 31) s_scan[wid] = inc;
    __syncthreads();  // (1) also: everyone has read s_next / s_desc / s_g0
    unsigned long long woff = 0, all = 0;
    for (int k = 0; k < THREADS / 32; ++k) { if (k < int(wid)) woff += s_scan[k]; all += s_scan[k]; }
    const unsigned long long exc = woff + inc - mine;
    const uint32_t off = hdr + uint32_t(exc);
    const uint32_t total = hdr + uint32_t(all);
    const bool first_live = live && uint32_t(exc >> 32) == 0;
    // the tile's size is all a successor needs from us until it resolves its own offset
    if (tid == 0 && t > 0) atomicExch(&R.tile_state[t], kStAgg | total);
    mbar_wait(bar, parity);  // staged labels have landed (issued a whole tile ago)
    parity ^= 1;
    // ---- pass 2: stream my line into the tile buffer as aligned 16-byte chunks
    if (live) {
      const uint32_t g = outb + off;                                            // first byte of my line
      const uint32_t b0 = g + name_len, b1 = b0 + lbl_len, b2 = b1 + vlen + 2;  // source boundaries; b2 = line end
      const uint32_t start = first_live ? (g & ~15u) : ((g + 15u) & ~15u);
      const uint32_t stop = (b2 + 15u) & ~15u;  // my chunks: [start, stop)
      const uint32_t name_sm = sm0 + lay.names + tile.family * kNameStride + 16;
      const uint32_t d0 = name_sm - g, d1 = lbl_src - b0, d2 = slot - b1, d3 = name_sm - b2;  // source = position + delta
      // `a4` holds the chunk at p, of which the first `e` bytes are already final.
      uint32_t p = start, e = 0;
      uint4 a4 = make_uint4(0, 0, 0, 0);
      auto append = [&](uint32_t d, uint32_t b) {  // source = position + d, up to position b (exclusive)
        const uint4 nw = gather16(p + d);
        a4 = e ? merge16(a4, nw, e, mask_sm) : nw;
        if (b >= p + 16) {
          sts128(p, a4);
          p += 16;
          while (p + 16 <= b) { sts128(p, gather16(p + d)); p += 16; }
          a4 = gather16(p + d);
        }
        e = b - p;
      };
      append(d0, b0);    // metric name (the first line of a tile also owns the chunk its line starts in)
      append(d1, b1);    // labels
      append(d2, b2);    // " <value>\n"
      append(d3, stop);  // the head of the next line's name completes my last chunk
    }
    __syncthreads();  // (2) tile text complete; label stage and value slots are free
    if (hdr) for (uint32_t i = tid; i < hdr; i += THREADS) smem[lay.out + i] = fm.header[i];
    // ---- global byte offset of the tile: decoupled look-back, 32 predecessors per step
    if (wid == 0) {
      unsigned long long base = 0;
      if (t > 0) {
        long long p = (long long)t - 1 - lane;  // lane 0 looks at the direct predecessor
        for (;;) {
          unsigned long long s = kStPrefix;     // before tile 0: an inclusive prefix of zero
          if (p >= 0) {
            do { s = *(volatile unsigned long long*)&R.tile_state[p]; } while ((s & kStMask) == 0);
          }
          const uint32_t pfx = __ballot_sync(0xffffffffu, (s & kStMask) == kStPrefix);
          const uint32_t first = pfx ? uint32_t(__ffs(pfx) - 1) : 32u;
          unsigned long long v = lane <= first ? (s & ~kStMask) : 0ull;
          for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
          base += v;
          if (pfx) break;
          p -= 32;
        }
      }
      if (lane == 0) {
        atomicExch(&R.tile_state[t], kStPrefix | (base + total));
        s_base = base;
        if (tile.flags & 2) R.node_body_off[tile.node] = base;
        if (t == R.n_tiles - 1) counters[C_TOTAL_LEN] = base + total;
      }
    }
    __syncthreads();  // (3)
    const unsigned long long base = s_base;
    if (wid == 0) fetch_next();  // next ticket, descriptor and label copy overlap the flush
    // ---- the tile leaves: re-align to the global address, 16 bytes per lane
    uint8_t* gout = R.out + base;
    const uint32_t head = min(total, uint32_t((16u - uint32_t(base & 15)) & 15u));
    const uint32_t nbody = (total - head) >> 4;
    const uint32_t tailb = total - head - 16u * nbody;
    {
      const uint32_t src0 = outb + head;  // shared address of the first body byte: same shift for the whole tile
      const uint32_t q = (src0 >> 2) & 3u, r = (src0 & 3u) * 8u;
      uint4* gb = reinterpret_cast<uint4*>(gout + head);
      for (uint32_t i = tid; i < nbody; i += THREADS) {
        const uint32_t a = (src0 & ~15u) + 16u * i;
        const uint4 x = lds128(a), y = lds128(a + 16);
        uint32_t w0, w1, w2, w3, w4;  // the five words that cover the 16 wanted bytes
        if (q == 0) { w0 = x.x; w1 = x.y; w2 = x.z; w3 = x.w; w4 = y.x; }
        else if (q == 1) { w0 = x.y; w1 = x.z; w2 = x.w; w3 = y.x; w4 = y.y; }
        else if (q == 2) { w0 = x.z; w1 = x.w; w2 = y.x; w3 = y.y; w4 = y.z; }
        else { w0 = x.w; w1 = y.x; w2 = y.y; w3 = y.z; w4 = y.w; }
        __stcs(gb + i, make_uint4(__funnelshift_r(w0, w1, r), __funnelshift_r(w1, w2, r), __funnelshift_r(w2, w3, r),
                                  __funnelshift_r(w3, w4, r)));
      }
    }
    if (tid < head) gout[tid] = smem[lay.out + tid];
    if (tid >= 16 && tid < 16 + tailb) {
      const uint32_t i = head + 16u * nbody + tid - 16;
      gout[i] = smem[lay.out + i];
    }
    __syncthreads();  // (4) tile buffer free; s_next / s_desc / s_g0 of the next tile visible
  }
}

}  // namespace kgx
