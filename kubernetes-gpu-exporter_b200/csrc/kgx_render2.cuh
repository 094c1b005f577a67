// kgx_render2.cuh — exposition writer, second generation (sm_100a).
//
// Same contract as k_render (byte-identical output; replaces what main.go:69-71
// serves through Registry.Gather + expfmt), rebuilt around how the bytes move:
//
//   * the label text of a tile is ONE contiguous byte range of the series label
//     blob: a single TMA bulk copy (cp.async.bulk global->shared, completion on
//     an mbarrier) stages it while the threads format their values;
//   * each thread owns one line and formats its value once, straight into a
//     shared-memory slot (" <value>\n");
//   * line lengths -> CTA scan -> inter-tile decoupled look-back give every
//     line its exact byte offset;
//   * each thread then streams its line into the shared output buffer as
//     16-byte chunks ALIGNED TO THE GLOBAL ADDRESS: a chunk belongs to the line
//     that holds its first byte, so every store is a full STS.128 and no two
//     threads touch the same chunk.  Sources (metric name, staged labels, value
//     slot, the next line's name prefix) are read with word loads + funnel
//     shifts; only the 2-3 chunks that straddle two sources need a byte merge;
//   * the finished tile leaves with one TMA bulk store (shared->global); the
//     few bytes before/after the 16-byte-aligned body go out as byte stores.
#pragma once
#include "kgx_kernels.cuh"

namespace kgx {

struct RenderLayout {  // byte offsets inside dynamic shared memory (all 16-aligned)
  uint32_t names;      // kMaxFamilies x 96: [16 pad][name (<=64)][16 pad]
  uint32_t slots;      // tile_lines x 28: " <value>\n"
  uint32_t stage;      // staged label text (+32 slack either side)
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
  uint4 m;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(m.x), "=r"(m.y), "=r"(m.z), "=r"(m.w) : "r"(mask_sm + 16u * e));
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
// TMA bulk copy shared -> global
__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS) k_render2(DevRender R, RenderLayout lay, DevAcc acc, unsigned long long* counters) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ unsigned long long s_scan[THREADS / 32];
  __shared__ uint32_t s_tile;
  __shared__ unsigned long long s_base;
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ __align__(16) uint32_t s_mask[16][4];
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
  uint32_t parity = 0;
  for (;;) {
    if (tid == 0) s_tile = uint32_t(atomicAdd(&counters[C_TICKET], 1ull));
    __syncthreads();
    const uint32_t t = s_tile;
    if (t >= R.n_tiles) {
      if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // my bulk stores have fully landed
      return;
    }
    const Tile tile = R.tiles[t];
    const FamMeta& fm = R.fam[tile.family];
    const uint32_t name_len = fm.name_len;
    // ---- stage the tile's label text: one bulk copy of an aligned superset
    const unsigned long long L0 = R.lbl_off[tile.first];
    const unsigned long long g0 = L0 & ~127ull;  // whole 128-byte lines: the copy engine's natural unit
    if (R.flags & 2) {  // staged by every thread with 16-byte cp.async (LDGSTS) instead of the TMA engine
      const unsigned long long L1 = R.lbl_off[tile.first + tile.count];
      const uint32_t chunks = uint32_t(((L1 + 15ull) & ~15ull) - g0) >> 4;
      const uint8_t* src = R.lbl_blob + g0;
      for (uint32_t i = tid; i < chunks; i += THREADS)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sm0 + lay.stage + 32 + 16 * i), "l"(src + 16 * size_t(i)) : "memory");
      asm volatile("cp.async.commit_group;" ::: "memory");
    } else if (tid == 0) {
      const unsigned long long L1 = R.lbl_off[tile.first + tile.count];
      const uint32_t bytes = uint32_t(((L1 + 15ull) & ~15ull) - g0);
      if (bytes) {
        mbar_expect_tx(bar, bytes);
        bulk_g2s(sm0 + lay.stage + 32, R.lbl_blob + g0, bytes, bar);
      } else {
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
      }
    }
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
    __syncthreads();
    unsigned long long woff = 0, all = 0;
    for (int k = 0; k < THREADS / 32; ++k) { if (k < int(wid)) woff += s_scan[k]; all += s_scan[k]; }
    const unsigned long long exc = woff + inc - mine;
    const uint32_t off = hdr + uint32_t(exc);
    const uint32_t total = hdr + uint32_t(all);
    const bool first_live = live && uint32_t(exc >> 32) == 0;
    const bool last_live = live && uint32_t(exc >> 32) + 1 == uint32_t(all >> 32);
    // ---- inter-tile offset: decoupled look-back over the tile chain
    // Warp 0 inspects 32 predecessors per step: everything after the nearest tile
    // that already knows its inclusive prefix contributes its own size.
    if (wid == 0) {
      unsigned long long base = 0;
      if (t > 0) {
        if (lane == 0) atomicExch(&R.tile_state[t], kStAgg | total);
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
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // previous tile's bulk store has read the output buffer
        if (tile.flags & 2) R.node_body_off[tile.node] = base;
        if (t == R.n_tiles - 1) counters[C_TOTAL_LEN] = base + total;
      }
    }
    if (R.flags & 2) asm volatile("cp.async.wait_group 0;" ::: "memory");  // my share of the label copy has landed
    __syncthreads();
    const unsigned long long base = s_base;
    const uint32_t sh = uint32_t(base & 15);   // shared offset == global address (mod 16)
    const uint32_t outb = sm0 + lay.out;       // byte p of the tile lives at outb + sh + p
    if (!(R.flags & 2)) { mbar_wait(bar, parity); parity ^= 1; }  // staged labels have landed
    // ---- pass 2: stream my line into the output buffer as aligned 16-byte chunks
    if (live) {
      const uint32_t g = outb + sh + off;                    // first byte of my line
      const uint32_t b0 = g + name_len, b1 = b0 + lbl_len, b2 = b1 + vlen + 2;  // source boundaries; b2 = line end
      const uint32_t start = first_live ? (g & ~15u) : ((g + 15u) & ~15u);
      const uint32_t stop = (b2 + 15u) & ~15u;               // my chunks: [start, stop)
      const uint32_t name_sm = sm0 + lay.names + tile.family * kNameStride + 16;
      const uint32_t d0 = name_sm - g, d1 = lbl_src - b0, d2 = slot - b1, d3 = name_sm - b2;  // source = position + delta
      (void)last_live;  // after the last line the chunk tail is slack: never copied out
      // The four sources are appended one after the other, so that all lanes of
      // a warp are in the same source at the same time (no per-chunk branching
      // on where the boundaries fall): `acc` holds the chunk at p, of which the
      // first `e` bytes are already final.
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
      append(d0, b0);
      append(d1, b1);
      append(d2, b2);
      append(d3, stop);
    }
    if (hdr) {  // the header goes in after the lines: the first line's leading chunk overlaps its tail
      __syncthreads();
      for (uint32_t i = tid; i < hdr; i += THREADS) smem[lay.out + sh + i] = fm.header[i];
    }
    if (R.flags & 1) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the bulk store
    __syncthreads();
    // ---- the tile leaves: byte head, bulk body, byte tail
    uint8_t* gout = R.out + base;
    const uint32_t head = min(total, (16u - sh) & 15u);
    const uint32_t body = (total - head) & ~15u;
    if (R.flags & 1) {  // body through the TMA engine (one bulk store)
      if (tid == 0 && body) {
        bulk_s2g(gout + head, outb + sh + head, body);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    } else {            // body through the LSU: 16 bytes per lane, fully coalesced
      const uint4* sb = reinterpret_cast<const uint4*>(smem + lay.out + sh + head);
      uint4* gb = reinterpret_cast<uint4*>(gout + head);
      for (uint32_t i = tid; i < body / 16; i += THREADS) __stcs(gb + i, sb[i]);
    }
    if (tid < head) gout[tid] = smem[lay.out + sh + tid];
    if (tid >= 16 && tid < 16 + (total - head - body)) {
      const uint32_t i = head + body + tid - 16;
      gout[i] = smem[lay.out + sh + i];
    }
    // no barrier here: the next tile's first barriers order everything, and the bulk
    // store's reads of the output buffer are awaited right before it is rewritten
  }
}

}  // namespace kgx
