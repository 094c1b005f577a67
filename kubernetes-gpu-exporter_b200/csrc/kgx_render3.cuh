// kgx_render3.cuh — exposition writer, direct-store variant (sm_100a).
//
// Same contract and same per-line streaming as k_render2 (kgx_render2.cuh), but
// the tile text is never staged in shared memory: once the tile's global byte
// offset is known, every lane stores its 16-byte chunks — aligned to the global
// address — straight to the output with streaming STG.128.  Two lanes never
// touch the same chunk (a chunk belongs to the line that holds its first byte),
// so the only partial chunks of a tile are the head of its first line and the
// tail of its last one; those few bytes go out as byte stores.
// What it buys: no output buffer (about 60 % of k_render2's shared memory, so
// twice the resident warps), no flush pass, one barrier fewer per tile.
// What it costs: 16-byte stores that do not coalesce across lanes.
#pragma once
#include "kgx_render2.cuh"

namespace kgx {

__device__ __forceinline__ void stg128_cs(uint8_t* p, uint4 v) {
  asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <int THREADS, bool STAGE>
__global__ void __launch_bounds__(THREADS) k_render3(DevRender R, RenderLayout lay, DevAcc acc, unsigned long long* counters) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ unsigned long long s_scan[THREADS / 32];
  __shared__ uint4 s_tail;
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ __align__(16) uint32_t s_mask[16][4];
  __shared__ uint32_t s_next;
  __shared__ Tile s_desc;
  __shared__ uint32_t s_l0;  // where the tile's first label starts inside the staged range
  const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint32_t sm0 = uint32_t(__cvta_generic_to_shared(smem));
  const uint32_t bar = uint32_t(__cvta_generic_to_shared(&s_bar));
  for (uint32_t i = tid; i < kMaxFamilies * kNameStride; i += THREADS) {
    const uint32_t f = i / kNameStride, o = i % kNameStride;
    smem[lay.names + i] = (o >= 16 && o < 16 + 64 && f < R.n_families) ? R.fam[f].name[o - 16] : uint8_t('#');
  }
  for (uint32_t i = tid; i < 64; i += THREADS) {
    const uint32_t e = i >> 2, w = i & 3;
    const int n = int(e) - int(4 * w);
    s_mask[e][w] = n <= 0 ? 0u : (n >= 4 ? 0xffffffffu : ((1u << (8 * n)) - 1u));
  }
  const uint32_t mask_sm = uint32_t(__cvta_generic_to_shared(&s_mask[0][0]));
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  auto fetch_next = [&]() {  // warp 0: next ticket, its descriptor, and the bulk copy of its label text
    uint32_t nt = 0;
    if (lane == 0) nt = uint32_t(atomicAdd(&counters[C_TICKET], 1ull));
    nt = __shfl_sync(0xffffffffu, nt, 0);
    if (nt < R.n_tiles && lane == 0) {
      const Tile d = R.tiles[nt];
      const unsigned long long L0 = R.lbl_off[d.first], L1 = R.lbl_off[d.first + d.count];
      const unsigned long long g0 = L0 & ~127ull;
      const uint32_t bytes = uint32_t(((L1 + 15ull) & ~15ull) - g0);
      if (bytes) {
        mbar_expect_tx(bar, bytes);
        bulk_g2s(sm0 + lay.stage + 32, R.lbl_blob + g0, bytes, bar);
      } else {
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
      }
      s_desc = d;
      s_l0 = uint32_t(L0 - g0);
    }
    if (lane == 0) s_next = nt;
  };
  if (wid == 0) fetch_next();
  __syncthreads();

  uint32_t parity = 0;
  for (;;) {
    const uint32_t t = s_next;
    if (t >= R.n_tiles) {
      if (STAGE && tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // my bulk stores have landed
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
      const double v = series_value(fm.kind, acc, R, r, live);
      if (live) {
        uint8_t* sp = smem + lay.slots + tid * kSlotBytes;
        sp[0] = ' ';
        vlen = uint32_t(format_value(v, sp + 1));
        sp[1 + vlen] = '\n';
        len = name_len + lbl_len + vlen + 2;
      }
    }
    const uint32_t hdr = ((tile.flags & 1) && acc.node_live[tile.node] != 0) ? fm.header_len : 0;
    // ---- exclusive scan over the CTA, three fields at once: output bytes (24 bits), label bytes of
    // every series, printed or not (24 bits), lines printed (16 bits)
    const unsigned long long mine = (uint64_t(live) << 48) | (uint64_t(lbl_len) << 24) | len;
    unsigned long long inc = mine;
    for (int o = 1; o < 32; o <<= 1) { unsigned long long u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
    if (lane == 31) s_scan[wid] = inc;
    if (STAGE && tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // last tile's bulk store has read the tile buffer
    __syncthreads();  // (1) also: everyone has read s_next / s_desc / s_g0
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
    if (tid == 0 && t > 0) atomicExch(&R.tile_state[t], kStAgg | total);
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
    if (tid == 0) {
      atomicExch(&R.tile_state[t], kStPrefix | (base + total));
      if (tile.flags & 2) R.node_body_off[tile.node] = base;
      if (t == R.n_tiles - 1) counters[C_TOTAL_LEN] = base + total;
    }
    uint8_t* gtile = R.out + base;                    // first byte of the tile in the output
    uint8_t* galn = R.out + (base & ~15ull);          // 16-aligned origin: position v lives at galn + v
    const uint32_t sh = uint32_t(base & 15);
    const uint32_t outb = sm0 + lay.out;  // STAGE: position v lives at outb + v in shared memory
    if (!STAGE && hdr) for (uint32_t i = tid; i < hdr; i += THREADS) gtile[i] = fm.header[i];
    mbar_wait(bar, parity);  // staged labels have landed (issued a whole tile ago)
    parity ^= 1;
    // ---- pass 2: stream my line to the output as 16-byte chunks aligned to the global address
    if (live) {
      const uint32_t g = sh + off;                                              // first byte of my line (position)
      const uint32_t b0 = g + name_len, b1 = b0 + lbl_len, b2 = b1 + vlen + 2;  // source boundaries; b2 = line end
      const uint32_t start = (STAGE && first_live) ? (g & ~15u) : ((g + 15u) & ~15u);
      const uint32_t stop = (!STAGE && last_live) ? (b2 & ~15u) : ((b2 + 15u) & ~15u);  // my chunks: [start, stop)
      const uint32_t name_sm = sm0 + lay.names + tile.family * kNameStride + 16;
      const uint32_t d0 = name_sm - g, d1 = lbl_src - b0, d2 = slot - b1, d3 = name_sm - b2;  // source = position + delta
      if (!STAGE && first_live) {  // the bytes before my first aligned chunk: nobody else writes them
        for (uint32_t q = g; q < start && q < b2; ++q) galn[q] = smem[lay.names + tile.family * kNameStride + 16 + (q - g)];
      }
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
        a4 = e ? merge16(a4, nw, e, mask_sm) : nw;
        if (b >= p + 16) {
          if (STAGE) sts128(outb + p, a4); else stg128_cs(galn + p, a4);
          p += 16;
          step();
          while (p + 16 <= b) {
            const uint4 c4 = chunk();
            if (STAGE) sts128(outb + p, c4); else stg128_cs(galn + p, c4);
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
      if (STAGE || !last_live) {
        append(d3, stop);  // the head of the next line's name completes my last chunk
      } else if (e) {      // the tile ends inside a chunk: its bytes belong to me, the rest to the next tile
        s_tail = a4;  // one thread per tile gets here: park the chunk, then byte-copy
        const uint8_t* sp = reinterpret_cast<const uint8_t*>(&s_tail);
        for (uint32_t q = 0; q < e; ++q) galn[p + q] = sp[q];
      }
    }
    __syncthreads();  // (2) label stage and value slots are free
    if (STAGE) {
      if (hdr) for (uint32_t i = tid; i < hdr; i += THREADS) smem[lay.out + sh + i] = fm.header[i];
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // my writes to the tile buffer -> visible to the bulk store
    }
    if (wid == 0) fetch_next();  // next ticket, descriptor, label copy
    __syncthreads();  // (3) s_next / s_desc / s_g0 visible
    if (STAGE) {  // the tile leaves: byte head, one TMA bulk store for the 16-byte aligned body, byte tail
      const uint32_t head = min(total, (16u - sh) & 15u);
      const uint32_t body = (total - head) & ~15u;
      if (tid == 0 && body) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gtile + head), "r"(outb + sh + head), "r"(body) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      if (tid >= 32 && tid < 32 + head) gtile[tid - 32] = smem[lay.out + sh + tid - 32];
      if (tid >= 48 && tid < 48 + (total - head - body)) {
        const uint32_t i = head + body + tid - 48;
        gtile[i] = smem[lay.out + sh + i];
      }
    }
  }
}

}  // namespace kgx
