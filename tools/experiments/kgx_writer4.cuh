// kgx_writer4.cuh — the exposition writer, chunk-major streaming (sm_100a).
//
// EXPERIMENT, NOT BUILT INTO libkgx.so.  It passed the GPU parity suite
// (tests/test_gpu_parity.py, tests/test_gpu_random_differential.py) but measured
// SLOWER than k_render3 on the bench workload: 2.24 ms (64-line tiles) / 2.14 ms
// (128) against 1.61 ms.  The coalesced stores do what they promise (global-store
// wavefronts 71 M -> 18 M per launch), but the shared-memory side pays it back:
// 2 LDS.128 per chunk at 2.1x bank conflicts (a quarter-warp of 8 consecutive
// chunks spans label | glue | label, three unrelated address ranges), a second
// pair for the 23 % of chunks that straddle a segment boundary, plus the chunk
// map.  Total L1 data-pipe wavefronts per SM end up the same (2.56 M) with 40 %
// more instructions.  ncu: profiles/r01_k_render4_chunk_major_raw.csv.
// To build it again: include it from kgx_api.cu, give RenderLayout4 the sizes
// listed in the struct comments (loc: one u16 per 16 output bytes of the
// largest tile), launch with the same arguments as k_render3.
//
// Same contract and same per-tile front end as k_render3 (kgx_writer.cuh: TMA-staged
// label text, one value formatted per thread, CTA scan, decoupled look-back), but
// the streaming stage is turned by 90 degrees.  k_render3 lets every lane walk its
// OWN line, so a warp's 16-byte stores land in 32 different 128-byte lines and its
// word gathers from shared memory collide in the banks: 14.5 L1 data-pipe
// wavefronts per 128 bytes of text, the pipe 81 % busy.  Here consecutive lanes
// take CONSECUTIVE 16-byte chunks of the tile's output:
//
//   * a tile's text is the concatenation  name | labels_0 | glue_0 | labels_1 |
//     glue_1 | ...  where glue_i = " <value_i>\n" + metric name (the head of the
//     next line), written by thread i next to its value; every segment is at
//     least 15 bytes long, so a chunk overlaps at most two of them;
//   * thread i records where its two segments start in the output and in shared
//     memory (`seg`), and marks, for every chunk whose first byte falls into one
//     of them, the segment number (`loc`);
//   * chunk c: segment = loc[c]; 16 bytes from it (two aligned LDS.128 + shift),
//     merged with the head of the following segment when that starts inside the
//     chunk; one STG.128.  A warp reads 512 contiguous bytes of shared memory
//     and writes 512 contiguous bytes of the output per iteration.
#pragma once
#include "../../kubernetes-gpu-exporter_b200/csrc/kgx_writer.cuh"

namespace kgx {

struct RenderLayout4 {  // byte offsets inside dynamic shared memory (all 16-aligned)
  uint32_t names;       // 32 pad + kMaxFamilies x 96 ([16 pad][name (<=64)][16 pad]) + 48 pad
  uint32_t slots;       // THREADS x slot_bytes: " <value>\n<name>"
  uint32_t slot_bytes;  // odd number of words >= 28 + longest name + 4
  uint32_t seg;         // (2 THREADS + 4) x {position, shared address}
  uint32_t loc;         // u16 per chunk of the tile: the segment holding its first byte
  uint32_t stage;       // staged label text (+32 slack before, +128+48 after)
  uint32_t total;
};

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t lds16(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ void sts64(uint32_t addr, uint32_t a, uint32_t b) {
  asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void sts16(uint32_t addr, uint32_t v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }

// 16 bytes from any shared-memory byte address: the two aligned quadwords around it, shifted
__device__ __forceinline__ uint4 load16(uint32_t addr) {
  const uint32_t a = addr & ~15u;
  const uint4 x = lds128(a), y = lds128(a + 16);
  const bool k1 = (addr & 4u) != 0, k2 = (addr & 8u) != 0;
  const uint32_t t0 = k1 ? x.y : x.x, t1 = k1 ? x.z : x.y, t2 = k1 ? x.w : x.z, t3 = k1 ? y.x : x.w, t4 = k1 ? y.y : y.x,
                 t5 = k1 ? y.z : y.y, t6 = k1 ? y.w : y.z;
  const uint32_t u0 = k2 ? t2 : t0, u1 = k2 ? t3 : t1, u2 = k2 ? t4 : t2, u3 = k2 ? t5 : t3, u4 = k2 ? t6 : t4;
  const uint32_t s = (addr & 3u) * 8u;
  return make_uint4(__funnelshift_r(u0, u1, s), __funnelshift_r(u1, u2, s), __funnelshift_r(u2, u3, s), __funnelshift_r(u3, u4, s));
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS) k_render4(DevRender R, RenderLayout4 lay, DevAcc acc, unsigned long long* counters) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ unsigned long long s_scan[THREADS / 32];
  __shared__ uint4 s_tail;
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ uint32_t s_next;  // ticket of the tile to process next
  __shared__ Tile s_desc;      // its descriptor
  __shared__ uint32_t s_l0;    // where its first label starts inside the staged range
  const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint32_t sm0 = uint32_t(__cvta_generic_to_shared(smem));
  const uint32_t bar = uint32_t(__cvta_generic_to_shared(&s_bar));
  const uint32_t names0 = lay.names + 32;  // family f: [16 pad][name][16 pad] at names0 + f * kNameStride
  // one-time: metric names of every family, with slack on both sides
  for (uint32_t i = tid; i < 32 + kMaxFamilies * kNameStride + 48; i += THREADS) {
    uint8_t ch = '#';
    if (i >= 32 && i < 32 + kMaxFamilies * kNameStride) {
      const uint32_t f = (i - 32) / kNameStride, o = (i - 32) % kNameStride;
      if (o >= 16 && o < 16 + 64 && f < R.n_families) ch = R.fam[f].name[o - 16];
    }
    smem[lay.names + i] = ch;
  }
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
  for (;;) {
    const uint32_t t = s_next;
    if (t >= R.n_tiles) return;
    const Tile tile = s_desc;
    const uint32_t l0 = s_l0;
    const FamMeta& fm = R.fam[tile.family];
    const uint32_t name_len = fm.name_len;
    const uint32_t name_off = names0 + tile.family * kNameStride + 16;  // inside smem[]
    const uint32_t name_sm = sm0 + name_off;
    // ---- pass 1: " <value>\n" + metric name (the glue between my labels and the next line's), exact line length
    const uint32_t slot_off = lay.slots + tid * lay.slot_bytes;
    const uint32_t slot = sm0 + slot_off;
    uint32_t len = 0, lbl_len = 0, vlen = 0;
    bool live = false;
    if (tid < tile.count) {
      const uint32_t r = tile.first + tid;
      lbl_len = R.ser_len[r];
      const double v = series_value(fm.kind, acc, R, r, live);
      if (live) {
        uint8_t* sp = smem + slot_off;
        sp[0] = ' ';
        vlen = uint32_t(format_value(v, sp + 1));
        sp[1 + vlen] = '\n';
        len = name_len + lbl_len + vlen + 2;
        // the name follows at byte o: up to 3 single bytes reach a word boundary, then shifted words
        const uint32_t o = vlen + 2, hb = (4u - (o & 3u)) & 3u;
        for (uint32_t q = 0; q < hb; ++q) sp[o + q] = smem[name_off + q];
        const uint32_t dst = slot + o + hb, s8 = hb * 8u;
        uint32_t w_prev = lds32(name_sm);
        for (uint32_t j = 0; 4u * j + hb < name_len; ++j) {
          const uint32_t w_next = lds32(name_sm + 4u * j + 4u);
          sts32(dst + 4u * j, __funnelshift_r(w_prev, w_next, s8));
          w_prev = w_next;
        }
      }
    }
    const uint32_t hdr = ((tile.flags & 1) && acc.node_live[tile.node] != 0) ? fm.header_len : 0;
    // ---- exclusive scan over the CTA, three fields at once: output bytes (24 bits), label bytes of
    // every series, printed or not (24 bits), lines printed (16 bits)
    const unsigned long long mine = (uint64_t(live) << 48) | (uint64_t(lbl_len) << 24) | len;
    unsigned long long inc = mine;
    for (int o = 1; o < 32; o <<= 1) { unsigned long long u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
    if (lane == 31) s_scan[wid] = inc;
    __syncthreads();  // (1) also: everyone has read s_next / s_desc / s_l0
    unsigned long long woff = 0, all = 0;
    for (int k = 0; k < THREADS / 32; ++k) { if (k < int(wid)) woff += s_scan[k]; all += s_scan[k]; }
    const unsigned long long exc = woff + inc - mine;
    const uint32_t off = hdr + (uint32_t(exc) & 0xffffffu);
    const uint32_t total = hdr + (uint32_t(all) & 0xffffffu);
    const uint32_t lbl_src = sm0 + lay.stage + 32 + l0 + (uint32_t(exc >> 24) & 0xffffffu);  // my label in the staged range
    const uint32_t k_live = uint32_t(exc >> 48), n_live = uint32_t(all >> 48);
    const bool first_live = live && k_live == 0;
    const bool last_live = live && k_live + 1 == n_live;
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
    uint8_t* gtile = R.out + base;            // first byte of the tile in the output
    uint8_t* galn = R.out + (base & ~15ull);  // 16-aligned origin: position v lives at galn + v
    const uint32_t sh = uint32_t(base & 15);
    if (hdr) for (uint32_t i = tid; i < hdr; i += THREADS) gtile[i] = fm.header[i];
    // ---- segment table and chunk -> segment map.  Positions are relative to galn.
    const uint32_t g_first = sh + hdr, end = sh + total;
    const uint32_t c_begin = (g_first + 15u) >> 4, c_end = end >> 4;  // full chunks of the lines: [c_begin, c_end)
    const uint32_t seg_sm = sm0 + lay.seg, loc_sm = sm0 + lay.loc;
    if (live) {
      const uint32_t g = sh + off, b0 = g + name_len, b1 = b0 + lbl_len, b2 = b1 + vlen + 2;
      sts64(seg_sm + (2u * k_live + 1u) * 8u, b0, lbl_src);
      sts64(seg_sm + (2u * k_live + 2u) * 8u, b1, slot);
      const uint32_t c0 = (b0 + 15u) >> 4, c1 = (b1 + 15u) >> 4;
      const uint32_t c2 = last_live ? ((end + 15u) >> 4) : ((b2 + name_len + 15u) >> 4);
      if (first_live) {
        sts64(seg_sm, g, name_sm);
        for (uint32_t c = c_begin; c < c0; ++c) sts16(loc_sm + (c - c_begin) * 2u, 0u);
        for (uint32_t q = g; q < 16u * c_begin && q < end; ++q) galn[q] = smem[name_off + (q - g)];  // before the first full chunk
      }
      if (last_live) sts64(seg_sm + (2u * k_live + 3u) * 8u, 0xffffffffu, 0u);
      for (uint32_t c = c0; c < c1; ++c) sts16(loc_sm + (c - c_begin) * 2u, 2u * k_live + 1u);
      for (uint32_t c = c1; c < c2; ++c) sts16(loc_sm + (c - c_begin) * 2u, 2u * k_live + 2u);
    }
    __syncthreads();         // (2) seg / loc / glue of every line are in place
    mbar_wait(bar, parity);  // staged labels have landed (issued a whole tile ago)
    parity ^= 1;
    // ---- pass 2: consecutive lanes, consecutive 16-byte chunks
    if (n_live) {
      for (uint32_t c = c_begin + tid; c <= c_end; c += THREADS) {
        const uint32_t P = c << 4;
        const uint32_t tail = c == c_end ? (end & 15u) : 16u;  // bytes of this chunk that belong to the tile
        if (tail == 0) break;
        const uint32_t s = lds16(loc_sm + (c - c_begin) * 2u);
        const uint2 sa = lds64(seg_sm + s * 8u), sb = lds64(seg_sm + s * 8u + 8u);
        uint4 v = load16(sa.y + (P - sa.x));
        const uint32_t e = sb.x - P;  // bytes of this chunk before the next segment
        if (e < 16u) v = merge16(v, load16(sb.y - e), e);
        if (tail == 16u) {
          stg128_cs(galn + P, v);
        } else {  // the tile ends inside this chunk: its first bytes are mine, the rest the next tile's
          s_tail = v;
          const uint8_t* sp = reinterpret_cast<const uint8_t*>(&s_tail);
          for (uint32_t q = 0; q < tail; ++q) galn[P + q] = sp[q];
        }
      }
    }
    __syncthreads();             // (3) label stage, glue, seg and loc are free
    if (wid == 0) fetch_next();  // next ticket, descriptor, label copy
    __syncthreads();             // (4) s_next / s_desc / s_l0 visible
  }
}

}  // namespace kgx
