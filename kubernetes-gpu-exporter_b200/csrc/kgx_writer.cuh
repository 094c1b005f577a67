// kgx_writer.cuh — building blocks of the exposition writer (sm_100a): shared-memory layout, the word gather /
// funnel-shift / 16-byte streaming-store helpers, and the phase-clock instrumentation of experiment builds.  The
// kernel itself is k_render5 in kgx_writer5.cuh.
//
// How a tile of <= THREADS series of one family of one node becomes text (rules O1-O3, G1-G2, F1-F5 of SURVEY.md 8c;
// replaces what promhttp serves at main.go:69-71 -> Registry.Gather + expfmt):
//   * the label text of a tile is ONE contiguous byte range of the series label blob: a single TMA bulk copy
//     (cp.async.bulk global->shared, completion on an mbarrier) stages it;
//   * each thread owns one line and formats its value once, straight into a shared-memory slot (" <value>\n");
//   * line lengths -> CTA scan; the tile's size is published and every warp resolves the tile's global byte offset
//     itself by decoupled look-back over 32 predecessors per step.  Tickets are handed out in output order, so a tile
//     never waits on one that has not started;
//   * each thread then streams its line to the output as 16-byte chunks aligned to the GLOBAL address: a chunk belongs
//     to the line that holds its first byte, so no two lanes ever write the same chunk and every store is a full
//     chunk — or two: chunks that fill one 32-byte sector leave together as STG.256 (a 16-byte store is half a sector
//     on its way to L2).  The four sources of a line (metric name, staged labels, value slot, head of the next line's
//     name) are appended one after the other, all lanes of a warp in the same source at the same time; word loads +
//     funnel shifts realign, computed byte masks merge the one chunk that straddles two sources.  Only the head of a
//     tile's first line and the tail of its last one are partial chunks: those few bytes are byte stores.
// Needs metric names of 15..64 bytes (the tail of a chunk is completed with the next line's name); other schemas take
// the generic k_render (kgx_kernels.cuh).
#pragma once
#include "kgx_async.cuh"
#include "kgx_kernels.cuh"

namespace kgx {

struct RenderLayout {  // byte offsets inside dynamic shared memory (all 16-aligned)
  uint32_t names;      // kMaxFamilies x 96: [16 pad][name (<=64)][16 pad]
  uint32_t slots;      // 2 x THREADS x 28: " <value>\n" of the tile being emitted and of the one being measured
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
// The chunk `hi` at `p` (16-byte aligned): if `pair`, stored together with `lo` as the 32 bytes at p - 16 (32-byte
// aligned; sm_100: STG.256); if `single`, stored alone; else nothing.  Predicated, no branch.
__device__ __forceinline__ void st_pair_or_single(uint8_t* p, const uint4& lo, const uint4& hi, bool pair, bool single) {
  asm volatile(
      "{\n\t.reg .pred q, s;\n\t.reg .u64 a;\n\t"
      "setp.ne.u32 q, %9, 0;\n\t"
      "setp.ne.u32 s, %10, 0;\n\t"
      "sub.u64 a, %0, 16;\n\t"
      "@q st.global.cs.v8.b32 [a], {%1, %2, %3, %4, %5, %6, %7, %8};\n\t"
      "@s st.global.cs.v4.b32 [%0], {%5, %6, %7, %8};\n\t}" ::"l"(p),
      "r"(lo.x), "r"(lo.y), "r"(lo.z), "r"(lo.w), "r"(hi.x), "r"(hi.y), "r"(hi.z), "r"(hi.w), "r"(uint32_t(pair)), "r"(uint32_t(single))
      : "memory");
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
#define KGX_PC_COUNT(i) do { if (tid == 0) pc_acc[i] += 1; } while (0)
#define KGX_PC_FLUSH do { if (tid == 0) { for (int i_ = 0; i_ < 10; ++i_) atomicAdd(&g_phase_clk[i_], pc_acc[i_]); atomicAdd(&g_phase_clk[10], pc_tiles); } } while (0)
#else
#define KGX_PC_DECL
#define KGX_PC(i)
#define KGX_PC_COUNT(i)
#define KGX_PC_FLUSH
#endif

}  // namespace kgx
