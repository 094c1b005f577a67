// Microbenchmark: can TMA do the writer's byte realignment?  (Answer, B200: no.)
//
// A 1-D tensor map over a byte blob takes an element (= byte) coordinate, so on
// paper a box could start at ANY byte of global memory and still land on an
// aligned shared-memory address; the writer could then fetch each line's label
// text already shifted to the line's alignment in the output, and push finished
// lines out with small bulk stores — no word gathers / funnel shifts / STG in
// the LSU.  Measured on B200 (`tma_realign check <dst skew> <box-1> <x0> <issuers>`):
//   * a box whose first byte is not 16-byte aligned in global memory (x0 = 5, or
//     the second of two issuers at x0 + 101) raises "illegal instruction";
//     x0 = 0, 16, 4096 copy correctly for boxes of 16, 128 and 256 bytes;
//   * a shared-memory destination that is 16- but not 128-byte aligned raises
//     "misaligned address".
// So cp.async.bulk.tensor cannot shift bytes either; the realignment stays in
// registers (word loads + funnel shifts).  The rate part of this program (per-
// line tensor loads / bulk stores, 64 per tile) therefore only runs with
// aligned coordinates and is kept for reference.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_realign tma_realign.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

struct alignas(64) Maps { CUtensorMap m[16]; };  // m[k]: box of 16*(k+1) bytes

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tensor_load_1d(uint32_t dst, const CUtensorMap* map, uint32_t x, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.1d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2}], [%3];" ::"r"(dst),
               "l"(map), "r"(x), "r"(bar)
               : "memory");
}

constexpr int kLines = 64;
constexpr int kSlot = 256;

// mode bit 0: per-line tensor loads; bit 1: per-line bulk stores; bit 2: one bulk load per tile instead
__global__ void __launch_bounds__(kLines) k_lines(const Maps* __restrict__ mapsp, const uint8_t* blob, uint8_t* out, size_t blob_bytes,
                                                  int iters, int mode, uint32_t box_k, uint32_t dst_skew) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) unsigned long long s_bar;
  const uint32_t tid = threadIdx.x;
  const uint32_t bar = uint32_t(__cvta_generic_to_shared(&s_bar));
  const uint32_t slot = uint32_t(__cvta_generic_to_shared(smem)) + tid * kSlot + dst_skew;
  const uint32_t box = 16 * (box_k + 1);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"((mode & 4) ? 1 : kLines));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  __syncthreads();
  uint32_t parity = 0;
  for (int it = 0; it < iters; ++it) {
    const size_t tile = (size_t(blockIdx.x) * iters + it) * 6400 % (blob_bytes - 65536);
    if (mode & 1) {
      const uint32_t x = uint32_t((tile + tid * 100) & ~size_t(15));  // must be 16-byte aligned (see above)
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(box) : "memory");
      tensor_load_1d(slot, &mapsp->m[box_k], x, bar);
      mbar_wait(bar, parity);
      parity ^= 1;
    } else if (mode & 4) {
      if (tid == 0) {
        const uint32_t bytes = 6400 + 128;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         uint32_t(__cvta_generic_to_shared(smem))),
                     "l"(blob + (tile & ~size_t(127))), "r"(bytes), "r"(bar)
                     : "memory");
      }
      mbar_wait(bar, parity);
      parity ^= 1;
    }
    if (mode & 2) {
      // a thread patches a few words of its slot (name / value text), then pushes the line out
      *reinterpret_cast<uint32_t*>(smem + tid * kSlot + 4) = it;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      uint8_t* g = out + ((size_t(blockIdx.x) * iters + it) * kLines + tid) * 144 % (blob_bytes - 65536);
      g = reinterpret_cast<uint8_t*>(reinterpret_cast<uintptr_t>(g) & ~uintptr_t(15));
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(slot - dst_skew), "r"(144) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
  }
  if (mode & 2) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// correctness: one CTA, every thread loads a box at its own odd coordinate; the slots are dumped with plain stores
__global__ void k_check(const Maps* __restrict__ mapsp, uint8_t* dump, uint32_t box_k, uint32_t dst_skew, uint32_t x0, uint32_t issuers) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) unsigned long long s_bar;
  const uint32_t tid = threadIdx.x;
  const uint32_t bar = uint32_t(__cvta_generic_to_shared(&s_bar));
  for (uint32_t i = tid; i < kLines * kSlot + 256; i += blockDim.x) smem[i] = 0xEE;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(issuers));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  __syncthreads();
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  const uint32_t box = 16 * (box_k + 1);
  if (tid < issuers) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(box) : "memory");
    tensor_load_1d(uint32_t(__cvta_generic_to_shared(smem)) + tid * kSlot + dst_skew, &mapsp->m[box_k], x0 + tid * 101, bar);
  }
  mbar_wait(bar, 0);
  __syncthreads();
  for (uint32_t i = tid; i < kLines * kSlot + 256; i += blockDim.x) dump[i] = smem[i];
}

int main(int argc, char** argv) {
  const bool only_check = argc > 3;
  const uint32_t arg_x0 = argc > 4 ? atoi(argv[4]) : 12345, arg_iss = argc > 5 ? atoi(argv[5]) : kLines;
  const uint32_t arg_skew = only_check ? atoi(argv[2]) : 0, arg_bk = only_check ? atoi(argv[3]) : 0;
  const bool only_rate = argc == 2;
  using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  EncodeFn encode = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaFree(0));
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", reinterpret_cast<void**>(&encode), cudaEnableDefault, &q));
  if (!encode) { printf("no cuTensorMapEncodeTiled\n"); return 1; }

  const size_t blob_bytes = size_t(1) << 30;
  uint8_t *blob, *out, *dump;
  CK(cudaMalloc(&blob, blob_bytes)); CK(cudaMalloc(&out, blob_bytes)); CK(cudaMalloc(&dump, kLines * kSlot + 256));
  std::vector<uint8_t> h(1 << 20);
  for (size_t i = 0; i < h.size(); ++i) h[i] = uint8_t((i * 2654435761u) >> 13);
  for (size_t o = 0; o < blob_bytes; o += h.size()) CK(cudaMemcpy(blob + o, h.data(), h.size(), cudaMemcpyHostToDevice));

  Maps maps;
  for (uint32_t k = 0; k < 16; ++k) {
    cuuint64_t dim[1] = {blob_bytes};
    cuuint64_t stride[1] = {0};
    cuuint32_t box[1] = {16 * (k + 1)};
    cuuint32_t estr[1] = {1};
    CUresult r = encode(&maps.m[k], CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, blob, dim, stride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode box %u failed: %d\n", box[0], int(r)); return 1; }
  }
  Maps* d_maps;
  CK(cudaMalloc(&d_maps, sizeof(Maps)));
  CK(cudaMemcpy(d_maps, &maps, sizeof(Maps), cudaMemcpyHostToDevice));
  const int smem = kLines * kSlot + 256;
  CK(cudaFuncSetAttribute(k_lines, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  CK(cudaFuncSetAttribute(k_check, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));

  // ---- correctness of an unaligned source coordinate, for 128- and 16-aligned destinations
  for (uint32_t skew : {arg_skew}) {
    for (uint32_t bk : {arg_bk}) {
      if (only_rate) break;
      const uint32_t x0 = arg_x0;
      k_check<<<1, kLines, smem>>>(d_maps, dump, bk, skew, x0, arg_iss);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("check skew %u box %u: CUDA error %s\n", skew, 16 * (bk + 1), cudaGetErrorString(e)); return 1; }
      std::vector<uint8_t> d(smem);
      CK(cudaMemcpy(d.data(), dump, smem, cudaMemcpyDeviceToHost));
      size_t bad = 0, spill = 0;
      for (int t = 0; t < int(arg_iss); ++t) {
        const uint32_t box = 16 * (bk + 1);
        for (uint32_t i = 0; i < kSlot; ++i) {
          const size_t so = size_t(t) * kSlot + skew + i;
          if (so >= size_t(smem)) continue;
          if (i < box) bad += d[so] != h[(x0 + t * 101 + i) % h.size()];
          else if (i + skew < kSlot) spill += d[so] != 0xEE;
        }
      }
      printf("check: x0 %u issuers %u dst skew %2u box %3u : %zu wrong bytes, %zu bytes written outside the box\n", x0, arg_iss, skew, 16 * (bk + 1), bad, spill);
    }
  }

  if (only_check) return 0;
  // ---- rates
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  int dev_clock_khz = 0;
  CK(cudaDeviceGetAttribute(&dev_clock_khz, cudaDevAttrClockRate, 0));
  const int iters = 400;
  struct Case { int mode; const char* what; };
  const Case cases[] = {{1, "tensor loads per line"}, {2, "bulk stores per line"}, {3, "loads + stores per line"}, {4, "one bulk load per tile"}, {6, "one bulk load + stores per line"}};
  for (const Case& c : cases)
    for (uint32_t bk : {6u, 8u})
      for (int per_sm : {1, 2, 4, 8, 12}) {
        const int grid = 148 * per_sm;
        k_lines<<<grid, kLines, smem>>>(d_maps, blob, out, blob_bytes, 20, c.mode, bk, 0);
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0));
        k_lines<<<grid, kLines, smem>>>(d_maps, blob, out, blob_bytes, iters, c.mode, bk, 0);
        CK(cudaEventRecord(e1));
        CK(cudaDeviceSynchronize());
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        const double tiles_per_sm = double(iters) * per_sm;
        const double us_per_tile = ms * 1e3 / tiles_per_sm;
        const double lines_per_s = double(grid) * iters * kLines / (ms * 1e-3);
        printf("%-32s box %3u  %2d CTA/SM : %7.3f ms  %6.3f us per 64-line tile per SM  %6.1f M lines/s/SM  %6.2f G lines/s chip\n", c.what,
               16 * (bk + 1), per_sm, ms, us_per_tile, lines_per_s / 148 / 1e6, lines_per_s / 1e9);
      }
  return 0;
}
