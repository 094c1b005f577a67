// Microbenchmark: cost of scattered streaming stores, as the exposition writer issues them — every
// lane writes its own "line" (144 bytes apart) chunk by chunk — with 16-byte (STG.128) or 32-byte
// (STG.256, sm_100) chunks, against fully coalesced 16-byte stores.  Reports GB/s.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>  // 0: scattered 16 B, 1: scattered 32 B, 2: coalesced 16 B
__global__ void k(uint8_t* out, size_t n_lines, uint32_t line_bytes) {
  const size_t line = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
  if (MODE == 2) {
    const size_t total = n_lines * line_bytes / 16;
    for (size_t i = line; i < total; i += size_t(gridDim.x) * blockDim.x) {
      uint4 v = make_uint4(uint32_t(i), 1, 2, 3);
      asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(out + 16 * i), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    }
    return;
  }
  for (size_t l = line; l < n_lines; l += size_t(gridDim.x) * blockDim.x) {
    uint8_t* p = out + l * line_bytes;
    if (MODE == 0) {
      for (uint32_t o = 0; o + 16 <= line_bytes; o += 16)
        asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p + o), "r"(o), "r"(1), "r"(2), "r"(3) : "memory");
    } else {
      for (uint32_t o = 0; o + 32 <= line_bytes; o += 32)
        asm volatile("st.global.cs.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p + o), "r"(o), "r"(1), "r"(2), "r"(3), "r"(4), "r"(5), "r"(6), "r"(7) : "memory");
    }
  }
}

int main() {
  const uint32_t line_bytes = 160;  // multiple of 32 so both variants write the same bytes
  const size_t n_lines = 20u << 20;
  uint8_t* out;
  cudaMalloc(&out, n_lines * line_bytes + 64);
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  const char* names[3] = {"scattered 16 B/lane", "scattered 32 B/lane", "coalesced 16 B/lane"};
  for (int threads : {64, 256})
    for (int mode = 0; mode < 3; ++mode) {
      const int grid = 148 * (2048 / threads);
      float best = 1e9f;
      for (int it = 0; it < 4; ++it) {
        cudaEventRecord(a);
        if (mode == 0) k<0><<<grid, threads>>>(out, n_lines, line_bytes);
        if (mode == 1) k<1><<<grid, threads>>>(out, n_lines, line_bytes);
        if (mode == 2) k<2><<<grid, threads>>>(out, n_lines, line_bytes);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (it) best = ms < best ? ms : best;
      }
      printf("%-22s threads %3d : %.3f ms  %.0f GB/s\n", names[mode], threads, best, n_lines * double(line_bytes) / best / 1e6);
    }
  return cudaDeviceSynchronize() != cudaSuccess;
}
