// Microbenchmark: do 64-bit reductions (RED, no return) that hit the SAME 32-byte sector from adjacent lanes cost one
// memory request or four?  The join issues six reductions per sample into six SoA arrays (six scattered requests).
// With the per-series state laid out AoS — {count, sum_used, sum_sm, sum_mem} in one 32-byte sector, {max_used,
// last_ts} in the next 16 bytes — four lanes can reduce the four add-fields of ONE sample with one instruction.
//   mode 0: SoA, one lane per sample, K reductions per lane (K instructions, K x 32 scattered words per warp)
//   mode 1: AoS, four lanes per sample, one instruction per 8 samples: lanes 4s..4s+3 hit one sector
//   mode 2: AoS, one lane per sample, four instructions (same sectors as mode 1, but lane-serial)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o red_sector red_sector.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

// every "sample" picks a series inside a window of W series that follows the CTA (as the node-grouped join does)
__device__ __forceinline__ uint32_t series_of(uint64_t i, uint32_t M, uint32_t W) {
  const uint32_t win = uint32_t((i / 512) % (M / W));
  return win * W + mix(uint32_t(i)) % W;
}

template <int MODE, int FIELDS>
__global__ void __launch_bounds__(512) k(unsigned long long* a, uint32_t M, uint32_t W, uint64_t n) {
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint32_t lane = threadIdx.x & 31;
  for (uint64_t i0 = blockIdx.x * uint64_t(blockDim.x) + (threadIdx.x & ~31u); i0 < n; i0 += stride) {
    if (MODE == 0) {
      const uint32_t s = series_of(i0 + lane, M, W);
#pragma unroll
      for (int f = 0; f < FIELDS; ++f) atomicAdd(a + size_t(f) * M + s, (unsigned long long)(lane + f));
    } else if (MODE == 2) {
      const uint32_t s = series_of(i0 + lane, M, W);
#pragma unroll
      for (int f = 0; f < FIELDS; ++f) atomicAdd(a + size_t(s) * 8 + f, (unsigned long long)(lane + f));
    } else {
      constexpr int PER = 32 / FIELDS;  // samples per instruction
#pragma unroll
      for (int g = 0; g < FIELDS; ++g) {
        const uint32_t s = series_of(i0 + g * PER + lane / FIELDS, M, W);
        atomicAdd(a + size_t(s) * 8 + (lane % FIELDS), (unsigned long long)(lane + g));
      }
    }
  }
}

// mode 3: the same four sums accumulated in SHARED memory (a window of 1 024 series x 4 words per CTA, as a join that
// kept a node's state on chip would) with 64-bit or 32-bit atomicAdd, lane per sample; the flush is not counted
template <typename T>
__global__ void __launch_bounds__(512) k_smem(T* out, uint64_t n) {
  __shared__ T s[1024 * 4];
  for (uint32_t i = threadIdx.x; i < 1024 * 4; i += blockDim.x) s[i] = 0;
  __syncthreads();
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n; i += stride) {
    const uint32_t r = mix(uint32_t(i)) % 1024;
#pragma unroll
    for (int f = 0; f < 4; ++f) atomicAdd(&s[r * 4 + f], T(f + 1));
  }
  __syncthreads();
  if (threadIdx.x == 0) out[blockIdx.x] = s[0];
}

template <typename T>
void run_smem(void* a, uint64_t n, const char* what) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    k_smem<T><<<148 * 2, 512>>>(static_cast<T*>(a), n);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  printf("%-64s %.3f ms per 10 M samples (4 fields) = %.1f G field-reductions/s\n", what, best, 4 * double(n) / best / 1e6);
}

template <int MODE, int FIELDS>
void run(unsigned long long* a, uint32_t M, uint32_t W, uint64_t n, const char* what) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    k<MODE, FIELDS><<<148 * 4, 512>>>(a, M, W, n);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  printf("%-64s %.3f ms per 10 M samples (%d fields) = %.1f G field-reductions/s\n", what, best, FIELDS, FIELDS * double(n) / best / 1e6);
}

int main() {
  const uint32_t M = 4u << 20;
  unsigned long long* a; cudaMalloc(&a, size_t(8) * M * 8); cudaMemset(a, 0, size_t(8) * M * 8);
  const uint64_t n = 10000000 / 32 * 32;
  run<0, 4>(a, M, 2048, n, "SoA, lane per sample, 4 instructions");
  run<2, 4>(a, M, 2048, n, "AoS, lane per sample, 4 instructions (same sector, lane-serial)");
  run<1, 4>(a, M, 2048, n, "AoS, 4 lanes per sample -> one sector, 4 instr per 32 samples");
  run<0, 2>(a, M, 2048, n, "SoA, lane per sample, 2 instructions");
  run<1, 2>(a, M, 2048, n, "AoS, 2 lanes per sample -> 16 bytes, 2 instr per 32 samples");
  run<0, 6>(a, M, 2048, n, "SoA, lane per sample, 6 instructions (the join today)");
  run<1, 8>(a, M, 2048, n, "AoS, 8 lanes per sample -> 64 bytes, 8 instr per 32 samples");
  run_smem<unsigned long long>(a, n, "shared memory, 64-bit atomicAdd, lane per sample, 4 instructions");
  run_smem<unsigned int>(a, n, "shared memory, 32-bit atomicAdd, lane per sample, 4 instructions");
  return 0;
}
