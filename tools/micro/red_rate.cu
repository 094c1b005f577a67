// Microbenchmark: throughput of 64-bit reductions (RED.ADD / RED.MAX, no return value) into an L2-resident array —
// the operation that bounds the join (six per sample).  Addresses: (a) uniformly random over M words, (b) as in the
// join: a CTA works on one "node" at a time, i.e. random inside a window of W consecutive words.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o red_rate red_rate.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

template <int OPS, bool MAX>
__global__ void __launch_bounds__(512) k_red(unsigned long long* a, uint32_t M, uint32_t W, uint64_t n) {
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n; i += stride) {
    uint32_t idx;
    if (W) {  // windowed: the window follows the CTA's progress
      const uint32_t win = uint32_t((i / blockDim.x) % (M / W));
      idx = win * W + mix(uint32_t(i)) % W;
    } else {
      idx = mix(uint32_t(i)) % M;
    }
#pragma unroll
    for (int k = 0; k < OPS; ++k) {
      unsigned long long* p = a + size_t(k) * M + idx;
      if (MAX) atomicMax(p, (unsigned long long)i); else atomicAdd(p, (unsigned long long)(i & 1023));
    }
  }
}

template <int OPS, bool MAX>
void run(unsigned long long* a, uint32_t M, uint32_t W, uint64_t n, const char* what) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    k_red<OPS, MAX><<<148 * 4, 512>>>(a, M, W, n);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  printf("%-40s %d x %llu ops: %.3f ms = %.1f G red/s\n", what, OPS, (unsigned long long)n, best, OPS * double(n) / best / 1e6);
}

int main() {
  const uint32_t M = 4u << 20;  // 4 M series
  unsigned long long* a; cudaMalloc(&a, size_t(6) * M * 8); cudaMemset(a, 0, size_t(6) * M * 8);
  const uint64_t n = 10000000;
  run<1, false>(a, M, 0, n, "add, 1 array, random over 4 M");
  run<6, false>(a, M, 0, n, "add, 6 arrays, random over 4 M");
  run<1, false>(a, M, 2048, n, "add, 1 array, window of 2048");
  run<6, false>(a, M, 2048, n, "add, 6 arrays, window of 2048");
  run<6, true>(a, M, 2048, n, "max, 6 arrays, window of 2048");
  run<3, false>(a, M, 2048, n, "add, 3 arrays, window of 2048");
  return 0;
}
