// Microbenchmark: how fast can a grid draw tickets from ONE global counter?  The writer hands its tiles out this
// way (atomicAdd on a single address, one per 64-line tile).  Each CTA loops: lane 0 draws a ticket, the CTA
// learns it through shared memory and a barrier (as the writer does), optionally reads a 32-byte descriptor
// indexed by the ticket, and goes again.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ticket_rate ticket_rate.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

__global__ void __launch_bounds__(64) k_tickets(unsigned long long* ctr, uint32_t n, const uint4* desc, int with_desc, unsigned long long* sink) {
  __shared__ uint32_t s_t;
  __shared__ uint4 s_d;
  unsigned long long acc = 0;
  for (;;) {
    if (threadIdx.x == 0) {
      const uint32_t t = uint32_t(atomicAdd(ctr, 1ull));
      s_t = t;
      if (with_desc && t < n) { s_d = desc[2 * size_t(t)]; }
    }
    __syncthreads();
    const uint32_t t = s_t;
    if (t >= n) break;
    acc += t + (with_desc ? s_d.x : 0u);
    __syncthreads();
  }
  if (acc == 0xdeadbeefdeadbeefull) *sink = acc;
}

int main() {
  const uint32_t n = 384000;
  unsigned long long *ctr, *sink;
  uint4* desc;
  cudaMalloc(&ctr, 8); cudaMalloc(&sink, 8); cudaMalloc(&desc, size_t(n) * 32);
  cudaMemset(desc, 1, size_t(n) * 32);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int with_desc = 0; with_desc < 2; ++with_desc)
    for (int per_sm : {1, 2, 4, 8, 16, 21, 32}) {
      float best = 1e9f;
      for (int rep = 0; rep < 5; ++rep) {
        cudaMemset(ctr, 0, 8);
        cudaEventRecord(e0);
        k_tickets<<<148 * per_sm, 64>>>(ctr, n, desc, with_desc, sink);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
      }
      printf("%s  %2d CTAs/SM: %u tickets in %.3f ms = %.1f ns per ticket, %.2f us per ticket per CTA\n", with_desc ? "ticket+descriptor" : "ticket only      ",
             per_sm, n, best, best * 1e6 / n, best * 1e3 / (double(n) / (148 * per_sm)));
    }
  return 0;
}
