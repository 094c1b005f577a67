// Microbenchmark: latency of a 1-D TMA bulk copy global->shared (and back) as a
// function of size and of source/destination alignment.  One CTA per SM.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}

__global__ void k(const uint8_t* src, uint8_t* dst, uint32_t bytes, uint32_t src_off, uint32_t dst_off, int iters,
                  size_t stride, long long* out_ld, long long* out_st) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) unsigned long long bar;
  const uint32_t b = uint32_t(__cvta_generic_to_shared(&bar));
  const uint32_t s = uint32_t(__cvta_generic_to_shared(smem)) + dst_off;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b));
    asm volatile("fence.mbarrier_init.release.cluster;");
    long long tl = 0, ts = 0;
    uint32_t parity = 0;
    for (int i = 0; i < iters; ++i) {
      const uint8_t* g = src + (size_t(blockIdx.x) * iters + i) * stride + src_off;
      long long t0 = clock64();
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s), "l"(g), "r"(bytes), "r"(b) : "memory");
      mbar_wait(b, parity);
      parity ^= 1;
      long long t1 = clock64();
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      uint8_t* gd = dst + (size_t(blockIdx.x) * iters + i) * stride + src_off;
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gd), "r"(s), "r"(bytes) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      long long t2 = clock64();
      tl += t1 - t0; ts += t2 - t1;
    }
    out_ld[blockIdx.x] = tl / iters; out_st[blockIdx.x] = ts / iters;
  }
}

int main() {
  const int nb = 148, iters = 64;
  const size_t stride = 64 * 1024;
  uint8_t *src, *dst; long long *ol, *os;
  cudaMalloc(&src, size_t(nb) * iters * stride + 4096); cudaMalloc(&dst, size_t(nb) * iters * stride + 4096);
  cudaMemset(src, 1, size_t(nb) * iters * stride + 4096);
  cudaMalloc(&ol, nb * 8); cudaMalloc(&os, nb * 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const uint32_t sizes[] = {1024, 4096, 16384, 26624, 40960};
  const uint32_t offs[][2] = {{0, 0}, {16, 16}, {0, 32}, {48, 32}, {64, 0}};
  for (int pass = 0; pass < 2; ++pass)
    for (uint32_t sz : sizes)
      for (auto& o : offs) {
        k<<<nb, 32, 100 * 1024>>>(src, dst, sz, o[0], o[1], iters, stride, ol, os);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        long long hl[nb], hs[nb];
        cudaMemcpy(hl, ol, nb * 8, cudaMemcpyDeviceToHost); cudaMemcpy(hs, os, nb * 8, cudaMemcpyDeviceToHost);
        long long sl = 0, ss = 0; for (int i = 0; i < nb; ++i) { sl += hl[i]; ss += hs[i]; }
        if (pass) printf("bytes %6u src_off %2u dst_off %2u : load %6lld cyc (%.1f B/cyc)  store(read-done) %6lld cyc\n", sz, o[0], o[1], sl / nb, double(sz) / (sl / nb), ss / nb);
      }
  return 0;
}
