// Micro-benchmark: throughput of 1-D bulk copies (cp.async.bulk global -> shared) issued by ONE thread per CTA with D copies in flight.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulk_copy_bench bulk_copy_bench.cu && ./bulk_copy_bench
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
  asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(a), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src), "r"(bytes), "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
// D stages in flight, PER copies of `bytes` per stage; sources scattered with `stride` bytes
__global__ void k(const char* src, size_t span, size_t stride, unsigned bytes, int D, int PER, int nstage, long long* cycles) {
  extern __shared__ __align__(128) char sm[];
  __shared__ unsigned long long bars[32];
  if (threadIdx.x == 0) {
    for (int i = 0; i < D; i++) mbar_init(bars + i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  size_t off = ((size_t)blockIdx.x * 7919 * stride) & (span - 1);
  const long long t0 = clock64();
  for (int s = 0; s < nstage + D; s++) {
    const int slot = s % D;
    if (s >= D) mbar_wait(bars + slot, ((s / D) - 1) & 1);
    if (s < nstage) {
      mbar_expect_tx(bars + slot, bytes * PER);
      for (int q = 0; q < PER; q++) {
        bulk_g2s(sm + ((size_t)slot * PER + q) * bytes, src + off, bytes, bars + slot);
        off = (off + stride) & (span - 1);
      }
    }
  }
  cycles[blockIdx.x] = clock64() - t0;
}
int main() {
  const size_t span = (size_t)8 << 30;
  char* src; cudaMalloc(&src, span + (1 << 20)); cudaMemset(src, 1, span);
  long long* cyc; cudaMallocManaged(&cyc, 1024 * sizeof(long long));
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int nstage = 400;
  printf("grid bytes PER D : us/stage  GB/s per CTA   (stride 92.8 KB, 8 GB span: DRAM)\n");
  for (int grid : {1, 148, 296})
    for (unsigned bytes : {1856u, 7424u})
      for (int PER : {1, 4})
        for (int D : {1, 2, 5, 10}) {
          if ((size_t)D * PER * bytes > 190 * 1024) continue;
          k<<<grid, 32, (size_t)D * PER * bytes>>>(src, span, 92800 - 92800 % 64 + 64, bytes, D, PER, nstage, cyc);
          if (cudaDeviceSynchronize() != cudaSuccess) { printf("error %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
          double mx = 0; for (int i = 0; i < grid; i++) mx = cyc[i] > mx ? cyc[i] : mx;
          const double us = mx / 1.965e3 / nstage;
          printf("%4d %5u %d %2d : %7.3f  %7.2f\n", grid, bytes, PER, D, us, bytes * PER / us / 1e3);
        }
  return 0;
}
