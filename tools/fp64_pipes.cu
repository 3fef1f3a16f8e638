// Micro-benchmark: FP64 throughput of the DFMA pipe vs the DMMA (mma.sync f64) shapes on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_pipes fp64_pipes.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma1688(double* c, const double* a, const double* b) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
__device__ __forceinline__ void dmma16816(double* c, const double* a, const double* b) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}

template <int MODE>
__global__ void __launch_bounds__(256) kern(double* out, int iters, double seed) {
  double acc[16];
#pragma unroll
  for (int i = 0; i < 16; i++) acc[i] = seed * (threadIdx.x + i);
  double a[8], b[4];
#pragma unroll
  for (int i = 0; i < 8; i++) a[i] = 1.0 + 1e-9 * (threadIdx.x + i);
#pragma unroll
  for (int i = 0; i < 4; i++) b[i] = 1.0 - 1e-9 * (threadIdx.x + i);
  for (int it = 0; it < iters; it++) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 16; i++) acc[i] = fma(acc[i], a[i & 7], b[i & 3]);
    } else if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 8; i++) dmma884(acc[2 * i], acc[2 * i + 1], a[i], b[i & 3]);
    } else if (MODE == 2) {
#pragma unroll
      for (int i = 0; i < 4; i++) dmma1688(acc + 4 * i, a + (i & 1) * 4, b + (i & 1) * 2);
    } else if (MODE == 3) {
#pragma unroll
      for (int i = 0; i < 4; i++) dmma16816(acc + 4 * i, a, b);
    } else if (MODE == 4) {  // mix: 8 DFMA + 4 DMMA.8x8x4 per iteration (do the two share one pipe?)
#pragma unroll
      for (int i = 0; i < 4; i++) {
        acc[8 + 2 * i] = fma(acc[8 + 2 * i], a[i], b[i]);
        dmma884(acc[2 * i], acc[2 * i + 1], a[i + 4], b[i]);
        acc[9 + 2 * i] = fma(acc[9 + 2 * i], a[i], b[i]);
      }
    } else if (MODE == 5) {  // one dependent DFMA chain per thread (latency)
#pragma unroll
      for (int i = 0; i < 16; i++) acc[0] = fma(acc[0], a[i & 7], b[i & 3]);
    } else {                 // one dependent DMMA chain per warp (latency)
#pragma unroll
      for (int i = 0; i < 16; i++) dmma884(acc[0], acc[1], a[i & 7], b[i & 3]);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
double run(int blocks, int iters, double flops_per_thread_iter, int threads = 256, float* ms_out = nullptr) {
  double* out;
  cudaMalloc(&out, sizeof(double) * blocks * 256);
  kern<MODE><<<blocks, threads>>>(out, 10, 1e-3);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  kern<MODE><<<blocks, threads>>>(out, iters, 1e-3);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  cudaFree(out);
  if (ms_out) *ms_out = ms;
  return flops_per_thread_iter * iters * blocks * (double)threads / (ms * 1e-3) / 1e12;
}

int main() {
  const int blocks = 148 * 8, iters = 20000;
  // per thread per iteration: DFMA 16*2 flops; m8n8k4: 8 mma * 512/32; m16n8k8: 4 * 2048/32; m16n8k16: 4 * 4096/32
  printf("{\"dfma_tflops\": %.2f, ", run<0>(blocks, iters, 32.0));
  printf("\"dmma_m8n8k4_tflops\": %.2f, ", run<1>(blocks, iters, 8 * 16.0));
  printf("\"dmma_m16n8k8_tflops\": %.2f, ", run<2>(blocks, iters, 4 * 64.0));
  printf("\"dmma_m16n8k16_tflops\": %.2f, ", run<3>(blocks, iters, 4 * 128.0));
  // mix: if DFMA and DMMA ran on separate pipes this would exceed either peak
  printf("\"mix_8dfma_4dmma_tflops\": %.2f, ", run<4>(blocks, iters, 8 * 2.0 + 4 * 16.0));
  // dependent-issue latencies: one warp per SM sub-partition, one chain, 16 ops per iteration
  int clk_khz = 0;
  cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  float ms = 0;
  run<5>(148, iters, 32.0, 32, &ms);
  printf("\"dfma_dependent_latency_cycles\": %.1f, ", ms * 1e-3 * clk_khz * 1e3 / (16.0 * iters));
  run<6>(148, iters, 32.0, 32, &ms);
  printf("\"dmma884_dependent_latency_cycles\": %.1f, \"clock_khz_assumed\": %d}\n", ms * 1e-3 * clk_khz * 1e3 / (16.0 * iters), clk_khz);
  return 0;
}
