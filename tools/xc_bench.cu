// Micro-benchmark of the device XC point evaluation (helfem_b200/csrc/xc_point.cuh) at the occupancy the
// grid kernel runs it at: 16 warps per SM, every lane one point per pass.  Reports cycles per warp pass.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 [-DHFB_...] -o tools/xc_bench tools/xc_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../helfem_b200/csrc/xc_point.cuh"

using namespace hfb;

// MODE 0: functional ids at run time; 1: compile-time ids; 2: compile-time ids, two points per lane per pass
template <int FAM, int MODE, int XID, int CID>
__global__ void __launch_bounds__(256, 2) kern(int xid, int cid, int passes, double dens_thr, double* out, long long* cyc) {
  const xc::XcThresholds thr = xc::make_thresholds(dens_thr);
  const int gt = blockIdx.x * blockDim.x + threadIdx.x;
  const int xf = MODE == 0 ? xid : XID, cf = MODE == 0 ? cid : CID;
  double acc = 0.0;
  const long long t0 = clock64();
  for (int it = 0; it < passes; it++) {
    constexpr int NP = MODE == 2 ? 2 : 1;
    double rho[NP], sg[NP], ta[NP], e[NP], vr[NP], vs[NP], vt[NP];
#pragma unroll
    for (int u = 0; u < NP; u++) {
      const int k = (gt * 7 + it * 13 + u * 101) & 1023;
      rho[u] = 1e-4 * exp2(0.02 * k);           // 1e-4 .. ~150
      sg[u] = rho[u] * rho[u] * (0.01 + 1e-3 * (k & 63));
      ta[u] = 0.3 * rho[u] * (1.0 + 0.01 * (k & 31));
      e[u] = vr[u] = vs[u] = vt[u] = 0.0;
    }
#pragma unroll
    for (int u = 0; u < NP; u++) xc::xc_point_unpol<FAM>(xf, thr, rho[u], sg[u], ta[u], e[u], vr[u], vs[u], vt[u]);
#pragma unroll
    for (int u = 0; u < NP; u++) xc::xc_point_unpol<FAM>(cf, thr, rho[u], sg[u], ta[u], e[u], vr[u], vs[u], vt[u]);
#pragma unroll
    for (int u = 0; u < NP; u++) acc += e[u] + vr[u] + vs[u] + vt[u];
  }
  const long long t1 = clock64();
  out[gt] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int FAM, int MODE, int XID, int CID>
void run(const char* name, int passes) {
  const int blocks = 148 * 2;
  double* out; long long* cyc;
  cudaMalloc(&out, sizeof(double) * blocks * 256);
  cudaMalloc(&cyc, sizeof(long long) * blocks);
  // 100 KB of dynamic shared memory per CTA pins the residency at 2 CTAs (16 warps) per SM, as in the grid kernel
  cudaFuncSetAttribute(kern<FAM, MODE, XID, CID>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  kern<FAM, MODE, XID, CID><<<blocks, 256, 100 * 1024>>>(XID, CID, 2, 1e-12, out, cyc);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  kern<FAM, MODE, XID, CID><<<blocks, 256, 100 * 1024>>>(XID, CID, passes, 1e-12, out, cyc);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[296];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double mean = 0; for (int i = 0; i < blocks; i++) mean += (double)h[i] / blocks;
  const int np = MODE == 2 ? 2 : 1;
  cudaFuncAttributes fa;
  cudaFuncGetAttributes(&fa, kern<FAM, MODE, XID, CID>);
  printf("%-28s regs %3d  cycles/warp-pass %8.0f  cycles/point-slot %8.0f  Gpoints/s %6.2f  err=%s\n", name, fa.numRegs,
         mean / passes, mean / passes / np, (double)blocks * 256 * passes * np / (ms * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  const int passes = 200;
  run<xc::FAM_LDA, 0, 1, 12>("lda   runtime ids", passes);
  run<xc::FAM_LDA, 1, 1, 12>("lda   fixed ids", passes);
  run<xc::FAM_GGA, 0, 101, 130>("pbe   runtime ids", passes);
  run<xc::FAM_GGA, 1, 101, 130>("pbe   fixed ids", passes);
  run<xc::FAM_GGA, 2, 101, 130>("pbe   fixed ids, 2 pts", passes);
  run<xc::FAM_MGGA, 0, 497, 498>("r2scan runtime ids", passes);
  run<xc::FAM_MGGA, 1, 497, 498>("r2scan fixed ids", passes);
  run<xc::FAM_MGGA, 2, 497, 498>("r2scan fixed ids, 2 pts", passes);
  return 0;
}
