// sm_100a kernels for exchange() of the spherically averaged atom.
//
// Reference: /root/reference/src/sadatom/basis.cpp:360-526.  For every output l the reference
// forms P_L = sum_lin (4 pi/(2L+1)) totcoup(lout,lin,L) P[lin] (:420-457, the coupling table is a
// constant of the basis and is uploaded once), then contracts per element pair: in-element
// blocks with the K-ordered integrals (:479-493), different elements with the two disjoint
// factors (:495-515).  Results are NEGATIVE (K -= ...), as the reference returns -K.
#pragma once
#include <cuda_runtime.h>

namespace hfb {

struct SadExchParams {
  const double* P;       // dense [nslices][nrad*ldp]
  int ldp, nrad, nslices, nL, nel, np_max;
  const int* first;      // [nel]
  const int* npe;        // [nel]
  const int* elem_of;    // [nrad][2] elements owning a radial function (-1 = none)
  const double* cpl;     // [nslices(out)][nslices(in)][nL]
  const double* dL;      // [nL][nel][np2]
  const double* dm1L;    // [nL][nel][np2]
  const double* ktei;    // [nL][nel][np2*np2]
  double* blocks;        // [nslices][nel][nel][np2]  per-element-pair contributions
  double* K;             // dense [nslices][nrad*ldk]
  int ldk;
};

// grid (nel*nel, nslices): one CTA per (iel, jel, lout)
__global__ void __launch_bounds__(256) sad_exchange_block_kernel(SadExchParams prm) {
  extern __shared__ double sm[];
  const int np = prm.np_max, np2 = np * np;
  double* sPs = sm;        // np2  coupled density block
  double* sT = sPs + np2;  // np2
  double* sK = sT + np2;   // np2  accumulator
  const int iel = blockIdx.x / prm.nel, jel = blockIdx.x % prm.nel, lout = blockIdx.y;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int fi = prm.first[iel], fj = prm.first[jel], ni = prm.npe[iel], nj = prm.npe[jel];
  for (int k = tid; k < np2; k += nthr) sK[k] = 0.0;
  for (int L = 0; L < prm.nL; L++) {
    // any coupling for this (lout, L)?
    bool any = false;
    for (int lin = 0; lin < prm.nslices; lin++) any = any || (prm.cpl[((size_t)lout * prm.nslices + lin) * prm.nL + L] != 0.0);
    if (!any) continue;  // uniform across the block
    __syncthreads();
    for (int k = tid; k < np2; k += nthr) {
      const int i = k % np, j = k / np;
      double v = 0.0;
      if (i < ni && j < nj)
        for (int lin = 0; lin < prm.nslices; lin++) {
          const double c = prm.cpl[((size_t)lout * prm.nslices + lin) * prm.nL + L];
          if (c != 0.0) v = fma(c, prm.P[(size_t)lin * prm.ldp * prm.nrad + (size_t)(fi + i) + (size_t)(fj + j) * prm.ldp], v);
        }
      sPs[k] = v;
    }
    __syncthreads();
    if (iel == jel) {
      const double* T = prm.ktei + ((size_t)L * prm.nel + iel) * np2 * np2;
      for (int row = tid; row < np2; row += nthr) {
        double a0 = 0.0, a1 = 0.0;
        int c = 0;
        for (; c + 1 < np2; c += 2) {
          a0 = fma(T[(size_t)c * np2 + row], sPs[c], a0);
          a1 = fma(T[(size_t)(c + 1) * np2 + row], sPs[c + 1], a1);
        }
        for (; c < np2; c++) a0 = fma(T[(size_t)c * np2 + row], sPs[c], a0);
        sK[row] += a0 + a1;
      }
    } else {
      const double* iint = ((iel > jel) ? prm.dm1L : prm.dL) + ((size_t)L * prm.nel + iel) * np2;
      const double* jint = ((iel > jel) ? prm.dL : prm.dm1L) + ((size_t)L * prm.nel + jel) * np2;
      // T = Psub * jint^T
      for (int k = tid; k < np2; k += nthr) {
        const int i = k % np, kk = k / np;
        double s = 0.0;
        for (int j = 0; j < np; j++) s = fma(sPs[i + j * np], jint[kk + j * np], s);
        sT[k] = s;
      }
      __syncthreads();
      for (int k = tid; k < np2; k += nthr) {
        const int a = k % np, kk = k / np;
        double s = 0.0;
        for (int i = 0; i < np; i++) s = fma(iint[a + i * np], sT[i + kk * np], s);
        sK[k] += s;
      }
    }
  }
  __syncthreads();
  double* out = prm.blocks + (((size_t)lout * prm.nel + iel) * prm.nel + jel) * np2;
  for (int k = tid; k < np2; k += nthr) out[k] = -sK[k];
}

// Dense assembly: K(a,b,l) = sum over elements containing a and b, fixed order (no atomics).
__global__ void sad_exchange_assemble_kernel(SadExchParams prm) {
  const int np = prm.np_max, np2 = np * np;
  const size_t total = (size_t)prm.nrad * prm.nrad * prm.nslices;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int a = (int)(idx % prm.nrad);
    const int b = (int)((idx / prm.nrad) % prm.nrad);
    const int l = (int)(idx / ((size_t)prm.nrad * prm.nrad));
    double s = 0.0;
    for (int ka = 0; ka < 2; ka++) {
      const int ie = prm.elem_of[2 * a + ka];
      if (ie < 0) continue;
      const int ia = a - prm.first[ie];
      for (int kb = 0; kb < 2; kb++) {
        const int je = prm.elem_of[2 * b + kb];
        if (je < 0) continue;
        const int jb = b - prm.first[je];
        s += prm.blocks[(((size_t)l * prm.nel + ie) * prm.nel + je) * np2 + ia + jb * np];
      }
    }
    prm.K[(size_t)l * prm.ldk * prm.nrad + a + (size_t)b * prm.ldk] = s;
  }
}

}  // namespace hfb
