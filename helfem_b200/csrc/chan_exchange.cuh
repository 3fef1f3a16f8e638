// sm_100a kernels for the Gaunt-coupled exchange operator of the `atomic` and `diatomic` programs
// (first correct version: one CTA per (element pair, output angular pair), direct coupling sums).
//
//   atomic    /root/reference/src/atomic/TwoDBasis.cpp:957-1140
//   diatomic  /root/reference/src/diatomic/basis.cpp:1532-1733
// For the output block (jang,kang) and every integral family r the reference accumulates
//   R_r^{ab} = fac_r sum_{(iang,lang): m_j-m_i = m_k-m_l} c_r^a(j,i) c_r^b(k,l) P_{il}       (:1023-1059 / :1604-1650)
// and contracts it per element pair: same element with the K-ordered in-element integrals,
// different elements with the two disjoint factors (:1063-1134 / :1653-1726).  a, b index the
// stacked families (atomic: one; diatomic: 0 = cos^2-modified coupling, 2 = plain coupling); signs
// and fac_r are folded into the coupling coefficients on the host.  The result is NEGATIVE (-K).
#pragma once
#include <cuda_runtime.h>

namespace hfb {

struct ChanXParams {
  int nel, np, np2, stack, nrad, nang;
  const int* first;      // [nel]
  const int* npe;        // [nel]
  const int* elem_of;    // [nrad][2]
  const double* ktei;    // [nr][nel][stack*stack][np2*np2]
  const double* small_f; // [nr][nel][stack][np2]
  const double* big_f;   // [nr][nel][stack][np2]
  // CSR over output pairs (j*nang + k): entries sorted by family
  const int* xptr;       // [nout + 1]  CSR over the listed output pairs
  const int* xin;        // [nnz] input pair i*nang + l
  const int* xr;         // [nnz] family
  const double* xcoef;   // [nnz][4] coefficient (a,b) at 2a+b
  const double* P;       // dense, ld = ldp
  int ldp;
  const int* out_list;   // [nout] output pairs with at least one coupling
  double* blocks;        // [nout][nel][nel][np2]
};

// grid (nel*nel, nout), 256 threads; dynamic smem (stack^2 + 2) * np2 doubles
static __global__ void __launch_bounds__(256) chan_exchange_block_kernel(ChanXParams prm) {
  extern __shared__ double sm[];
  const int np = prm.np, np2 = prm.np2, st = prm.stack, st2 = st * st;
  double* sR = sm;               // st2 * np2
  double* sT = sR + st2 * np2;   // np2
  double* sK = sT + np2;         // np2
  const int iel = blockIdx.x / prm.nel, jel = blockIdx.x % prm.nel;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int fi = prm.first[iel], fj = prm.first[jel], ni = prm.npe[iel], nj = prm.npe[jel];
  for (int k = tid; k < np2; k += nthr) sK[k] = 0.0;
  int e = prm.xptr[blockIdx.y];            // CSR indexed by position in out_list
  const int eend = prm.xptr[blockIdx.y + 1];
  while (e < eend) {
    const int r = prm.xr[e];
    int e1 = e;
    while (e1 < eend && prm.xr[e1] == r) e1++;
    __syncthreads();
    // coupled density blocks of this family
    for (int k = tid; k < np2; k += nthr) {
      const int i = k % np, j = k / np;
      double v[4] = {0.0, 0.0, 0.0, 0.0};
      if (i < ni && j < nj) {
        for (int q = e; q < e1; q++) {
          const int ip = prm.xin[q];
          const int ia = ip / prm.nang, la = ip % prm.nang;
          const double p = prm.P[(size_t)(ia * prm.nrad + fi + i) + (size_t)(la * prm.nrad + fj + j) * prm.ldp];
#pragma unroll
          for (int ab = 0; ab < 4; ab++)
            if (ab < st2) v[ab] = fma(prm.xcoef[4 * q + ab], p, v[ab]);
        }
      }
      for (int ab = 0; ab < st2; ab++) sR[ab * np2 + k] = v[ab];
    }
    __syncthreads();
    for (int a = 0; a < st; a++)
      for (int b = 0; b < st; b++) {
        const double* R = sR + (a * st + b) * np2;
        if (iel == jel) {
          const double* T = prm.ktei + ((((size_t)r * prm.nel + iel) * st2) + a * st + b) * np2 * np2;
          for (int row = tid; row < np2; row += nthr) {
            double a0 = 0.0, a1 = 0.0;
            int c = 0;
            for (; c + 1 < np2; c += 2) {
              a0 = fma(T[(size_t)c * np2 + row], R[c], a0);
              a1 = fma(T[(size_t)(c + 1) * np2 + row], R[c + 1], a1);
            }
            for (; c < np2; c++) a0 = fma(T[(size_t)c * np2 + row], R[c], a0);
            sK[row] += a0 + a1;
          }
        } else {
          const double* iint = ((iel > jel) ? prm.big_f : prm.small_f) + (((size_t)r * prm.nel + iel) * st + a) * np2;
          const double* jint = ((iel > jel) ? prm.small_f : prm.big_f) + (((size_t)r * prm.nel + jel) * st + b) * np2;
          __syncthreads();
          for (int k = tid; k < np2; k += nthr) {   // T = R * jint^T
            const int i = k % np, kk = k / np;
            double s = 0.0;
            for (int j = 0; j < np; j++) s = fma(R[i + j * np], jint[kk + j * np], s);
            sT[k] = s;
          }
          __syncthreads();
          for (int k = tid; k < np2; k += nthr) {   // K += iint * T
            const int aa = k % np, kk = k / np;
            double s = 0.0;
            for (int i = 0; i < np; i++) s = fma(iint[aa + i * np], sT[i + kk * np], s);
            sK[k] += s;
          }
        }
      }
    e = e1;
  }
  __syncthreads();
  double* out = prm.blocks + (((size_t)blockIdx.y * prm.nel + iel) * prm.nel + jel) * np2;
  for (int k = tid; k < np2; k += nthr) out[k] = -sK[k];
}

// Dense assembly of the listed output pairs (fixed order, no atomics): K(j*nrad + a, k*nrad + b).
static __global__ void chan_exchange_assemble_kernel(ChanXParams prm, int nout, double* K, int ldk) {
  const int np = prm.np, np2 = prm.np2;
  const size_t per = (size_t)prm.nrad * prm.nrad;
  const size_t total = per * nout;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int o = (int)(idx / per);
    const int a = (int)(idx % prm.nrad), b = (int)((idx / prm.nrad) % prm.nrad);
    const int op = prm.out_list[o];
    const int ja = op / prm.nang, ka = op % prm.nang;
    double s = 0.0;
    for (int x = 0; x < 2; x++) {
      const int ie = prm.elem_of[2 * a + x];
      if (ie < 0) continue;
      for (int y = 0; y < 2; y++) {
        const int je = prm.elem_of[2 * b + y];
        if (je < 0) continue;
        s += prm.blocks[(((size_t)o * prm.nel + ie) * prm.nel + je) * np2 + (a - prm.first[ie]) + (b - prm.first[je]) * np];
      }
    }
    K[(size_t)(ja * prm.nrad + a) + (size_t)(ka * prm.nrad + b) * ldk] = s;
  }
}

}  // namespace hfb
