// sm_100a kernels for the Gaunt-coupled exchange operator of the `atomic` and `diatomic` programs.
//
//   atomic    /root/reference/src/atomic/TwoDBasis.cpp:957-1140
//   diatomic  /root/reference/src/diatomic/basis.cpp:1532-1733
// For the output block (jang,kang) and every integral family r the reference accumulates
//   R_r^{ab} = fac_r sum_{(iang,lang): m_j-m_i = m_k-m_l} c_r^a(j,i) c_r^b(k,l) P_{il}       (:1023-1059 / :1604-1650)
// and contracts it per element pair: same element with the K-ordered in-element integrals,
// different elements with the two disjoint factors (:1063-1134 / :1653-1726).  a, b index the
// stacked families (atomic: one; diatomic: 0 = cos^2-modified coupling, 2 = plain coupling).  The
// result is NEGATIVE (-K).
//
// The reference's quadruple angular loop costs O(#(j,k) #(i,l) #L) block AXPYs.  The coupling
// coefficient of a channel c = (L,M) FACTORISES, c_c(j,i) c_c(k,l), and the radial operator of a
// family commutes with the angular contraction, so the operator is evaluated family by family as
//   x1  W^a_c[j,l]  = sum_i c^a_c(j,i) P[i,l]                        (left angular contraction, sparse rows)
//   x2  R^{ab}[j,k] = fac_r (+-) sum_{c in r} sum_l c^b_c(k,l) W^a_c[j,l]   (right angular contraction), written
//                     per element pair as zero-padded (np x np) sub-blocks
//   x3  Ksub(ie,je) = sum_ab T^{ab}_r(ie,je)[Rsub^{ab}(ie,je)]       FP64 tensor pipe (DMMA.8x8x4):
//         ie == je:  (np^2 x np^2) K-ordered integrals x [np^2 x NI items]  -- a GEMM whose N dimension
//                    batches NI output blocks, integrals pre-swizzled into A-fragment order
//         ie != je:  iint . Rsub . jint^T  as two chained (np x np) DMMA products per item
//   x4  K[j,k] -= the (at most four) sub-blocks that overlap each radial index pair   (fixed order)
// Zero blocks of P are skipped like the reference does (Frobenius norm < 10 eps, atomic/TwoDBasis.cpp:1036):
// x1 consults a block mask, x2 the mask of W blocks x1 produced, and the list of output blocks
// ("items") of every family is derived from the mask on the host once per sparsity pattern.
#pragma once
#include <cuda_runtime.h>
#include "sadatom_grid_mma.cuh"  // dmma()

namespace hfb {

struct ChanXParams {
  int nel, np, np2, stack, nrad, nang;
  int MT, KS, LDK;        // m-tiles / k-steps of the swizzled K-ordered integrals; smem row stride of the item tile
  const int* first;       // [nel]
  const int* npe;         // [nel]
  const int* elem_of;     // [nrad][2]
  const double* teiK;     // [nr][nel][stack*stack][MT][KS][32]   A-fragment order
  const double* small_f;  // [nr][nel][stack][np2]
  const double* big_f;    // [nr][nel][stack][np2]
  // coupling rows: row (c, j) lists (i, c^0_c(j,i), c^1_c(j,i))
  const int* rowptr;      // [nchan*nang + 1]
  const int* row_i;       // [nnz]
  const double* row_coef; // [nnz][2]
  // per call
  const double* P;        // dense, ld = ldp
  int ldp;
  const unsigned char* nzP;  // [nang*nang] block (i,l) at i*nang + l
  double* W;              // [nch_max][stack][nang*nang][nrad*nrad]
  unsigned char* Wmask;   // [nch_max][nang*nang]
  double* Rsub;           // [items][stack*stack][nel*nel][np2]
  double* blocks;         // [items][nel*nel][np2]
  double* K;              // dense, ld = ldk
  int ldk;
};

// Frobenius-norm mask of the angular blocks of P.  grid (nang*nang), 128 threads.
static __global__ void __launch_bounds__(128) chan_x0_mask_kernel(const double* P, int ldp, int nrad, int nang, unsigned char* nz) {
  __shared__ double red[4];
  const int i = blockIdx.x / nang, l = blockIdx.x % nang;
  const double* blk = P + (size_t)i * nrad + (size_t)l * nrad * ldp;
  double s = 0.0;
  for (int k = threadIdx.x; k < nrad * nrad; k += blockDim.x) {
    const double v = blk[(k % nrad) + (size_t)(k / nrad) * ldp];
    s = fma(v, v, s);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    const double t = (red[0] + red[1]) + (red[2] + red[3]);
    nz[blockIdx.x] = (sqrt(t) >= 10.0 * 2.220446049250313e-16) ? 1 : 0;
  }
}

// x1.  grid (rows of the family, l chunks), 256 threads.  rows: (channel, j) with a non-empty coupling row.
static __global__ void __launch_bounds__(256) chan_x1_kernel(ChanXParams prm, const int2* rows, int c0) {
  const int c = rows[blockIdx.x].x, j = rows[blockIdx.x].y, cl = c - c0;
  const int nang = prm.nang, nrad = prm.nrad, nrad2 = nrad * nrad, st = prm.stack;
  const int e0 = prm.rowptr[c * nang + j], e1 = prm.rowptr[c * nang + j + 1];
  const size_t nang2 = (size_t)nang * nang;
  for (int l = blockIdx.y; l < nang; l += gridDim.y) {
    bool any = false;
    for (int e = e0; e < e1; e++) any = any || prm.nzP[prm.row_i[e] * nang + l];
    if (!any) continue;  // block-uniform
    double* W0 = prm.W + (((size_t)cl * st) * nang2 + (size_t)j * nang + l) * nrad2;
    double* W1 = W0 + nang2 * nrad2;
    for (int k = threadIdx.x; k < nrad2; k += blockDim.x) {
      const int A = k % nrad, B = k / nrad;
      const double* src = prm.P + A + (size_t)(l * nrad + B) * prm.ldp;
      double w0 = 0.0, w1 = 0.0;
      for (int e = e0; e < e1; e++) {
        const int i = prm.row_i[e];
        if (!prm.nzP[i * nang + l]) continue;
        const double p = src[(size_t)i * nrad];
        w0 = fma(prm.row_coef[2 * e], p, w0);
        w1 = fma(prm.row_coef[2 * e + 1], p, w1);
      }
      W0[k] = w0;
      if (st == 2) W1[k] = w1;
    }
    if (threadIdx.x == 0) prm.Wmask[(size_t)cl * nang2 + (size_t)j * nang + l] = 1;
  }
}

// x2.  grid (items of the family, element chunks of 256*EPT), 256 threads.
constexpr int CHX_EPT = 8;
static __global__ void __launch_bounds__(256) chan_x2_kernel(ChanXParams prm, const int2* items, int c0, int nch, double fac) {
  const int j = items[blockIdx.x].x, k = items[blockIdx.x].y;
  const int nang = prm.nang, nrad = prm.nrad, nrad2 = nrad * nrad, st = prm.stack, st2 = st * st;
  const size_t nang2 = (size_t)nang * nang;
  const int base = blockIdx.y * (256 * CHX_EPT) + threadIdx.x;
  double acc[CHX_EPT][4];
#pragma unroll
  for (int q = 0; q < CHX_EPT; q++) { acc[q][0] = acc[q][1] = acc[q][2] = acc[q][3] = 0.0; }
  for (int cl = 0; cl < nch; cl++) {
    const int c = c0 + cl;
    const int e0 = prm.rowptr[c * nang + k], e1 = prm.rowptr[c * nang + k + 1];
    if (prm.rowptr[c * nang + j] == prm.rowptr[c * nang + j + 1]) continue;  // j has no row in this channel: no W
    for (int e = e0; e < e1; e++) {
      const int l = prm.row_i[e];
      if (!prm.Wmask[(size_t)cl * nang2 + (size_t)j * nang + l]) continue;
      const double cb0 = prm.row_coef[2 * e], cb1 = prm.row_coef[2 * e + 1];
      const double* W0 = prm.W + (((size_t)cl * st) * nang2 + (size_t)j * nang + l) * nrad2;
      const double* W1 = W0 + nang2 * nrad2;
#pragma unroll
      for (int q = 0; q < CHX_EPT; q++) {
        const int el = base + 256 * q;
        if (el < nrad2) {
          const double w0 = W0[el];
          acc[q][0] = fma(cb0, w0, acc[q][0]);
          if (st == 2) {
            const double w1 = W1[el];
            acc[q][1] = fma(cb1, w0, acc[q][1]);  // (a,b) = (0,1)
            acc[q][2] = fma(cb0, w1, acc[q][2]);  // (1,0)
            acc[q][3] = fma(cb1, w1, acc[q][3]);  // (1,1)
          }
        }
      }
    }
  }
  // scatter into the element-pair sub-blocks (a radial function belongs to one or two elements)
  const int nel = prm.nel, nel2 = nel * nel, np = prm.np, np2 = prm.np2;
  double* out = prm.Rsub + (size_t)blockIdx.x * st2 * nel2 * np2;
#pragma unroll
  for (int q = 0; q < CHX_EPT; q++) {
    const int el = base + 256 * q;
    if (el >= nrad2) continue;
    const int A = el % nrad, B = el / nrad;
    for (int x = 0; x < 2; x++) {
      const int ie = prm.elem_of[2 * A + x];
      if (ie < 0) continue;
      for (int y = 0; y < 2; y++) {
        const int je = prm.elem_of[2 * B + y];
        if (je < 0) continue;
        const size_t off = (size_t)(ie * nel + je) * np2 + (A - prm.first[ie]) + (B - prm.first[je]) * np;
        for (int ab = 0; ab < st2; ab++) {
          const int a = ab / st, b = ab % st;
          out[(size_t)ab * nel2 * np2 + off] = (a != b ? -fac : fac) * acc[q][st == 2 ? ab : 0];
        }
      }
    }
  }
}

// x3.  grid (item groups of NI, nel*nel), 256 threads; dynamic smem NI*LDK doubles.
// NTL = 8-row tiles of an (np x np) block (np <= 8 NTL).
template <int NI, int NTL>
__global__ void __launch_bounds__(256) chan_x3_kernel(ChanXParams prm, int r, int nitems) {
  extern __shared__ double sm[];
  const int nel = prm.nel, nel2 = nel * nel, np = prm.np, np2 = prm.np2, st = prm.stack, st2 = st * st;
  const int ie = blockIdx.y / nel, je = blockIdx.y % nel;
  const int n0 = blockIdx.x * NI, nn = min(NI, nitems - n0);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, t4 = lane & 3;
  const size_t istride = (size_t)st2 * nel2 * np2;  // Rsub stride between items
  const double* Rbase = prm.Rsub + (size_t)n0 * istride + (size_t)(ie * nel + je) * np2;
  double* Obase = prm.blocks + ((size_t)n0 * nel2 + ie * nel + je) * np2;

  if (ie == je) {
    // ---------------- same element: C[np2 x NI] = sum_ab teiK^{ab} [np2 x np2] . Rsub^{ab} [np2 x NI] ----------------
    constexpr int NTN = NI / 8;
    const int MT = prm.MT, KS = prm.KS, LDK = prm.LDK;
    double acc[4][NTN][2];
#pragma unroll
    for (int mi = 0; mi < 4; mi++)
#pragma unroll
      for (int nt = 0; nt < NTN; nt++) { acc[mi][nt][0] = 0.0; acc[mi][nt][1] = 0.0; }
    for (int mt0 = 0; mt0 < MT; mt0 += 32) {  // 32 m-tiles per sweep (4 per warp); np2 <= 256 needs one sweep
      for (int ab = 0; ab < st2; ab++) {
        __syncthreads();
        for (int idx = tid; idx < NI * LDK; idx += 256) {
          const int n = idx / LDK, k = idx - n * LDK;
          sm[idx] = (n < nn && k < np2) ? Rbase[(size_t)n * istride + (size_t)ab * nel2 * np2 + k] : 0.0;
        }
        __syncthreads();
        const double* A = prm.teiK + ((((size_t)r * nel + ie) * st2 + ab) * MT) * KS * 32 + lane;
        const double* bs = sm + gid * LDK + t4;
#pragma unroll 2
        for (int ks = 0; ks < KS; ks++) {
          double b[NTN], a[4];
#pragma unroll
          for (int nt = 0; nt < NTN; nt++) b[nt] = bs[8 * nt * LDK + 4 * ks];
#pragma unroll
          for (int mi = 0; mi < 4; mi++) {
            const int mt = mt0 + warp + 8 * mi;
            a[mi] = (mt < MT) ? __ldg(A + ((size_t)mt * KS + ks) * 32) : 0.0;
          }
#pragma unroll
          for (int mi = 0; mi < 4; mi++)
#pragma unroll
            for (int nt = 0; nt < NTN; nt++) dmma(acc[mi][nt][0], acc[mi][nt][1], a[mi], b[nt]);
        }
      }
#pragma unroll
      for (int mi = 0; mi < 4; mi++) {
        const int row = 8 * (mt0 + warp + 8 * mi) + gid;
#pragma unroll
        for (int nt = 0; nt < NTN; nt++)
#pragma unroll
          for (int cc = 0; cc < 2; cc++) {
            const int n = 8 * nt + 2 * t4 + cc;
            if (row < np2 && n < nn) Obase[(size_t)n * nel2 * np2 + row] = acc[mi][nt][cc];
            acc[mi][nt][cc] = 0.0;
          }
      }
    }
  } else {
    // ---------------- different elements: Ksub = sum_ab iint^a . Rsub^{ab} . (jint^b)^T, one item per warp pass ----------------
    constexpr int LDY = 8 * NTL + 1;
    double* Y = sm + warp * (8 * NTL * LDY);
    const bool lower = ie > je;
    for (int n = warp; n < nn; n += 8) {
      double kacc[NTL][NTL][2];
#pragma unroll
      for (int mt = 0; mt < NTL; mt++)
#pragma unroll
        for (int qt = 0; qt < NTL; qt++) { kacc[mt][qt][0] = 0.0; kacc[mt][qt][1] = 0.0; }
      for (int ab = 0; ab < st2; ab++) {
        const int a = ab / st, b = ab % st;
        const double* iint = (lower ? prm.big_f : prm.small_f) + (((size_t)r * nel + ie) * st + a) * np2;
        const double* jint = (lower ? prm.small_f : prm.big_f) + (((size_t)r * nel + je) * st + b) * np2;
        const double* Rn = Rbase + (size_t)n * istride + (size_t)ab * nel2 * np2;
        // step 1: Y = iint . Rsub
        double y[NTL][NTL][2];
#pragma unroll
        for (int mt = 0; mt < NTL; mt++)
#pragma unroll
          for (int nt = 0; nt < NTL; nt++) { y[mt][nt][0] = 0.0; y[mt][nt][1] = 0.0; }
#pragma unroll
        for (int ks = 0; ks < 2 * NTL; ks++) {
          const int kk = 4 * ks + t4;
          double af[NTL], bf[NTL];
#pragma unroll
          for (int mt = 0; mt < NTL; mt++) af[mt] = (kk < np && 8 * mt + gid < np) ? __ldg(iint + (8 * mt + gid) + kk * np) : 0.0;
#pragma unroll
          for (int nt = 0; nt < NTL; nt++) bf[nt] = (kk < np && 8 * nt + gid < np) ? Rn[kk + (8 * nt + gid) * np] : 0.0;
#pragma unroll
          for (int mt = 0; mt < NTL; mt++)
#pragma unroll
            for (int nt = 0; nt < NTL; nt++) dmma(y[mt][nt][0], y[mt][nt][1], af[mt], bf[nt]);
        }
        __syncwarp();
#pragma unroll
        for (int mt = 0; mt < NTL; mt++)
#pragma unroll
          for (int nt = 0; nt < NTL; nt++) {
            Y[(8 * mt + gid) * LDY + 8 * nt + 2 * t4] = y[mt][nt][0];
            Y[(8 * mt + gid) * LDY + 8 * nt + 2 * t4 + 1] = y[mt][nt][1];
          }
        __syncwarp();
        // step 2: Ksub += Y . jint^T   (A = Y from shared memory, B[k][q] = jint[q][k])
#pragma unroll
        for (int ks = 0; ks < 2 * NTL; ks++) {
          const int kk = 4 * ks + t4;
          double af[NTL], bf[NTL];
#pragma unroll
          for (int mt = 0; mt < NTL; mt++) af[mt] = Y[(8 * mt + gid) * LDY + kk];
#pragma unroll
          for (int qt = 0; qt < NTL; qt++) bf[qt] = (kk < np && 8 * qt + gid < np) ? __ldg(jint + (8 * qt + gid) + kk * np) : 0.0;
#pragma unroll
          for (int mt = 0; mt < NTL; mt++)
#pragma unroll
            for (int qt = 0; qt < NTL; qt++) dmma(kacc[mt][qt][0], kacc[mt][qt][1], af[mt], bf[qt]);
        }
      }
      double* On = Obase + (size_t)n * nel2 * np2;
#pragma unroll
      for (int mt = 0; mt < NTL; mt++)
#pragma unroll
        for (int qt = 0; qt < NTL; qt++)
#pragma unroll
          for (int cc = 0; cc < 2; cc++) {
            const int row = 8 * mt + gid, col = 8 * qt + 2 * t4 + cc;
            if (row < np && col < np) On[row + col * np] = kacc[mt][qt][cc];
          }
    }
  }
}

// x4.  K[j,k] -= overlapping sub-blocks, fixed order.  grid-stride over items x nrad^2.
static __global__ void chan_x4_kernel(ChanXParams prm, const int2* items, int nitems) {
  const int np = prm.np, np2 = prm.np2, nel = prm.nel, nel2 = nel * nel, nrad = prm.nrad;
  const size_t per = (size_t)nrad * nrad, total = per * nitems;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int o = (int)(idx / per);
    const int a = (int)(idx % nrad), b = (int)((idx / nrad) % nrad);
    const int2 jk = items[o];
    double s = 0.0;
    for (int x = 0; x < 2; x++) {
      const int ie = prm.elem_of[2 * a + x];
      if (ie < 0) continue;
      for (int y = 0; y < 2; y++) {
        const int je = prm.elem_of[2 * b + y];
        if (je < 0) continue;
        s += prm.blocks[((size_t)o * nel2 + ie * nel + je) * np2 + (a - prm.first[ie]) + (b - prm.first[je]) * np];
      }
    }
    prm.K[(size_t)(jk.x * nrad + a) + (size_t)(jk.y * nrad + b) * prm.ldk] -= s;
  }
}

}  // namespace hfb
