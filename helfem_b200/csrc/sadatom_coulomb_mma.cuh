// sm_100a kernels for TwoDBasis::coulomb of the spherically averaged atom (sadatom/basis.cpp:296-358,
// L = 0 only, :306) on the FP64 tensor pipe.
//
// Algebra.  For element e the reference forms  Jsub = Lfac * prim_tei_e * vec(Psub)  with
// prim_tei(j*N+i, l*N+k) = (ij|kl), plus the disjoint terms  dL_e * sum_{j>e} jbig_j  and
// dm1L_e * sum_{j<e} jsmall_j  where  jsmall/jbig = Lfac tr(dL/dm1L * Psub)  (:314-352).  (ij|kl) is
// symmetric in i<->j and in k<->l, so the product only sees the symmetric part of Psub and only the
// i<=j rows are distinct: the integrals are stored TRIANGLE-PACKED, Tc[(i<=j),(k<=l)] (ntri = N(N+1)/2
// per side, 3.5x fewer bytes than the reference's N^2 x N^2 block) against the folded density
// pc[(k<=l)] = P_kl + P_lk (k<l), P_kk.  This is exact for ANY input matrix, symmetric or not.
//
//   sad_jpack_kernel   (element, build): pc and the two traces (only when no grid kernel runs in the same build)
//   sad_jmma_kernel    (element, 32-row tile, build chunk): J_c = Tc_e pc as DMMA.8x8x4 with the A
//                      operand pre-swizzled on the host into fragment order (one coalesced 256 B load
//                      per warp per fragment, straight from L2/HBM into registers), then the prefix /
//                      suffix trace sums, the disjoint terms and the scatter to the element band.
//                      SPLITK=false: warps split the builds (batched GEMM, FP64-pipe bound);
//                      SPLITK=true : warps split K and reduce through shared memory (one or a few
//                      builds: a GEMV that streams every stored integral once, HBM bound).
#pragma once
#include <cuda_runtime.h>
#include "sadatom_grid_mma.cuh"

namespace hfb {

struct SadJParams {
  const double* teiF;   // [nel][MT][KS][32]  fragment-ordered Tc: lane -> Tc[8 mt + lane/4][(lane%4) KS + ks]  (KS even:
                        // a lane's B operands of consecutive k-steps are consecutive doubles of pc -> 16-byte loads)
  const double* dLc;    // [nel][MP]  triangle-packed disjoint_L   (L = 0)
  const double* dm1Lc;  // [nel][MP]
  const int2* tri;      // [ntri] (i, j), i <= j, m = i + j(j+1)/2
  int nel, np_max, ntri, MT, KS, nslices, nspin, nb;
  const double* P;      // [nb][nspin][nslices][nel][np2]
  double* pc;           // [nb][nel][4 KS]   scratch
  double* traces;       // [nb][nel][2]      scratch: jsmall, jbig
  double* J;            // [nb][nel][np2]
};

// grid (nel, ceil(nb/4)), 128 threads: one WARP per (element, build).  Stand-alone form of the by-product the grid
// kernel emits when J and Vxc are built together (sadatom_grid_mma.cuh, stage 0) -- same summation order, so J does
// not depend on which operators were requested in the same call: per spin the l slices are summed in ascending
// order for P_ij and for P_ji separately, then added; spins are added in order; traces by lane-strided fma + xor tree.
static __global__ void __launch_bounds__(128) sad_jpack_kernel(SadJParams prm) {
  const int el = blockIdx.x, lane = threadIdx.x & 31;
  const int b = blockIdx.y * 4 + (threadIdx.x >> 5);
  if (b >= prm.nb) return;
  const int np = prm.np_max, np2 = np * np, KP = 4 * prm.KS, MP = 8 * prm.MT;
  const size_t sstride = (size_t)prm.nel * np2;
  double ts = 0.0, tb = 0.0;
  for (int m = lane; m < KP; m += 32) {
    double v = 0.0, dl = 0.0, dm = 0.0;
    if (m < prm.ntri) {
      const int2 ij = prm.tri[m];
      const int a = ij.x + ij.y * np, c = ij.y + ij.x * np;
      for (int s = 0; s < prm.nspin; s++) {
        const double* gP = prm.P + ((size_t)(b * prm.nspin + s) * prm.nslices * prm.nel + el) * np2;
        double sa = 0.0, sc = 0.0;
        for (int l = 0; l < prm.nslices; l++) {
          sa += gP[(size_t)l * sstride + a];
          sc += gP[(size_t)l * sstride + c];
        }
        double vs = sa;
        if (ij.x != ij.y) vs += sc;
        v = (s == 0) ? vs : vs + v;
      }
      dl = prm.dLc[(size_t)el * MP + m];
      dm = prm.dm1Lc[(size_t)el * MP + m];
    }
    ts = fma(dl, v, ts);
    tb = fma(dm, v, tb);
    prm.pc[((size_t)b * prm.nel + el) * KP + m] = v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ts += __shfl_xor_sync(0xffffffffu, ts, o);
    tb += __shfl_xor_sync(0xffffffffu, tb, o);
  }
  if (lane == 0) {
    const double Lfac = 4.0 * HFB_PI;
    prm.traces[((size_t)b * prm.nel + el) * 2 + 0] = Lfac * ts;
    prm.traces[((size_t)b * prm.nel + el) * 2 + 1] = Lfac * tb;
  }
}

// grid (nel * ceil(MT/4), nchunks), 256 threads.  NTW: 8-build tiles per warp pass.
template <bool SPLITK, int NTW>
__global__ void __launch_bounds__(256) sad_jmma_kernel(SadJParams prm) {
  __shared__ double sred[SPLITK ? 8 * 4 * NTW * 64 : 1];
  const int rts = (prm.MT + 3) >> 2;
  const int el = blockIdx.x / rts, rt = blockIdx.x % rts;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gid = lane >> 2, t4 = lane & 3;
  const int KS = prm.KS, KP = 4 * KS, MP = 8 * prm.MT, np = prm.np_max, np2 = np * np;
  const int mt0 = 4 * rt, nmt = min(4, prm.MT - mt0);
  // builds of this CTA / warp
  const int chunk = SPLITK ? 8 * NTW : 8 * 8 * NTW;
  const int bw0 = blockIdx.y * chunk + (SPLITK ? 0 : warp * 8 * NTW);
  if (!SPLITK && bw0 >= prm.nb) return;

  double acc[4][NTW][2];
#pragma unroll
  for (int m = 0; m < 4; m++)
#pragma unroll
    for (int n = 0; n < NTW; n++) { acc[m][n][0] = 0.0; acc[m][n][1] = 0.0; }

  const double* A = prm.teiF + (((size_t)el * prm.MT + mt0) * KS) * 32 + lane;
  // B operand: Bop[k][n] = pc[b = bw0 + 8 n + gid][el][t4 KS + ks]
  const double* Bp[NTW];
  bool bok[NTW];
#pragma unroll
  for (int n = 0; n < NTW; n++) {
    const int bb = bw0 + 8 * n + gid;
    bok[n] = bb < prm.nb;
    Bp[n] = prm.pc + ((size_t)(bok[n] ? bb : 0) * prm.nel + el) * KP + t4 * KS;
  }
  const int ks0 = SPLITK ? warp : 0, kstep = SPLITK ? 8 : 1;
  // software pipeline: fragments of step ks+kstep are in flight while step ks is multiplied
  double a[4], bf[NTW];
  auto load = [&](int ks, double* aa, double* bb) {
#pragma unroll
    for (int m = 0; m < 4; m++) aa[m] = (m < nmt && ks < KS) ? __ldg(A + ((size_t)m * KS + ks) * 32) : 0.0;
#pragma unroll
    for (int n = 0; n < NTW; n++) bb[n] = (bok[n] && ks < KS) ? __ldg(Bp[n] + ks) : 0.0;
  };
  load(ks0, a, bf);
#pragma unroll 4
  for (int ks = ks0; ks < KS; ks += kstep) {
    double an[4], bn[NTW];
    load(ks + kstep, an, bn);
#pragma unroll
    for (int m = 0; m < 4; m++)
#pragma unroll
      for (int n = 0; n < NTW; n++) dmma(acc[m][n][0], acc[m][n][1], a[m], bf[n]);
#pragma unroll
    for (int m = 0; m < 4; m++) a[m] = an[m];
#pragma unroll
    for (int n = 0; n < NTW; n++) bf[n] = bn[n];
  }
  if (SPLITK) {
    // cross-warp reduction in a fixed order (deterministic)
#pragma unroll
    for (int m = 0; m < 4; m++)
#pragma unroll
      for (int n = 0; n < NTW; n++) {
        sred[((warp * 4 + m) * NTW + n) * 64 + lane] = acc[m][n][0];
        sred[((warp * 4 + m) * NTW + n) * 64 + 32 + lane] = acc[m][n][1];
      }
    __syncthreads();
    if (warp != 0) return;
#pragma unroll
    for (int m = 0; m < 4; m++)
#pragma unroll
      for (int n = 0; n < NTW; n++) {
        double s0 = 0.0, s1 = 0.0;
        for (int w = 0; w < 8; w++) {
          s0 += sred[((w * 4 + m) * NTW + n) * 64 + lane];
          s1 += sred[((w * 4 + m) * NTW + n) * 64 + 32 + lane];
        }
        acc[m][n][0] = s0;
        acc[m][n][1] = s1;
      }
  }
  // epilogue: thread holds Jc[m = 8 (mt0+mm) + gid][b = bw0 + 8 n + 2 t4 + c]
  const double Lfac = 4.0 * HFB_PI;
#pragma unroll
  for (int n = 0; n < NTW; n++)
#pragma unroll
    for (int c = 0; c < 2; c++) {
      const int bb = bw0 + 8 * n + 2 * t4 + c;
      if (bb >= prm.nb) continue;
      const double* tr = prm.traces + (size_t)bb * prm.nel * 2;
      double below = 0.0, above = 0.0;  // sum of jsmall over j < el, of jbig over j > el (reference order)
      for (int j = 0; j < el; j++) below += tr[2 * j];
      for (int j = el + 1; j < prm.nel; j++) above += tr[2 * j + 1];
      double* gJ = prm.J + ((size_t)bb * prm.nel + el) * np2;
#pragma unroll
      for (int mm = 0; mm < 4; mm++) {
        const int m = 8 * (mt0 + mm) + gid;
        if (mm < nmt && m < prm.ntri) {
          double v = Lfac * acc[mm][n][c];
          v = fma(above, prm.dLc[(size_t)el * MP + m], v);
          v = fma(below, prm.dm1Lc[(size_t)el * MP + m], v);
          const int2 ij = prm.tri[m];
          gJ[ij.x + ij.y * np] = v;
          gJ[ij.y + ij.x * np] = v;
        }
      }
    }
}

// Batched form (many builds): grid (nel * ceil(MT/4), build chunks of 64 NTW), 256 threads.  The CTA's 4 row tiles of
// Tc for ALL k-steps (KS KB) are staged once in shared memory and shared by its 8 warps, each of which multiplies
// them into its own 8 NTW builds; a lane's B operands stream from pc with 16-byte loads, two k-steps each,
// prefetched one pair ahead.  Same epilogue as sad_jmma_kernel.
template <int NTW>
__global__ void __launch_bounds__(256) sad_jmma_smem_kernel(SadJParams prm) {
  static_assert(NTW <= 2, "epilogue constants are staged by threads [0, 64 NTW) and [224, 256)");
  extern __shared__ double sA[];  // [4][KS][32]
  __shared__ double sBelow[64 * NTW], sAbove[64 * NTW], sdl[32], sdm[32];
  __shared__ int2 stri[32];
  const int rts = (prm.MT + 3) >> 2;
  const int el = blockIdx.x / rts, rt = blockIdx.x % rts;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gid = lane >> 2, t4 = lane & 3;
  const int KS = prm.KS, KP = 4 * KS, MP = 8 * prm.MT, np = prm.np_max, np2 = np * np;
  const int mt0 = 4 * rt, nmt = min(4, prm.MT - mt0);
  const int bw0 = (blockIdx.y * 8 + warp) * 8 * NTW;

  const double2* Bp[NTW];
  bool bok[NTW];
#pragma unroll
  for (int n = 0; n < NTW; n++) {
    const int bb = bw0 + 8 * n + gid;
    bok[n] = bb < prm.nb;
    Bp[n] = reinterpret_cast<const double2*>(prm.pc + ((size_t)(bok[n] ? bb : 0) * prm.nel + el) * KP + t4 * KS);
  }
  const int KS2 = KS >> 1;
  double2 bcur[NTW], bnxt[NTW];
#pragma unroll
  for (int n = 0; n < NTW; n++) bcur[n] = bok[n] ? __ldg(Bp[n]) : make_double2(0.0, 0.0);
  // epilogue constants: the disjoint-term prefix / suffix trace sums of the CTA's builds (reference order: ascending
  // element index), and disjoint_L / disjoint_m1L / (i,j) of the CTA's 32 rows
  if (tid < 64 * NTW) {
    const int bb = blockIdx.y * 64 * NTW + tid;
    double below = 0.0, above = 0.0;
    if (bb < prm.nb) {
      const double* tr = prm.traces + (size_t)bb * prm.nel * 2;
      for (int j = 0; j < el; j++) below += tr[2 * j];
      for (int j = el + 1; j < prm.nel; j++) above += tr[2 * j + 1];
    }
    sBelow[tid] = below;
    sAbove[tid] = above;
  } else if (tid >= 224) {
    const int m = 8 * mt0 + (tid - 224);
    const bool ok = m < prm.ntri;
    sdl[tid - 224] = ok ? prm.dLc[(size_t)el * MP + m] : 0.0;
    sdm[tid - 224] = ok ? prm.dm1Lc[(size_t)el * MP + m] : 0.0;
    stri[tid - 224] = ok ? prm.tri[m] : make_int2(-1, -1);
  }
  {  // stage A: contiguous [nmt][KS][32] doubles
    const double2* gA = reinterpret_cast<const double2*>(prm.teiF + (((size_t)el * prm.MT + mt0) * KS) * 32);
    double2* s2 = reinterpret_cast<double2*>(sA);
    for (int k = tid; k < nmt * KS * 16; k += 256) s2[k] = __ldg(gA + k);
    for (int k = nmt * KS * 16 + tid; k < 4 * KS * 16; k += 256) s2[k] = make_double2(0.0, 0.0);
  }
  __syncthreads();
  if (bw0 >= prm.nb) return;

  double acc[4][NTW][2];
#pragma unroll
  for (int m = 0; m < 4; m++)
#pragma unroll
    for (int n = 0; n < NTW; n++) { acc[m][n][0] = 0.0; acc[m][n][1] = 0.0; }
  const double* sAl = sA + lane;
#pragma unroll 2
  for (int k2 = 0; k2 < KS2; k2++) {
#pragma unroll
    for (int n = 0; n < NTW; n++) bnxt[n] = (bok[n] && k2 + 1 < KS2) ? __ldg(Bp[n] + k2 + 1) : make_double2(0.0, 0.0);
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int ks = 2 * k2 + h;
      double a[4];
#pragma unroll
      for (int m = 0; m < 4; m++) a[m] = sAl[(m * KS + ks) * 32];
#pragma unroll
      for (int m = 0; m < 4; m++)
#pragma unroll
        for (int n = 0; n < NTW; n++) dmma(acc[m][n][0], acc[m][n][1], a[m], h ? bcur[n].y : bcur[n].x);
    }
#pragma unroll
    for (int n = 0; n < NTW; n++) bcur[n] = bnxt[n];
  }
  // epilogue: thread holds Jc[m = 8 (mt0+mm) + gid][b = bw0 + 8 n + 2 t4 + c]
  const double Lfac = 4.0 * HFB_PI;
#pragma unroll
  for (int n = 0; n < NTW; n++)
#pragma unroll
    for (int c = 0; c < 2; c++) {
      const int bl = warp * 8 * NTW + 8 * n + 2 * t4 + c, bb = blockIdx.y * 64 * NTW + bl;
      if (bb >= prm.nb) continue;
      const double below = sBelow[bl], above = sAbove[bl];
      double* gJ = prm.J + ((size_t)bb * prm.nel + el) * np2;
#pragma unroll
      for (int mm = 0; mm < 4; mm++) {
        const int2 ij = stri[8 * mm + gid];
        if (ij.x >= 0) {
          double v = Lfac * acc[mm][n][c];
          v = fma(above, sdl[8 * mm + gid], v);
          v = fma(below, sdm[8 * mm + gid], v);
          gJ[ij.x + ij.y * np] = v;
          gJ[ij.y + ij.x * np] = v;
        }
      }
    }
}

}  // namespace hfb
