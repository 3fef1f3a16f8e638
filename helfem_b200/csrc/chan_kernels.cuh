// sm_100a kernels for the Gaunt-coupled Coulomb operator of the `atomic` and `diatomic` programs.
//
// Both references have the same three-step structure
//   atomic    /root/reference/src/atomic/TwoDBasis.cpp:817-955
//   diatomic  /root/reference/src/diatomic/basis.cpp:1359-1530
//   (i)   ket contraction: per coupling channel c = (L,M) a radial density  Paux_c = sum_{(k,l)} G P_{kl}
//   (ii)  radial Coulomb operator of the channel's integral family r (atomic: r = L; diatomic: r = (L,|M|)):
//         in-element block  fac * T_r * vec(Psub)  + disjoint terms driven by two traces per element
//   (iii) bra contraction:  J_{ij} += G * Jaux_c
// which is re-derived here as three kernels over device-resident constants:
//   chan_pre_kernel    (element, channel, build): CSR gather of the angular-pair band into a
//                      triangle-packed, symmetry-folded density vector + the two traces
//   chan_jmma_kernel   (element x 32-row tile, column chunk, family): DMMA.8x8x4 product with the
//                      fragment-ordered packed integrals; columns = the channels of the family x builds
//   chan_post_kernel   (element, angular pair, build): CSR gather of the channel results into the
//                      angular-pair band of J
// "Stacking" (diatomic): the two coupled densities (Paux0, Paux2) and results (Jaux0, Jaux2) of a
// channel are stacked into one vector of length 2 ntri, the four integral families into one matrix
// fac [[T00, -T02], [-T20, T22]], and the four disjoint factors into s = fac [P0; -P2], b = [Q0; -Q2];
// then the diatomic update rules (:1432-1489) are literally the atomic ones (:879-923) on stacked data.
#pragma once
#include <cuda_runtime.h>
#include "sadatom_grid_mma.cuh"  // dmma()

namespace hfb {

struct ChanParams {
  // ---- constants ----
  int nel, np, np2, ntri, stack, MT, KS, nr, nchan, npair;
  const int2* tri;        // [ntri] (i<=j)
  const double* teiF;     // [nr][nel][MT][KS][32]   fac folded in
  const double* svec;     // [nr][nel][8 MT]         "small" disjoint factor, fac folded in
  const double* bvec;     // [nr][nel][8 MT]         "big" disjoint factor
  const int* chan_r;      // [nchan] family of a channel (channels sorted by family)
  const int* fam_c0;      // [nr+1]  channel range of a family
  const int* pre_ptr;     // [nchan+1] CSR: channel -> (pair, coef[stack])
  const int* pre_pair;
  const double* pre_coef; // [nnz][2]
  const int* post_ptr;    // [npair+1] CSR: pair -> (channel, coef[stack])
  const int* post_chan;
  const double* post_coef;
  // ---- per call ----
  int nb;
  const double* P;        // [nb][npair][nel][np2]
  double* pc;             // [nchan][nb][nel][4 KS]
  double* traces;         // [nchan][nb][nel][2]
  double* jc;             // [nchan][nb][nel][8 MT]
  double* J;              // [nb][npair][nel][np2]
  // zero screening (exact: a zero term adds nothing): pairs whose density band vanishes are skipped by the ket contraction,
  // channels whose radial result vanishes by the bra contraction, pairs whose J band vanishes by the unpack
  const unsigned char* pnz;  // [npair] the density band of the pair has a non-zero entry (nullptr: no screening)
  unsigned char* cnz;        // [nchan] the channel result jc has a non-zero entry
  unsigned char* onz;        // [npair] the J band of the pair has a non-zero entry
  const unsigned char* fam_own;  // [nr] this rank computes the family (sharded builds; the results are summed over ranks)
};

// grid (nel, nchan, nb), 128 threads.  The couplings of the channel whose pair band is non-zero (pnz) are first compacted, IN ORDER,
// into shared memory segment by segment (a linear molecule populates 2 % of the pairs), then every thread accumulates its entries
// over the compacted list only: the summation order is that of the full list with the zero terms left out.
constexpr int CHAN_PRE_SEG = 512;
static __global__ void __launch_bounds__(128) chan_pre_kernel(ChanParams prm) {
  __shared__ double red[8];
  __shared__ int s_pair[CHAN_PRE_SEG];
  __shared__ double s_cf[CHAN_PRE_SEG][2];
  __shared__ int s_wcnt[4], s_n;
  const int el = blockIdx.x, c = blockIdx.y, b = blockIdx.z, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int np = prm.np, np2 = prm.np2, KP = 4 * prm.KS, MP = 8 * prm.MT, r = prm.chan_r[c];
  const int e0 = prm.pre_ptr[c], e1 = prm.pre_ptr[c + 1];
  const double* gP = prm.P + ((size_t)b * prm.npair * prm.nel + el) * np2;
  const size_t pstride = (size_t)prm.nel * np2;
  const double* sv = prm.svec + ((size_t)r * prm.nel + el) * MP;
  const double* bv = prm.bvec + ((size_t)r * prm.nel + el) * MP;
  double* out = prm.pc + (((size_t)c * prm.nb + b) * prm.nel + el) * KP;
  // this thread's entries m = tid, tid + 128, ... (KP <= 4 * 128 is checked by the engine: at most 4 per thread)
  double v[4] = {0.0, 0.0, 0.0, 0.0};
  int ia[4], iat[4], hh[4];
  bool diag[4];
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const int m = tid + 128 * q;
    ia[q] = -1; iat[q] = 0; hh[q] = 0; diag[q] = true;
    if (m < prm.stack * prm.ntri) {
      hh[q] = m >= prm.ntri ? 1 : 0;
      const int2 ij = prm.tri[m - hh[q] * prm.ntri];
      ia[q] = ij.x + ij.y * np;
      iat[q] = ij.y + ij.x * np;
      diag[q] = ij.x == ij.y;
    }
  }
  for (int base = e0; base < e1;) {
    // ---- ordered compaction of up to CHAN_PRE_SEG active couplings starting at `base` ----
    if (tid == 0) s_n = 0;
    __syncthreads();
    int consumed = base;
    while (consumed < e1) {
      const int e = consumed + tid;
      bool act = false;
      int pr = 0;
      if (e < e1) {
        pr = prm.pre_pair[e];
        act = (prm.pnz == nullptr || prm.pnz[pr]) && (prm.pre_coef[2 * e] != 0.0 || prm.pre_coef[2 * e + 1] != 0.0);
      }
      const unsigned bal = __ballot_sync(0xffffffffu, act);
      if (lane == 0) s_wcnt[warp] = __popc(bal);
      __syncthreads();
      const int n0 = s_n;
      const int tot = s_wcnt[0] + s_wcnt[1] + s_wcnt[2] + s_wcnt[3];
      if (n0 + tot > CHAN_PRE_SEG) break;  // block-uniform: this chunk goes into the next segment
      int off = n0;
      for (int w = 0; w < warp; w++) off += s_wcnt[w];
      if (act) {
        const int pos = off + __popc(bal & ((1u << lane) - 1u));
        s_pair[pos] = pr;
        s_cf[pos][0] = prm.pre_coef[2 * e];
        s_cf[pos][1] = prm.pre_coef[2 * e + 1];
      }
      __syncthreads();
      if (tid == 0) s_n = n0 + tot;
      consumed += 128;
      __syncthreads();
    }
    const int nact = s_n;
    // ---- accumulate over the compacted segment ----
#pragma unroll
    for (int q = 0; q < 4; q++) {
      if (ia[q] < 0) continue;
      double acc = v[q];
      for (int x = 0; x < nact; x++) {
        const double cf = s_cf[x][hh[q]];
        if (cf == 0.0) continue;
        const double* src = gP + (size_t)s_pair[x] * pstride;
        double pv = src[ia[q]];
        if (!diag[q]) pv += src[iat[q]];
        acc = fma(cf, pv, acc);
      }
      v[q] = acc;
    }
    __syncthreads();
    base = consumed;
  }
  double ts = 0.0, tb = 0.0;
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const int m = tid + 128 * q;
    if (m >= KP) continue;
    if (ia[q] >= 0) {
      ts = fma(sv[m], v[q], ts);
      tb = fma(bv[m], v[q], tb);
    }
    out[m] = v[q];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ts += __shfl_xor_sync(0xffffffffu, ts, o);
    tb += __shfl_xor_sync(0xffffffffu, tb, o);
  }
  if ((tid & 31) == 0) { red[tid >> 5] = ts; red[4 + (tid >> 5)] = tb; }
  __syncthreads();
  if (tid == 0) {
    double* tr = prm.traces + (((size_t)c * prm.nb + b) * prm.nel + el) * 2;
    tr[0] = (red[0] + red[1]) + (red[2] + red[3]);
    tr[1] = (red[4] + red[5]) + (red[6] + red[7]);
  }
}

// grid (nel, nchan, nb), 128 threads.  Any vector length (used when the stacked triangle exceeds 512 entries: np > 21).
static __global__ void __launch_bounds__(128) chan_pre_generic_kernel(ChanParams prm) {
  __shared__ double red[8];
  const int el = blockIdx.x, c = blockIdx.y, b = blockIdx.z, tid = threadIdx.x;
  const int np = prm.np, np2 = prm.np2, KP = 4 * prm.KS, MP = 8 * prm.MT, r = prm.chan_r[c];
  const int e0 = prm.pre_ptr[c], e1 = prm.pre_ptr[c + 1];
  const double* gP = prm.P + ((size_t)b * prm.npair * prm.nel + el) * np2;
  const size_t pstride = (size_t)prm.nel * np2;
  const double* sv = prm.svec + ((size_t)r * prm.nel + el) * MP;
  const double* bv = prm.bvec + ((size_t)r * prm.nel + el) * MP;
  double* out = prm.pc + (((size_t)c * prm.nb + b) * prm.nel + el) * KP;
  double ts = 0.0, tb = 0.0;
  for (int m = tid; m < KP; m += blockDim.x) {
    double v = 0.0;
    if (m < prm.stack * prm.ntri) {
      const int h = m >= prm.ntri ? 1 : 0;
      const int2 ij = prm.tri[m - h * prm.ntri];
      const int a = ij.x + ij.y * np, at = ij.y + ij.x * np;
      for (int e = e0; e < e1; e++) {
        const double cf = prm.pre_coef[2 * e + h];
        if (cf == 0.0) continue;
        if (prm.pnz && !prm.pnz[prm.pre_pair[e]]) continue;
        const double* src = gP + (size_t)prm.pre_pair[e] * pstride;
        double p = src[a];
        if (ij.x != ij.y) p += src[at];
        v = fma(cf, p, v);
      }
      ts = fma(sv[m], v, ts);
      tb = fma(bv[m], v, tb);
    }
    out[m] = v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ts += __shfl_xor_sync(0xffffffffu, ts, o);
    tb += __shfl_xor_sync(0xffffffffu, tb, o);
  }
  if ((tid & 31) == 0) { red[tid >> 5] = ts; red[4 + (tid >> 5)] = tb; }
  __syncthreads();
  if (tid == 0) {
    double* tr = prm.traces + (((size_t)c * prm.nb + b) * prm.nel + el) * 2;
    tr[0] = (red[0] + red[1]) + (red[2] + red[3]);
    tr[1] = (red[4] + red[5]) + (red[6] + red[7]);
  }
}

// grid (nel * ceil(MT/4), column chunks, nr), 256 threads.  Columns n of family r: (channel - c0) * nb + build.
template <bool SPLITK, int NTW>
__global__ void __launch_bounds__(256) chan_jmma_kernel(ChanParams prm) {
  __shared__ double sred[SPLITK ? 8 * 4 * NTW * 64 : 1];
  const int rts = (prm.MT + 3) >> 2;
  const int el = blockIdx.x / rts, rt = blockIdx.x % rts, r = blockIdx.z;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gid = lane >> 2, t4 = lane & 3;
  const int KS = prm.KS, KP = 4 * KS, MP = 8 * prm.MT;
  const int mt0 = 4 * rt, nmt = min(4, prm.MT - mt0);
  if (prm.fam_own && !prm.fam_own[r]) return;  // block-uniform: another rank's family (the launcher zeroed jc)
  const int c0 = prm.fam_c0[r];
  const int ncol = (prm.fam_c0[r + 1] - c0) * prm.nb;
  const int chunk = SPLITK ? 8 * NTW : 8 * 8 * NTW;
  if (blockIdx.y * chunk >= ncol) return;  // block-uniform
  const int bw0 = blockIdx.y * chunk + (SPLITK ? 0 : warp * 8 * NTW);
  if (!SPLITK && bw0 >= ncol) return;

  double acc[4][NTW][2];
#pragma unroll
  for (int m = 0; m < 4; m++)
#pragma unroll
    for (int n = 0; n < NTW; n++) { acc[m][n][0] = 0.0; acc[m][n][1] = 0.0; }

  const double* A = prm.teiF + ((((size_t)r * prm.nel + el) * prm.MT + mt0) * KS) * 32 + lane;
  const double* colbase = prm.pc + (size_t)c0 * prm.nb * prm.nel * KP;  // column n at n * nel * KP
  const double* Bp[NTW];
  bool bok[NTW];
#pragma unroll
  for (int n = 0; n < NTW; n++) {
    const int col = bw0 + 8 * n + gid;
    bok[n] = col < ncol;
    Bp[n] = colbase + ((size_t)(bok[n] ? col : 0) * prm.nel + el) * KP + t4;
  }
  const int ks0 = SPLITK ? warp : 0, kstep = SPLITK ? 8 : 1;
  double a[4], bf[NTW];
  auto load = [&](int ks, double* aa, double* bb) {
#pragma unroll
    for (int m = 0; m < 4; m++) aa[m] = (m < nmt && ks < KS) ? __ldg(A + ((size_t)m * KS + ks) * 32) : 0.0;
#pragma unroll
    for (int n = 0; n < NTW; n++) bb[n] = (bok[n] && ks < KS) ? __ldg(Bp[n] + 4 * ks) : 0.0;
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
  // epilogue: thread holds Jc[m = 8 (mt0+mm) + gid][column bw0 + 8 n + 2 t4 + cc]
  const double* sv = prm.svec + ((size_t)r * prm.nel + el) * MP;
  const double* bv = prm.bvec + ((size_t)r * prm.nel + el) * MP;
#pragma unroll
  for (int n = 0; n < NTW; n++)
#pragma unroll
    for (int cc = 0; cc < 2; cc++) {
      const int col = bw0 + 8 * n + 2 * t4 + cc;
      if (col >= ncol) continue;
      const double* tr = prm.traces + ((size_t)c0 * prm.nb + col) * prm.nel * 2;
      double below = 0.0, above = 0.0;  // sum_{j<el} jsmall_j ; sum_{j>el} jbig_j
      for (int j = 0; j < el; j++) below += tr[2 * j];
      for (int j = el + 1; j < prm.nel; j++) above += tr[2 * j + 1];
      double* out = prm.jc + (((size_t)c0 * prm.nb + col) * prm.nel + el) * MP;
#pragma unroll
      for (int mm = 0; mm < 4; mm++) {
        const int m = 8 * (mt0 + mm) + gid;
        if (mm < nmt) {
          // J_el += jbig_j * s_el (j > el) + jsmall_j * b_el (j < el); fac is folded into s and T
          double v = acc[mm][n][cc];
          v = fma(above, sv[m], v);
          v = fma(below, bv[m], v);
          out[m] = v;
        }
      }
    }
}

// [nchan] flags of the channel results (after the all-reduce of a sharded run): grid (nchan), 128 threads
static __global__ void __launch_bounds__(128) chan_flag_kernel(ChanParams prm) {
  const int c = blockIdx.x;
  const double* src = prm.jc + (size_t)c * prm.nb * prm.nel * 8 * prm.MT;
  int any = 0;
  for (int k = threadIdx.x; k < prm.nb * prm.nel * 8 * prm.MT; k += blockDim.x) any |= (src[k] != 0.0);
  any = __syncthreads_or(any);
  if (threadIdx.x == 0) prm.cnz[c] = any ? 1 : 0;
}

// grid (npair, 1, nb), 256 threads: all elements of one angular pair.  The couplings to channels with a non-zero result (cnz) are
// compacted in order into shared memory; a pair without any leaves its band untouched and its onz flag clear (the unpack then
// writes zeros without reading the band).
constexpr int CHAN_POST_SEG = 256;
static __global__ void __launch_bounds__(256) chan_post_kernel(ChanParams prm) {
  __shared__ int s_chan[CHAN_POST_SEG];
  __shared__ double s_cf[CHAN_POST_SEG][2];
  __shared__ int s_wcnt[8], s_n;
  const int pr = blockIdx.x, b = blockIdx.z, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int np = prm.np, np2 = prm.np2, MP = 8 * prm.MT, nel = prm.nel;
  const int e0 = prm.post_ptr[pr], e1 = prm.post_ptr[pr + 1];
  int any_act = 0;
  for (int e = e0 + tid; e < e1; e += 256) any_act |= (prm.cnz == nullptr || prm.cnz[prm.post_chan[e]]);
  if (!__syncthreads_or(any_act)) return;
  double* out = prm.J + ((size_t)b * prm.npair + pr) * nel * np2;
  const int total = nel * np2;
  int any = 0;
  for (int k0 = 0; k0 < total; k0 += 2048) {  // sweeps of 2048 band entries (8 per thread); one sweep unless nel np^2 > 2048
  // this thread's entries k = k0 + tid, + 256, ... of the pair's band [element][np2]
  double v[8];
  int mm[8], ee[8];
#pragma unroll
  for (int q = 0; q < 8; q++) {
    const int k = k0 + tid + 256 * q;
    v[q] = 0.0; mm[q] = -1; ee[q] = 0;
    if (k < total) {
      const int el = k / np2, kk = k - el * np2, i = kk % np, j = kk / np;
      const int lo = min(i, j), hi = max(i, j);
      mm[q] = lo + hi * (hi + 1) / 2;
      ee[q] = el;
    }
  }
  for (int base = e0; base < e1;) {
    __syncthreads();
    if (tid == 0) s_n = 0;
    __syncthreads();
    int consumed = base;
    while (consumed < e1) {
      const int e = consumed + tid;
      bool act = false;
      int c = 0;
      if (e < e1) {
        c = prm.post_chan[e];
        act = prm.cnz == nullptr || prm.cnz[c];
      }
      const unsigned bal = __ballot_sync(0xffffffffu, act);
      if (lane == 0) s_wcnt[warp] = __popc(bal);
      __syncthreads();
      const int n0 = s_n;
      int tot = 0, off = n0;
      for (int w = 0; w < 8; w++) { if (w < warp) off += s_wcnt[w]; tot += s_wcnt[w]; }
      if (n0 + tot > CHAN_POST_SEG) break;  // block-uniform
      if (act) {
        const int pos = off + __popc(bal & ((1u << lane) - 1u));
        s_chan[pos] = c;
        s_cf[pos][0] = prm.post_coef[2 * e];
        s_cf[pos][1] = prm.post_coef[2 * e + 1];
      }
      __syncthreads();
      if (tid == 0) s_n = n0 + tot;
      consumed += 256;
      __syncthreads();
    }
    const int nact = s_n;
#pragma unroll
    for (int q = 0; q < 8; q++) {
      if (mm[q] < 0) continue;
      double acc = v[q];
      for (int x = 0; x < nact; x++) {
        const double* src = prm.jc + (((size_t)s_chan[x] * prm.nb + b) * nel + ee[q]) * MP;
        acc = fma(s_cf[x][0], src[mm[q]], acc);
        if (prm.stack == 2) acc = fma(s_cf[x][1], src[prm.ntri + mm[q]], acc);
      }
      v[q] = acc;
    }
    base = consumed;
  }
#pragma unroll
  for (int q = 0; q < 8; q++) {
    const int k = k0 + tid + 256 * q;
    if (k < total) { out[k] = v[q]; any |= (v[q] != 0.0); }
  }
  }  // sweeps
  if (prm.onz) {
    any = __syncthreads_or(any);
    if (tid == 0 && any) prm.onz[pr] = 1;
  }
}

// Device-resident operator calls: gather the element-diagonal band of every angular pair out of a dense matrix,
// grid (nel * npair), and the inverse (a radial index pair lies in at most two common elements), grid-stride.
static __global__ void __launch_bounds__(128) chan_band_pack_kernel(ChanParams prm, const int* first, const int* npe, int nang, int nrad,
                                                                    const double* D, int ld, double* band, unsigned char* pnz,
                                                                    const unsigned char* nzx = nullptr) {
  const int el = blockIdx.x % prm.nel, pr = blockIdx.x / prm.nel, k = pr / nang, l = pr % nang;
  if (nzx && !nzx[pr]) return;  // the whole angular block of D is exactly zero (mask of the same build): pnz stays clear, the band is never read
  const int f = first[el], n = npe[el], np = prm.np;
  double* dst = band + ((size_t)pr * prm.nel + el) * prm.np2;
  int any = 0;
  for (int q = threadIdx.x; q < prm.np2; q += blockDim.x) {
    const int i = q % np, j = q / np;
    const double v = (i < n && j < n) ? D[(size_t)(k * nrad + f + i) + (size_t)(l * nrad + f + j) * ld] : 0.0;
    dst[q] = v;
    any |= (v != 0.0);
  }
  any = __syncthreads_or(any);
  if (threadIdx.x == 0 && any) pnz[pr] = 1;  // preset to zero; every writer stores the same value
}
// grid (npair), 256 threads: the (nrad x nrad) block of one angular pair; zeros without touching the band when the pair's
// J band vanishes (onz)
static __global__ void __launch_bounds__(256) chan_band_unpack_kernel(ChanParams prm, const int* first, const int* elem_of, int nang, int nrad,
                                                                      const double* band, double* D, int ld, const unsigned char* onz) {
  const int pr = blockIdx.x;
  double* dst = D + (size_t)(pr / nang) * nrad + (size_t)(pr % nang) * nrad * ld;
  const bool live = onz == nullptr || onz[pr] != 0;
  const double* src = band + (size_t)pr * prm.nel * prm.np2;
  for (int idx = threadIdx.x; idx < nrad * nrad; idx += 256) {
    const int a = idx % nrad, b = idx / nrad;
    double s = 0.0;
    if (live) {
      for (int x = 0; x < 2; x++) {
        const int e = elem_of[2 * a + x];
        if (e < 0) continue;
        if (e == elem_of[2 * b] || e == elem_of[2 * b + 1]) s += src[(size_t)e * prm.np2 + (a - first[e]) + (b - first[e]) * prm.np];
      }
    }
    dst[a + (size_t)b * ld] = s;
  }
}

}  // namespace hfb
