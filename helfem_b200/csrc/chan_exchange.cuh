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
//   x1  W^a_c[j,l]  = sum_i c^a_c(j,i) P[i,l]                        (left angular contraction, sparse rows; tiled:
//                     rows that share their partners stream every P block once)
//   x2  R^{ab}[j,k] = fac_r (+-) sum_{c in r} sum_l c^b_c(k,l) W^a_c[j,l]   (right angular contraction), written
//                     per element pair as zero-padded (np x np) sub-blocks
//   x3  Ksub(ie,je) = sum_ab T^{ab}_r(ie,je)[Rsub^{ab}(ie,je)]       FP64 tensor pipe (DMMA.8x8x4):
//         ie == je:  (np^2 x np^2) K-ordered integrals x [np^2 x NI items]  -- a GEMM whose N dimension
//                    batches NI output blocks, integrals pre-swizzled into A-fragment order
//         ie != je:  iint . Rsub . jint^T  as two chained (np x np) DMMA products per item
//   x4  K[j,k] -= the (at most four) sub-blocks that overlap each radial index pair   (fixed order)
// Zero blocks of P are skipped like the reference does (Frobenius norm < 10 eps, atomic/TwoDBasis.cpp:1036):
// the block mask of P fixes, once per sparsity pattern and on the host, the set L(c,j) of W blocks every coupling
// row produces (W is stored compactly, one block per member), the output blocks ("items") of every family, and a
// partition of the families into BATCHES whose W / R workspaces fit a fixed budget.  One launch of each stage
// covers a whole batch of families, so that the grid fills the 148 SMs even when a single family is small.
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
  const int4* eltab;      // [nrad*nrad] offsets of the radial pair (A,B) inside the (<= 4) element-pair sub-blocks it belongs to, -1 = none
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
  // family tables and the per-sparsity-pattern lists (all families; a launch covers a BATCH of families)
  const int* fam_c0;      // [nr+1] channel range of a family
  const double* fam_fac;  // [nr]
  const int2* items;      // [nitems] output blocks (j,k), grouped by family
  const int* item_fam;    // [nitems]
  // compact storage of W: row (c,j) holds one block per l of its set L(c,j) = { l : some i of the row has P[i,l] != 0 }
  const long long* wbase; // [nchan*nang] first block of the row inside the batch's W buffer
  const unsigned long long* wbits;  // [nchan*nang][nw] bitset of L(c,j)
  int nw;
  double* W;              // [blocks of the batch][stack][nrad*nrad]
  int item0;              // first item of the batch (Rsub / blocks are indexed relative to it)
  double* Rsub;           // [items of the batch][stack*stack][nel*nel][np2]
  double* blocks;         // [items of the batch][nel*nel][np2]
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

// Tiling of the two angular contractions: a CTA owns up to CHX_T rows that share their coupled partners (same m),
// builds the UNION of the partners in shared memory (flags -> ballot scan -> slots) with a dense (slot x row)
// coefficient table, and streams every input block ONCE for all its rows instead of once per row.
constexpr int CHX_T = 8;      // rows (x1: j, x2: k) per tile
constexpr int CHX_UMAX = 64;  // union slots per pass
#ifndef HFB_X2_STAGES
#define HFB_X2_STAGES 6
#endif
#ifndef HFB_X2_MINB
#define HFB_X2_MINB 2
#endif
constexpr int CHX_X2_STAGES = HFB_X2_STAGES;  // k-steps in flight per warp in the x2 stream (cp.async ring)
constexpr int CHX_LDC = 12;   // row stride of the DMMA coefficient tables: the 4 x 8 fragment read of a half-warp hits 32 distinct banks

// slots for the flagged entries of dom[0..n): -1 = absent, 0 = present on entry; slot number on exit.  Warp 0 only.
__device__ __forceinline__ int chx_scan_slots(int* dom, int* ulist, int n, int lane) {
  int base = 0;
  for (int q0 = 0; q0 < n; q0 += 32) {
    const int q = q0 + lane;
    const bool f = q < n && dom[q] == 0;
    const unsigned m = __ballot_sync(0xffffffffu, f);
    if (f) {
      const int sl = base + __popc(m & ((1u << lane) - 1u));
      dom[q] = sl;
      ulist[sl] = q;
    }
    base += __popc(m);
  }
  return base;
}

// Block of W for (row, l), or -1 when l is not in the row's set.
__device__ __forceinline__ long long chx_wblock(const ChanXParams& prm, int row, int l) {
  const unsigned long long* bits = prm.wbits + (size_t)row * prm.nw;
  const int w = l >> 6;
  const unsigned long long m = bits[w];
  if (!((m >> (l & 63)) & 1ull)) return -1;
  long long idx = prm.wbase[row] + __popcll(m & ((1ull << (l & 63)) - 1ull));
  for (int q = 0; q < w; q++) idx += __popcll(bits[q]);
  return idx;
}

// x1.  grid (tiles of the batch, l chunks), 256 threads.  tiles: [c, nj, j_0 .. j_{T-1}] per tile.
// On the FP64 tensor pipe: W_{(c,j)}[l][k] = sum_v coef[u_v][j][b] P_{i_v l}[k] as a GEMM with M = the radial pairs k (A operand read from
// the dense P in fragment order), N = the tile's <= 8 rows j (one n-tile per stack half b), K = the active partners of this l.
// dynamic smem: coef[2][UMAX+1][CHX_LDC] doubles (slot UMAX stays zero: padding) | wblk[8][T] long long | ulist[nang] | dom[nang] |
//               act[8][UMAX+4] | actl[8][UMAX+4] ints | done[nang] | cur[nang] bytes
static __global__ void __launch_bounds__(256, 3) chan_x1_mma_kernel(ChanXParams prm, const int* tiles) {
  extern __shared__ double sm[];
  __shared__ int s_cnt, s_na[8];
  const int nang = prm.nang, nrad = prm.nrad, nrad2 = nrad * nrad, st = prm.stack;
  constexpr int CB = (CHX_UMAX + 1) * CHX_LDC;  // one stack half of the coefficient table
  double* coef = sm;
  constexpr int NA = CHX_UMAX + 4;
  long long* wblk_all = reinterpret_cast<long long*>(coef + 2 * CB);  // [8 l of a group][T]
  int* ulist = reinterpret_cast<int*>(wblk_all + 8 * CHX_T);
  int* dom = ulist + nang;
  int* act_all = dom + nang;         // [8][NA] coefficient slot of the v-th active partner
  int* actl_all = act_all + 8 * NA;  // [8][NA] its angular index times nrad (row offset of its P block)
  unsigned char* done = reinterpret_cast<unsigned char*>(actl_all + 8 * NA);  // [nang] an EARLIER pass wrote the blocks of l
  unsigned char* cur = done + nang;                                           // [nang] this pass writes them
  const int* T = tiles + (size_t)blockIdx.x * (2 + CHX_T);
  const int c = T[0], nj = T[1];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, t4 = lane & 3;
  for (int q = tid; q < nang; q += 256) { dom[q] = -1; done[q] = 0; cur[q] = 0; }
  __syncthreads();
  for (int jj = 0; jj < nj; jj++) {
    const int row = c * nang + T[2 + jj];
    for (int e = prm.rowptr[row] + tid; e < prm.rowptr[row + 1]; e += 256) dom[prm.row_i[e]] = 0;
  }
  __syncthreads();
  if (warp == 0) {
    const int n = chx_scan_slots(dom, ulist, nang, lane);
    if (lane == 0) s_cnt = n;
  }
  __syncthreads();
  const int cnt = s_cnt;
  for (int ubase = 0; ubase < cnt; ubase += CHX_UMAX) {
    const int nu = min(CHX_UMAX, cnt - ubase);
    for (int q = tid; q < 2 * CB; q += 256) coef[q] = 0.0;
    __syncthreads();
    for (int jj = 0; jj < nj; jj++) {
      const int row = c * nang + T[2 + jj];
      for (int e = prm.rowptr[row] + tid; e < prm.rowptr[row + 1]; e += 256) {
        const int sl = dom[prm.row_i[e]] - ubase;
        if (sl >= 0 && sl < nu) {
          coef[sl * CHX_LDC + jj] = prm.row_coef[2 * e];
          coef[CB + sl * CHX_LDC + jj] = prm.row_coef[2 * e + 1];
        }
      }
    }
    // groups of eight l: every warp compacts the partner list of one of them (no serial section, two block barriers per group
    // instead of per l), then the CTA streams the eight l one after the other
    for (int lg = blockIdx.y; lg < nang; lg += gridDim.y * 8) {
      __syncthreads();  // coef complete / previous group done with its lists
      {
        const int l = lg + warp * gridDim.y;
        int* act = act_all + warp * NA;
        int* actl = actl_all + warp * NA;
        int base = 0;
        if (l < nang) {  // warp-uniform
          for (int q0 = 0; q0 < nu; q0 += 32) {
            const int q = q0 + lane;
            const int lq = q < nu ? ulist[ubase + q] : 0;
            const bool f = q < nu && prm.nzP[(size_t)lq * nang + l];
            const unsigned mk = __ballot_sync(0xffffffffu, f);
            if (f) {
              const int v = base + __popc(mk & ((1u << lane) - 1u));
              act[v] = q;
              actl[v] = lq * nrad;
            }
            base += __popc(mk);
          }
          __syncwarp();
          // padded to k-steps of four with the zero slot
          if (base > 0 && lane < ((4 - (base & 3)) & 3)) { act[base + lane] = CHX_UMAX; actl[base + lane] = actl[0]; }
          if (lane < CHX_T) wblk_all[warp * CHX_T + lane] = lane < nj ? chx_wblock(prm, c * nang + T[2 + lane], l) : -1;
          if (lane == 0 && base > 0) cur[l] = 1;
        }
        if (lane == 0) s_na[warp] = base;
      }
      __syncthreads();
      for (int w = 0; w < 8; w++) {
      const int l = lg + w * gridDim.y;
      if (l >= nang) break;
      const int na = s_na[w];
      if (na == 0) continue;  // block-uniform
      const int* act = act_all + w * NA;
      const int* actl = actl_all + w * NA;
      const bool prev = done[l] != 0;
      const long long wb0 = wblk_all[w * CHX_T + 2 * t4], wb1 = wblk_all[w * CHX_T + 2 * t4 + 1];
      for (int k0 = warp * 32; k0 < nrad2; k0 += 256) {
        const int kg = k0 + gid;
        unsigned so[4];  // element offsets inside the dense P fit 32 bits (ldp^2 < 2^32 is checked by the engine)
#pragma unroll
        for (int mt = 0; mt < 4; mt++) {
          const int k = min(kg + 8 * mt, nrad2 - 1), A = k % nrad, B = k / nrad;
          so[mt] = (unsigned)A + (unsigned)(l * nrad + B) * (unsigned)prm.ldp;
        }
        double acc[4][2][2];
#pragma unroll
        for (int mt = 0; mt < 4; mt++) { acc[mt][0][0] = acc[mt][0][1] = acc[mt][1][0] = acc[mt][1][1] = 0.0; }
#pragma unroll 2
        for (int v0 = 0; v0 < na; v0 += 4) {
          const int ro = actl[v0 + t4], u = act[v0 + t4];
          double af[4];
#pragma unroll
          for (int mt = 0; mt < 4; mt++) af[mt] = prm.P[so[mt] + (unsigned)ro];
          const double b0 = coef[u * CHX_LDC + gid], b1 = coef[CB + u * CHX_LDC + gid];
#pragma unroll
          for (int mt = 0; mt < 4; mt++) {
            dmma(acc[mt][0][0], acc[mt][0][1], af[mt], b0);
            dmma(acc[mt][1][0], acc[mt][1][1], af[mt], b1);
          }
        }
#pragma unroll
        for (int mt = 0; mt < 4; mt++) {
          const int k = kg + 8 * mt;
          if (k >= nrad2) continue;
#pragma unroll
          for (int i = 0; i < 2; i++) {
            const long long wb = i ? wb1 : wb0;
            if (wb < 0) continue;  // l is not in the set of this row (its partners' blocks are all zero): nothing stored
            double* W0 = prm.W + (size_t)wb * st * nrad2 + k;
            double v0 = acc[mt][0][i], v1 = acc[mt][1][i];
            if (prev) { v0 += W0[0]; if (st == 2) v1 += W0[nrad2]; }
            __stcs(W0, v0);  // streaming: W is far larger than L2 and is next read by x2
            if (st == 2) __stcs(W0 + nrad2, v1);
          }
        }
      }
      }  // w
    }
    __syncthreads();
    for (int q = tid; q < nang; q += 256) { done[q] |= cur[q]; cur[q] = 0; }
  }
}

// x2.  grid (tiles of the batch, element chunks, stack half a), 256 threads.  A tile = <= CHX_T consecutive items (j, k_0..) of one family,
// one j.  On the FP64 tensor pipe: R[el][(item, b)] = sum_u W_u[el] coef[u][item][b] is a GEMM with M = the radial pairs (A operand, streamed
// from HBM directly in fragment order: lane (gid, t4) reads W_{u0+t4}[el0 + gid], 64-byte runs), N = the tile's <= 8 items (one n-tile per
// stack half b; B operand = the coefficient table in shared memory), K = the union slots: one DMMA pair per four streamed values instead
// of 8 LDS.128 + 16 DFMA per value.  The accumulator layout (lane holds radial pair gid, items
// 2 t4, 2 t4 + 1) stores 64-byte runs per item, and the scatter offsets come from the precomputed eltab.
// dynamic smem: coef[2][UMAX][CHX_LDC] doubles | woff[UMAX] (size_t) | ulist[D] | dom[D] ints |
//   stage[8 warps][CHX_X2_STAGES][4][32] doubles
static __global__ void __launch_bounds__(256, HFB_X2_MINB) chan_x2_mma_kernel(ChanXParams prm, const int2* tiles, int nch_max) {
  extern __shared__ double sm[];
  __shared__ int s_cnt;
  const int nang = prm.nang, nrad = prm.nrad, nrad2 = nrad * nrad, st = prm.stack, st2 = st * st;
  double* coef = sm;
  size_t* woff = reinterpret_cast<size_t*>(coef + 2 * CHX_UMAX * CHX_LDC);
  int* ulist = reinterpret_cast<int*>(woff + CHX_UMAX);
  int* dom = ulist + nch_max * nang;
  double* stage = reinterpret_cast<double*>(dom + nch_max * nang);  // [8 warps][STAGES][4][32]
  const int it0 = tiles[blockIdx.x].x, nk = tiles[blockIdx.x].y;
  const int j = prm.items[it0].x, r = prm.item_fam[it0];
  const int c0 = prm.fam_c0[r], nch = prm.fam_c0[r + 1] - c0, D = nch * nang;
  const double fac = prm.fam_fac[r];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, t4 = lane & 3;
  for (int q = tid; q < D; q += 256) dom[q] = -1;
  __syncthreads();
  for (int kk = warp; kk < nk; kk += 8) {  // one warp per item: the rows are short, the chains of dependent loads run side by side
    const int k = prm.items[it0 + kk].y;
    for (int cl = 0; cl < nch; cl++) {
      const int row = (c0 + cl) * nang + k, jrow = (c0 + cl) * nang + j;
      const unsigned long long* bits = prm.wbits + (size_t)jrow * prm.nw;
      for (int e = prm.rowptr[row] + lane; e < prm.rowptr[row + 1]; e += 32) {
        const int l = prm.row_i[e];
        if ((bits[l >> 6] >> (l & 63)) & 1ull) dom[cl * nang + l] = 0;
      }
    }
  }
  __syncthreads();
  if (warp == 0) {
    const int n = chx_scan_slots(dom, ulist, D, lane);
    if (lane == 0) s_cnt = n;
  }
  __syncthreads();
  const int cnt = s_cnt;
  const int nel = prm.nel, nel2 = nel * nel, np2 = prm.np2;
  const size_t istride = (size_t)st2 * nel2 * np2, bstride = (size_t)nel2 * np2;
  for (int ubase = 0; ubase < max(cnt, 1); ubase += CHX_UMAX) {  // cnt == 0: one pass that writes zeros
    const int nu = max(0, min(CHX_UMAX, cnt - ubase));
    const int nup = (nu + 3) & ~3;  // k-steps of four slots; the padding slots keep zero coefficients
    __syncthreads();
    for (int q = tid; q < 2 * CHX_UMAX * CHX_LDC; q += 256) coef[q] = 0.0;
    if (tid < nup) {
      const int q = ulist[ubase + (tid < nu ? tid : 0)], cl = q / nang, l = q - cl * nang;
      woff[tid] = (size_t)chx_wblock(prm, (c0 + cl) * nang + j, l) * st * nrad2;
    }
    __syncthreads();
    for (int kk = warp; kk < nk; kk += 8) {
      const int k = prm.items[it0 + kk].y;
      for (int cl = 0; cl < nch; cl++) {
        const int row = (c0 + cl) * nang + k;
        for (int e = prm.rowptr[row] + lane; e < prm.rowptr[row + 1]; e += 32) {
          const int d = dom[cl * nang + prm.row_i[e]];
          const int sl = d - ubase;
          if (d >= 0 && sl >= 0 && sl < nu) {
            coef[sl * CHX_LDC + kk] = prm.row_coef[2 * e];
            coef[(CHX_UMAX + sl) * CHX_LDC + kk] = prm.row_coef[2 * e + 1];
          }
        }
      }
    }
    __syncthreads();
    const bool it_ok0 = 2 * t4 < nk, it_ok1 = 2 * t4 + 1 < nk;
    for (int a = blockIdx.z; a < st; a += gridDim.z) {  // both stack halves reuse the tables built above
    const double* Wa = prm.W + (size_t)a * nrad2;
    // output rows of this lane's two items, stack block (a, b = 0), as 32-bit element offsets (the engine caps Rsub below 2^32 doubles)
    const unsigned ob0 = (unsigned)((size_t)(it0 - prm.item0 + 2 * t4) * istride + (size_t)(a * st) * bstride);
    const unsigned ist = (unsigned)istride, bst = (unsigned)bstride;
    const double sg0 = (a != 0) ? -fac : fac, sg1 = (a != 1) ? -fac : fac;  // sign of the stack block (a, b)
    // The union of a tile is small (a few k-steps per radial tile), so the loop is a stream of short load -> DMMA -> store rounds.  The
    // fragments are therefore staged through a per-lane cp.async ring (CHX_X2_STAGES k-steps deep, each lane reads back exactly what it
    // copied: no barrier), with the producer running ahead ACROSS radial tiles, so that the loads of the next tiles are in flight during
    // the scatter of the current one.
    const int ksn = nup >> 2;
    const int el_first = blockIdx.y * 256 + warp * 32, el_stride = gridDim.y * 256;
    const double* stg = stage + (size_t)warp * (CHX_X2_STAGES * 128) + lane;
    const unsigned stg_s = (unsigned)__cvta_generic_to_shared(stg);
    int p_el = el_first, p_ks = 0, p_slot = 0;
    auto produce = [&]() {
      if (p_el < nrad2) {
        // radial pairs past the end are clamped to the last one: their accumulator rows are never stored
        const int pg = p_el + gid;
        const double* src = Wa + woff[4 * p_ks + t4];
#pragma unroll
        for (int mt = 0; mt < 4; mt++)
          asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(stg_s + (unsigned)((p_slot * 4 + mt) * 256)),
                       "l"(src + min(pg + 8 * mt, nrad2 - 1)) : "memory");
        if (++p_ks == ksn) { p_ks = 0; p_el += el_stride; }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      p_slot = (p_slot + 1 == CHX_X2_STAGES) ? 0 : p_slot + 1;
    };
    if (ksn > 0)
      for (int q = 0; q < CHX_X2_STAGES - 1; q++) produce();
    int c_slot = 0;
    for (int el0 = el_first; el0 < nrad2; el0 += el_stride) {
      const int elg = el0 + gid;
      int4 e4[4];  // scatter offsets of this lane's four radial pairs (their latency hides behind the k loop)
#pragma unroll
      for (int mt = 0; mt < 4; mt++) e4[mt] = (elg + 8 * mt < nrad2) ? prm.eltab[elg + 8 * mt] : make_int4(-1, -1, -1, -1);
      double acc[4][2][2];
#pragma unroll
      for (int mt = 0; mt < 4; mt++) { acc[mt][0][0] = acc[mt][0][1] = acc[mt][1][0] = acc[mt][1][1] = 0.0; }
#pragma unroll 1
      for (int ks = 0; ks < ksn; ks++) {
        produce();
        asm volatile("cp.async.wait_group %0;" ::"n"(CHX_X2_STAGES - 1) : "memory");
        double af[4];
#pragma unroll
        for (int mt = 0; mt < 4; mt++) af[mt] = stg[(c_slot * 4 + mt) * 32];
        const double b0 = coef[(4 * ks + t4) * CHX_LDC + gid], b1 = coef[(CHX_UMAX + 4 * ks + t4) * CHX_LDC + gid];
#pragma unroll
        for (int mt = 0; mt < 4; mt++) {
          dmma(acc[mt][0][0], acc[mt][0][1], af[mt], b0);
          dmma(acc[mt][1][0], acc[mt][1][1], af[mt], b1);
        }
        c_slot = (c_slot + 1 == CHX_X2_STAGES) ? 0 : c_slot + 1;
      }
      // scatter into the element-pair sub-blocks: a radial function belongs to one or two elements, so a pair has 1, 2 or 4 homes
      // (eltab lists the valid offsets first); the loop is rolled, the common single-home case runs its body once
#pragma unroll
      for (int mt = 0; mt < 4; mt++) {
        int4 o4 = e4[mt];  // all -1 past the end
        const double v00 = sg0 * acc[mt][0][0], v01 = sg0 * acc[mt][0][1], v10 = sg1 * acc[mt][1][0], v11 = sg1 * acc[mt][1][1];
#pragma unroll 1
        while (o4.x >= 0) {
          const unsigned o = ob0 + (unsigned)o4.x;
          if (ubase == 0) {  // streaming stores: Rsub is far larger than L2 and is next read by x3
            if (it_ok0) { __stcs(prm.Rsub + o, v00); if (st == 2) __stcs(prm.Rsub + o + bst, v10); }
            if (it_ok1) { __stcs(prm.Rsub + o + ist, v01); if (st == 2) __stcs(prm.Rsub + o + ist + bst, v11); }
          } else {
            if (it_ok0) { prm.Rsub[o] += v00; if (st == 2) prm.Rsub[o + bst] += v10; }
            if (it_ok1) { prm.Rsub[o + ist] += v01; if (st == 2) prm.Rsub[o + ist + bst] += v11; }
          }
          o4 = make_int4(o4.y, o4.z, o4.w, -1);
        }
      }
    }
    }  // a
  }
}

// x3.  Two kernels, 256 threads, over (item groups of the batch) x (element pairs); groups: (first item, count <= NI) of
// ONE family.  OFFD = false: same element, grid.y = nel, dynamic smem NI*LDK doubles.  OFFD = true: ordered pairs of
// different elements, grid.y = nel*(nel-1), dynamic smem 8 warps x (8 NTL)(8 NTL + 1) doubles.
// NTL = 8-row tiles of an (np x np) block (np <= 8 NTL).
template <int NI, int NTL, bool OFFD>
__global__ void __launch_bounds__(256, 2) chan_x3_kernel(ChanXParams prm, const int2* groups) {
  extern __shared__ double sm[];
  const int nel = prm.nel, nel2 = nel * nel, np = prm.np, np2 = prm.np2, st = prm.stack, st2 = st * st;
  int ie, je;
  if (OFFD) {
    ie = blockIdx.y / (nel - 1);
    const int jj = blockIdx.y % (nel - 1);
    je = jj + (jj >= ie ? 1 : 0);
  } else {
    ie = je = blockIdx.y;
  }
  const int r = prm.item_fam[groups[blockIdx.x].x];
  const int n0 = groups[blockIdx.x].x - prm.item0, nn = groups[blockIdx.x].y;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, t4 = lane & 3;
  const size_t istride = (size_t)st2 * nel2 * np2;  // Rsub stride between items
  const double* Rbase = prm.Rsub + (size_t)n0 * istride + (size_t)(ie * nel + je) * np2;
  double* Obase = prm.blocks + ((size_t)n0 * nel2 + ie * nel + je) * np2;

  if constexpr (!OFFD) {
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
        // B tile: every element goes global -> shared by cp.async (zero fill for the padding), all copies of the CTA in flight
        // together instead of one load -> store round trip per element
        {
          const unsigned sm_s = (unsigned)__cvta_generic_to_shared(sm);
          const double* Rab = Rbase + (size_t)ab * nel2 * np2;
          for (int idx = tid; idx < NI * LDK; idx += 256) {
            const int n = idx / LDK, k = idx - n * LDK;
            const bool ok = n < nn && k < np2;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(sm_s + (unsigned)idx * 8u),
                         "l"(ok ? Rab + (size_t)n * istride + k : Rbase), "r"(ok ? 8 : 0) : "memory");
          }
          asm volatile("cp.async.commit_group;" ::: "memory");
          asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        const double* A = prm.teiK + ((((size_t)r * nel + ie) * st2 + ab) * MT) * KS * 32 + lane;
        const double* bs = sm + gid * LDK + t4;
        // software pipeline: the A fragments of k-step ks + 1 are requested before the DMMAs of step ks are issued, so the L2
        // latency of the integral stream hides behind a full step of tensor work instead of surfacing at every unroll boundary
        double a[4], an[4];
#pragma unroll
        for (int mi = 0; mi < 4; mi++) {
          const int mt = mt0 + warp + 8 * mi;
          a[mi] = (mt < MT) ? __ldg(A + (size_t)mt * KS * 32) : 0.0;
        }
#pragma unroll 2
        for (int ks = 0; ks < KS; ks++) {
          double b[NTN];
#pragma unroll
          for (int mi = 0; mi < 4; mi++) {
            const int mt = mt0 + warp + 8 * mi;
            an[mi] = (mt < MT && ks + 1 < KS) ? __ldg(A + ((size_t)mt * KS + ks + 1) * 32) : 0.0;
          }
#pragma unroll
          for (int nt = 0; nt < NTN; nt++) b[nt] = bs[8 * nt * LDK + 4 * ks];
#pragma unroll
          for (int mi = 0; mi < 4; mi++) {
            if (mt0 + warp + 8 * mi < MT) {  // warp-uniform: a missing m-tile (np2 = 225: 29 tiles on 32 slots) costs no DMMA
#pragma unroll
              for (int nt = 0; nt < NTN; nt++) dmma(acc[mi][nt][0], acc[mi][nt][1], a[mi], b[nt]);
            }
          }
#pragma unroll
          for (int mi = 0; mi < 4; mi++) a[mi] = an[mi];
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
            if (row < np2 && n < nn) __stcs(Obase + (size_t)n * nel2 * np2 + row, acc[mi][nt][cc]);
            acc[mi][nt][cc] = 0.0;
          }
      }
    }
  } else {
    // ---------------- different elements: Ksub = sum_ab iint^a . Rsub^{ab} . (jint^b)^T ----------------
    // A warp owns the items n = warp, warp + 8, ...; the fragments of iint / jint are loaded once per (a,b) and the
    // accumulators of all its items stay in registers (small element blocks), or one item at a time (large ones).
    // row stride = 4 (mod 16) doubles: the A-fragment reads of step 2 (row gid, column 4 ks + t4) are bank-conflict free
    constexpr int LDY = (8 * NTL + 15) / 16 * 16 + 4;
    constexpr int NIT = (NTL <= 2 && NI >= 16) ? 2 : 1;  // items a warp carries through the ab loop together
    double* Y = sm + warp * (8 * NTL * LDY);
    const bool lower = ie > je;
    for (int nbase = warp; nbase < nn; nbase += 8 * NIT) {
      double kacc[NIT][NTL][NTL][2];
#pragma unroll
      for (int it = 0; it < NIT; it++)
#pragma unroll
        for (int mt = 0; mt < NTL; mt++)
#pragma unroll
          for (int qt = 0; qt < NTL; qt++) { kacc[it][mt][qt][0] = 0.0; kacc[it][mt][qt][1] = 0.0; }
      for (int ab = 0; ab < st2; ab++) {
        const int a = ab / st, b = ab % st;
        const double* iint = (lower ? prm.big_f : prm.small_f) + (((size_t)r * nel + ie) * st + a) * np2;
        const double* jint = (lower ? prm.small_f : prm.big_f) + (((size_t)r * nel + je) * st + b) * np2;
        double af[2 * NTL][NTL], jf[2 * NTL][NTL];
#pragma unroll
        for (int ks = 0; ks < 2 * NTL; ks++) {
          const int kk = 4 * ks + t4;
#pragma unroll
          for (int mt = 0; mt < NTL; mt++) {
            const bool ok = kk < np && 8 * mt + gid < np;
            af[ks][mt] = ok ? __ldg(iint + (8 * mt + gid) + kk * np) : 0.0;
            jf[ks][mt] = ok ? __ldg(jint + (8 * mt + gid) + kk * np) : 0.0;
          }
        }
#pragma unroll
        for (int it = 0; it < NIT; it++) {
          const int n = nbase + 8 * it;
          if (n >= nn) break;  // warp-uniform
          const double* Rn = Rbase + (size_t)n * istride + (size_t)ab * nel2 * np2;
          // step 1: Y = iint . Rsub
          double rf[2 * NTL][NTL];
#pragma unroll
          for (int ks = 0; ks < 2 * NTL; ks++) {
            const int kk = 4 * ks + t4;
#pragma unroll
            for (int nt = 0; nt < NTL; nt++) rf[ks][nt] = (kk < np && 8 * nt + gid < np) ? __ldcs(Rn + kk + (8 * nt + gid) * np) : 0.0;
          }
          double y[NTL][NTL][2];
#pragma unroll
          for (int mt = 0; mt < NTL; mt++)
#pragma unroll
            for (int nt = 0; nt < NTL; nt++) { y[mt][nt][0] = 0.0; y[mt][nt][1] = 0.0; }
#pragma unroll
          for (int ks = 0; ks < 2 * NTL; ks++)
#pragma unroll
            for (int mt = 0; mt < NTL; mt++)
#pragma unroll
              for (int nt = 0; nt < NTL; nt++) dmma(y[mt][nt][0], y[mt][nt][1], af[ks][mt], rf[ks][nt]);
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
            double yf[NTL];
#pragma unroll
            for (int mt = 0; mt < NTL; mt++) yf[mt] = Y[(8 * mt + gid) * LDY + kk];
#pragma unroll
            for (int mt = 0; mt < NTL; mt++)
#pragma unroll
              for (int qt = 0; qt < NTL; qt++) dmma(kacc[it][mt][qt][0], kacc[it][mt][qt][1], yf[mt], jf[ks][qt]);
          }
        }
      }
#pragma unroll
      for (int it = 0; it < NIT; it++) {
        const int n = nbase + 8 * it;
        if (n >= nn) break;
        double* On = Obase + (size_t)n * nel2 * np2;
#pragma unroll
        for (int mt = 0; mt < NTL; mt++)
#pragma unroll
          for (int qt = 0; qt < NTL; qt++)
#pragma unroll
            for (int cc = 0; cc < 2; cc++) {
              const int row = 8 * mt + gid, col = 8 * qt + 2 * t4 + cc;
              if (row < np && col < np) __stcs(On + row + col * np, kacc[it][mt][qt][cc]);
            }
      }
    }
  }
}

// x4.  K[j,k] -= sum over the batch's items with that (j,k) (several families) of the (at most four) sub-blocks that
// overlap each radial index pair; fixed order.  out_ptr/out_item: CSR over the distinct output blocks of the batch.
static __global__ void chan_x4_kernel(ChanXParams prm, const int* out_ptr, const int* out_item, int nout) {
  const int np = prm.np, np2 = prm.np2, nel = prm.nel, nel2 = nel * nel, nrad = prm.nrad;
  const size_t per = (size_t)nrad * nrad, total = per * nout;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int o = (int)(idx / per);
    const int a = (int)(idx % nrad), b = (int)((idx / nrad) % nrad);
    const int q0 = out_ptr[o], q1 = out_ptr[o + 1];
    const int2 jk = prm.items[out_item[q0]];
    double s = 0.0;
    for (int x = 0; x < 2; x++) {
      const int ie = prm.elem_of[2 * a + x];
      if (ie < 0) continue;
      for (int y = 0; y < 2; y++) {
        const int je = prm.elem_of[2 * b + y];
        if (je < 0) continue;
        const size_t off = (size_t)(ie * nel + je) * np2 + (a - prm.first[ie]) + (b - prm.first[je]) * np;
        for (int q = q0; q < q1; q++) s += prm.blocks[(size_t)(out_item[q] - prm.item0) * nel2 * np2 + off];
      }
    }
    prm.K[(size_t)(jk.x * nrad + a) + (size_t)(jk.y * nrad + b) * prm.ldk] -= s;
  }
}

}  // namespace hfb
