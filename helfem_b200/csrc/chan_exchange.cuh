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
//                     rows that share their partners stream every P block once), stored element pair by element pair
//   xf  (chan_exchange_fused.cuh) per element pair (ie, je), in ONE kernel:
//         R^{ab}[j,k] = fac_r (+-) sum_{c in r} sum_l c^b_c(k,l) W^a_c[j,l]   (right angular contraction) built in shared
//         memory from bulk copies of the W sub-blocks, then
//         Ksub(ie,je) = sum_ab T^{ab}_r(ie,je)[R^{ab}(ie,je)]      on the FP64 tensor pipe (DMMA.8x8x4):
//         ie == je:  (np^2 x np^2) K-ordered integrals x [np^2 x NI items]  -- a GEMM whose N dimension
//                    batches NI output blocks, integrals pre-swizzled into A-fragment order
//         ie != je:  iint . R . jint^T  as two chained (np x np) DMMA products per item
//   x4  K[j,k] -= the (at most four) sub-blocks that overlap each radial index pair   (fixed order)
// Zero blocks of P are skipped like the reference does (Frobenius norm < 10 eps, atomic/TwoDBasis.cpp:1036):
// the block mask of P fixes, once per sparsity pattern and on the host, the set L(c,j) of W blocks every coupling
// row produces (W is stored compactly, one block per member), the output blocks ("items") of every family, and a
// partition of the families into BATCHES whose W / output workspaces fit a fixed budget.  One launch of each stage
// covers a whole batch of families, so that the grid fills the 148 SMs even when a single family is small.
// When P is EXACTLY symmetric (what every SCF driver passes: P = C C^T) K[k,j] = K[j,k]^T, so only the output blocks
// j <= k are contracted and x4 writes both; any other matrix takes the full item list, like the reference.
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
  // family tables and the per-sparsity-pattern lists (all families; a launch covers a BATCH of families)
  const int* fam_c0;      // [nr+1] channel range of a family
  const double* fam_fac;  // [nr]
  const int2* items;      // [nitems] output blocks (j,k), grouped by family
  const int* item_fam;    // [nitems]
  // compact storage of W: row (c,j) holds one block per l of its set L(c,j) = { l : some i of the row has P[i,l] != 0 }
  const long long* wbase; // [nchan*nang] first block of the row inside the batch's W buffer
  const unsigned long long* wbits;  // [nchan*nang][nw] bitset of L(c,j)
  int nw;
  double* W;              // [blocks of the batch][stack][WB]: element-pair-major, sub-block (ie,je) at (ie*nel + je)*PB, entry (i, l) at i + l*np
  int PB, WB;             // doubles per sub-block (np2 rounded up to whole 8-row m-tiles) and per stack half of a block (nel*nel*PB)
  const unsigned* x1tab;  // [WB] offset A + B*ldp inside an angular block of the dense P of W index k, 0xffffffff for padding
  int item0;              // first item of the batch (blocks are indexed relative to it)
  double* blocks;         // [items of the batch][nel*nel][np2]
  double* K;              // dense, ld = ldk
  int ldk;
  // fused right-angular + radial contraction (chan_exchange_fused.cuh): per-tile slot tables built once per sparsity pattern
  const int2* tile_u;          // [tiles] (first slot, slot count padded to a multiple of 4)
  const long long* slot_wblk;  // [slots] block of the batch's W buffer
  const double* slot_coef;     // [slots][2 stack halves][8 items of the tile]
  int sym;                     // P is exactly symmetric: only the items j <= k exist, x4 also writes the transposed block
};

// Frobenius-norm mask of the angular blocks of P, and whether P is EXACTLY symmetric (checked on the non-zero blocks only:
// a zero block whose transposed partner is non-zero is caught from the partner's side).  grid (nang*nang), 128 threads.
// sym_flag: one byte, preset to 1, cleared on the first mismatch.
// Sharded run (nranks > 1): rank r scans the blocks b = r, r + nranks, ... and writes 0 / 1 (and a mismatch mark at [nang^2]) into the
// zeroed double array flagsD, which the engine all-reduces and chan_x0_flags_kernel turns back into the byte mask.
// nzx (optional, single rank) / flagsD[nang^2 + 1 ...] (sharded): the block has an entry that is not exactly zero -- the Coulomb pack
// of the same Fock build skips the others (exact; the norm mask above may also drop tiny non-zero blocks, as the reference's
// exchange does and its coulomb does not).
static __global__ void __launch_bounds__(128) chan_x0_mask_kernel(const double* P, int ldp, int nrad, int nang, unsigned char* nz,
                                                                  unsigned char* sym_flag, int rank = 0, int nranks = 1, double* flagsD = nullptr,
                                                                  unsigned char* nzx = nullptr) {
  __shared__ double red[4];
  __shared__ int s_nz;
  const int blk_id = blockIdx.x * nranks + rank;
  if (blk_id >= nang * nang) return;
  const int i = blk_id / nang, l = blk_id % nang;
  const double* blk = P + (size_t)i * nrad + (size_t)l * nrad * ldp;
  double s = 0.0;
  int any = 0;
  for (int k = threadIdx.x; k < nrad * nrad; k += blockDim.x) {
    const double v = blk[(k % nrad) + (size_t)(k / nrad) * ldp];
    s = fma(v, v, s);
    any |= (v != 0.0);
  }
  any = __syncthreads_or(any);
  if (threadIdx.x == 0) {
    if (flagsD) flagsD[nang * nang + 1 + blk_id] = any ? 1.0 : 0.0;
    else if (nzx) nzx[blk_id] = any ? 1 : 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    const double t = (red[0] + red[1]) + (red[2] + red[3]);
    const int f = (sqrt(t) >= 10.0 * 2.220446049250313e-16) ? 1 : 0;
    if (flagsD) flagsD[blk_id] = (double)f;
    else nz[blk_id] = (unsigned char)f;
    s_nz = f;
  }
  __syncthreads();
  if (!s_nz || (sym_flag == nullptr && flagsD == nullptr)) return;
  const double* tb = P + (size_t)l * nrad + (size_t)i * nrad * ldp;  // block (l, i)
  bool same = true;
  for (int k = threadIdx.x; k < nrad * nrad; k += blockDim.x) {
    const int A = k % nrad, B = k / nrad;
    same = same && (blk[A + (size_t)B * ldp] == tb[B + (size_t)A * ldp]);
  }
  if (!same) {
    if (flagsD) flagsD[nang * nang] = 1.0;
    else *sym_flag = 0;
  }
}
static __global__ void __launch_bounds__(256) chan_x0_flags_kernel(const double* flagsD, int n, unsigned char* nz, int allow_sym, unsigned char* nzx) {
  const int q = blockIdx.x * 256 + threadIdx.x;
  if (q < n) {
    nz[q] = flagsD[q] != 0.0 ? 1 : 0;
    nzx[q] = flagsD[n + 1 + q] != 0.0 ? 1 : 0;
  } else if (q == n) nz[n] = (allow_sym && flagsD[n] == 0.0) ? 1 : 0;
}

// Tiling of the two angular contractions: a CTA owns up to CHX_T rows that share their coupled partners (same m),
// builds the UNION of the partners in shared memory (flags -> ballot scan -> slots) with a dense (slot x row)
// coefficient table, and streams every input block ONCE for all its rows instead of once per row.
constexpr int CHX_T = 8;      // rows j of an x1 tile / items (j, k_0..) of a tile of the fused kernels
#ifndef HFB_CHX_UMAX
#define HFB_CHX_UMAX 64
#endif
constexpr int CHX_UMAX = HFB_CHX_UMAX;  // union slots per pass (a multiple of 4; -DHFB_CHX_UMAX=4 exercises the multi-pass accumulation, tools/gpu_umax_variant.sh)
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
// On the FP64 tensor pipe: W_{(c,j)}[l][k] = sum_v coef[u_v][j][b] P_{i_v l}[k] as a GEMM with M = the radial pairs k, enumerated ELEMENT PAIR BY
// ELEMENT PAIR (zero-padded np x np sub-blocks, a boundary function shared by two elements appears in both: x1tab; the fused kernels then
// fetch a whole sub-block with one bulk copy) (A operand read from the dense P in fragment order), N = the tile's <= 8 rows j (one n-tile per stack half b), K = the active partners of this l.
// dynamic smem: coef[2][UMAX+1][CHX_LDC] doubles (slot UMAX stays zero: padding) | wblk[8][T] long long | ulist[nang] | dom[nang] |
//               act[8][UMAX+4] | actl[8][UMAX+4] ints | done[nang] | cur[nang] bytes
static __global__ void __launch_bounds__(256, 3) chan_x1_mma_kernel(ChanXParams prm, const int* tiles) {
  extern __shared__ double sm[];
  __shared__ int s_cnt, s_na[8];
  const int nang = prm.nang, nrad = prm.nrad, WB = prm.WB, st = prm.stack;
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
      const bool has0 = wb0 >= 0, has1 = wb1 >= 0;
      double* const Wp0 = prm.W + (size_t)(has0 ? wb0 : 0) * st * WB + gid;  // this lane's rows 2 t4, 2 t4 + 1 of the tile, W index gid
      double* const Wp1 = prm.W + (size_t)(has1 ? wb1 : 0) * st * WB + gid;
      for (int k0 = warp * 32; k0 < WB; k0 += 256) {
        const int kg = k0 + gid;
        unsigned so[4];  // element offsets inside the dense P fit 32 bits (ldp^2 < 2^32 is checked by the engine)
        bool ok[4];
#pragma unroll
        for (int mt = 0; mt < 4; mt++) {
          const unsigned tb = __ldg(prm.x1tab + min(kg + 8 * mt, WB - 1));
          ok[mt] = tb != 0xffffffffu;  // padding rows of a sub-block are written as zeros
          so[mt] = (ok[mt] ? tb : 0u) + (unsigned)(l * nrad) * (unsigned)prm.ldp;
        }
        double acc[4][2][2];
#pragma unroll
        for (int mt = 0; mt < 4; mt++) { acc[mt][0][0] = acc[mt][0][1] = acc[mt][1][0] = acc[mt][1][1] = 0.0; }
#pragma unroll 2
        for (int v0 = 0; v0 < na; v0 += 4) {
          const int ro = actl[v0 + t4], u = act[v0 + t4];
          double af[4];
#pragma unroll
          for (int mt = 0; mt < 4; mt++) af[mt] = ok[mt] ? prm.P[so[mt] + (unsigned)ro] : 0.0;
          const double b0 = coef[u * CHX_LDC + gid], b1 = coef[CB + u * CHX_LDC + gid];
#pragma unroll
          for (int mt = 0; mt < 4; mt++) {
            dmma(acc[mt][0][0], acc[mt][0][1], af[mt], b0);
            dmma(acc[mt][1][0], acc[mt][1][1], af[mt], b1);
          }
        }
        // stores (streaming: W is far larger than L2 and is next read by the fused kernels).  wb < 0: l is not in the set of the row
        // (its partners' blocks are all zero), nothing stored.  WB and k0 are multiples of 8, so the range test is warp-uniform.
        if (!prev) {  // block-uniform: the common single-pass case, straight predicated stores
#pragma unroll
          for (int mt = 0; mt < 4; mt++) {
            if (k0 + 8 * mt >= WB) break;
            if (has0) {
              __stcs(Wp0 + k0 + 8 * mt, acc[mt][0][0]);
              if (st == 2) __stcs(Wp0 + k0 + 8 * mt + WB, acc[mt][1][0]);
            }
            if (has1) {
              __stcs(Wp1 + k0 + 8 * mt, acc[mt][0][1]);
              if (st == 2) __stcs(Wp1 + k0 + 8 * mt + WB, acc[mt][1][1]);
            }
          }
        } else {
#pragma unroll
          for (int mt = 0; mt < 4; mt++) {
            if (k0 + 8 * mt >= WB) break;
#pragma unroll
            for (int i = 0; i < 2; i++) {
              if (!(i ? has1 : has0)) continue;
              double* W0 = (i ? Wp1 : Wp0) + k0 + 8 * mt;
              double v0 = acc[mt][0][i] + W0[0], v1 = acc[mt][1][i];
              if (st == 2) v1 += W0[WB];
              __stcs(W0, v0);
              if (st == 2) __stcs(W0 + WB, v1);
            }
          }
        }
      }
      }  // w
    }
    __syncthreads();
    for (int q = tid; q < nang; q += 256) { done[q] |= cur[q]; cur[q] = 0; }
  }
}

// x4.  K[j,k] -= sum over the batch's items with that (j,k) (several families) of the (at most four) sub-blocks that
// overlap each radial index pair; fixed order.  out_ptr/out_item: CSR over the distinct output blocks of the batch.
// One CTA per output block at a time, 32 x 32 tiles of the (nrad x nrad) block; with prm.sym the transposed block
// K[k,j] = K[j,k]^T is written from the same tile through shared memory (both writes coalesced).  256 threads.
static __global__ void __launch_bounds__(256) chan_x4_kernel(ChanXParams prm, const int* out_ptr, const int* out_item, int nout) {
  __shared__ double tile[32][33];
  const int np = prm.np, np2 = prm.np2, nel = prm.nel, nel2 = nel * nel, nrad = prm.nrad;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int ntile = (nrad + 31) / 32;
  for (int o = blockIdx.x; o < nout; o += gridDim.x) {
    const int q0 = out_ptr[o], q1 = out_ptr[o + 1];
    const int2 jk = prm.items[out_item[q0]];
    const bool mirror = prm.sym && jk.x != jk.y;
    for (int tb = 0; tb < ntile; tb++)
      for (int ta = 0; ta < ntile; ta++) {
        const int a = ta * 32 + tx;
        // this thread's four elements (a, b_v), b_v = 32 tb + ty + 8 v.  Every element has a first home (the sub-block of the
        // first elements its row and column belong to); rows / columns on an element boundary have up to three more.  One
        // pass over the items with the four first-home loads independent of each other, the rare extra homes afterwards.
        int off0[4];
        bool extra = false;
#pragma unroll
        for (int v = 0; v < 4; v++) {
          const int b = tb * 32 + ty + 8 * v;
          off0[v] = -1;
          if (a < nrad && b < nrad) {
            const int ie = prm.elem_of[2 * a], je = prm.elem_of[2 * b];
            off0[v] = (ie * nel + je) * np2 + (a - prm.first[ie]) + (b - prm.first[je]) * np;
            extra = extra || prm.elem_of[2 * a + 1] >= 0 || prm.elem_of[2 * b + 1] >= 0;
          }
        }
        double s[4] = {0.0, 0.0, 0.0, 0.0};
        const size_t istride = (size_t)nel2 * np2;
        int q = q0;
        for (; q + 4 <= q1; q += 4) {  // four items at a time: 16 independent loads in flight, added in item order
          const double* B0 = prm.blocks + (size_t)(out_item[q] - prm.item0) * istride;
          const double* B1 = prm.blocks + (size_t)(out_item[q + 1] - prm.item0) * istride;
          const double* B2 = prm.blocks + (size_t)(out_item[q + 2] - prm.item0) * istride;
          const double* B3 = prm.blocks + (size_t)(out_item[q + 3] - prm.item0) * istride;
          double t[4][4];
#pragma unroll
          for (int v = 0; v < 4; v++) {
            const bool ok = off0[v] >= 0;
            const int o = ok ? off0[v] : 0;
            t[v][0] = ok ? __ldcs(B0 + o) : 0.0;
            t[v][1] = ok ? __ldcs(B1 + o) : 0.0;
            t[v][2] = ok ? __ldcs(B2 + o) : 0.0;
            t[v][3] = ok ? __ldcs(B3 + o) : 0.0;
          }
#pragma unroll
          for (int v = 0; v < 4; v++) s[v] = (((s[v] + t[v][0]) + t[v][1]) + t[v][2]) + t[v][3];
        }
        for (; q < q1; q++) {
          const double* Bq = prm.blocks + (size_t)(out_item[q] - prm.item0) * istride;
#pragma unroll
          for (int v = 0; v < 4; v++)
            if (off0[v] >= 0) s[v] += __ldcs(Bq + off0[v]);
        }
        if (extra) {
#pragma unroll
          for (int v = 0; v < 4; v++) {
            const int b = tb * 32 + ty + 8 * v;
            if (a >= nrad || b >= nrad) continue;
            for (int h = 1; h < 4; h++) {
              const int ie = prm.elem_of[2 * a + (h >> 1)], je = prm.elem_of[2 * b + (h & 1)];
              if (ie < 0 || je < 0) continue;
              const int o = (ie * nel + je) * np2 + (a - prm.first[ie]) + (b - prm.first[je]) * np;
              for (int q = q0; q < q1; q++) s[v] += prm.blocks[(size_t)(out_item[q] - prm.item0) * istride + o];
            }
          }
        }
#pragma unroll
        for (int v = 0; v < 4; v++) {
          const int b = tb * 32 + ty + 8 * v;
          if (a < nrad && b < nrad) prm.K[(size_t)(jk.x * nrad + a) + (size_t)(jk.y * nrad + b) * prm.ldk] -= s[v];
          tile[ty + 8 * v][tx] = s[v];
        }
        if (mirror) {  // block-uniform
          __syncthreads();
          const int bb = tb * 32 + tx;
          for (int yy = ty; yy < 32; yy += 8) {
            const int aa = ta * 32 + yy;
            if (aa < nrad && bb < nrad) prm.K[(size_t)(jk.y * nrad + bb) + (size_t)(jk.x * nrad + aa) * prm.ldk] -= tile[tx][yy];
          }
          __syncthreads();
        }
      }
  }
}

// x4, shared-memory form (used when an (nrad x nrad) block fits): one CTA per output block at a time.  The items' sub-blocks are
// read SUB-BLOCK BY SUB-BLOCK -- each is one contiguous run of np^2 doubles per item, so every 32-byte sector that is fetched is
// used (the tiled form above cuts the runs at tile and element borders and fetched ~2x the algorithmic bytes, ncu r02s) -- summed
// over the items in item order and added into the block held in shared memory (a boundary function belongs to two elements: the
// sub-blocks overlap there, hence the barrier between sub-blocks and the fixed sub-block order); the finished block and, with
// prm.sym, its transpose are then subtracted from K with coalesced accesses.  256 threads, dynamic smem nrad * (nrad + 1) doubles.
static __global__ void __launch_bounds__(256) chan_x4_smem_kernel(ChanXParams prm, const int* out_ptr, const int* out_item, int nout) {
  extern __shared__ double blk[];  // [nrad][nrad + 1]: entry (a, b) at a + b * (nrad + 1)
  const int np = prm.np, np2 = prm.np2, nel = prm.nel, nel2 = nel * nel, nrad = prm.nrad, ldb = nrad + 1;
  const int tid = threadIdx.x;
  const size_t istride = (size_t)nel2 * np2;
  const int li = tid % np, ll = tid / np;  // this thread's entry of a sub-block (np2 <= 256 is checked by the launcher)
  for (int o = blockIdx.x; o < nout; o += gridDim.x) {
    const int q0 = out_ptr[o], q1 = out_ptr[o + 1];
    const int2 jk = prm.items[out_item[q0]];
    const bool mirror = prm.sym && jk.x != jk.y;
    for (int e = tid; e < nrad * ldb; e += 256) blk[e] = 0.0;
    __syncthreads();
    for (int pe = 0; pe < nel2; pe++) {
      const int ie = pe / nel, je = pe % nel;
      const int ni = prm.npe[ie], nj = prm.npe[je];
      if (tid < np2 && li < ni && ll < nj) {
        const double* B = prm.blocks + (size_t)pe * np2 + tid;
        double s = 0.0;
        int q = q0;
        for (; q + 8 <= q1; q += 8) {  // eight items in flight, added in item order
          double t[8];
#pragma unroll
          for (int u = 0; u < 8; u++) t[u] = __ldcs(B + (size_t)(out_item[q + u] - prm.item0) * istride);
#pragma unroll
          for (int u = 0; u < 8; u++) s += t[u];
        }
        for (; q < q1; q++) s += __ldcs(B + (size_t)(out_item[q] - prm.item0) * istride);
        blk[(prm.first[ie] + li) + (prm.first[je] + ll) * ldb] += s;
      }
      __syncthreads();
    }
    double* Kjk = prm.K + (size_t)jk.x * nrad + (size_t)jk.y * nrad * prm.ldk;
    for (int e = tid; e < nrad * nrad; e += 256) {
      const int a = e % nrad, b = e / nrad;
      Kjk[a + (size_t)b * prm.ldk] -= blk[a + b * ldb];
    }
    if (mirror) {
      double* Kkj = prm.K + (size_t)jk.y * nrad + (size_t)jk.x * nrad * prm.ldk;
      for (int e = tid; e < nrad * nrad; e += 256) {
        const int b = e % nrad, a = e / nrad;  // K[k,j](b, a) = K[j,k](a, b): coalesced in b, conflict-free in shared memory (ldb odd or even + 1)
        Kkj[b + (size_t)a * prm.ldk] -= blk[a + b * ldb];
      }
    }
    __syncthreads();
  }
}

// Packed copies of a list of (nrad x nrad) angular blocks (j,k) of a dense matrix: the payload of the exchange all-reduce of a
// sharded run.  grid (blocks), 256 threads.  mirror: also write block (k,j) = block(j,k)^T (symmetric operator, j <= k listed).
static __global__ void __launch_bounds__(256) chan_blocks_gather_kernel(const double* M, int ld, int nrad, const int2* blk, double* packed) {
  const int2 jk = blk[blockIdx.x];
  const double* src = M + (size_t)jk.x * nrad + (size_t)jk.y * nrad * ld;
  double* dst = packed + (size_t)blockIdx.x * nrad * nrad;
  for (int e = threadIdx.x; e < nrad * nrad; e += 256) dst[e] = src[(e % nrad) + (size_t)(e / nrad) * ld];
}
static __global__ void __launch_bounds__(256) chan_blocks_scatter_kernel(double* M, int ld, int nrad, const int2* blk, const double* packed, int mirror) {
  const int2 jk = blk[blockIdx.x];
  const double* src = packed + (size_t)blockIdx.x * nrad * nrad;
  double* dst = M + (size_t)jk.x * nrad + (size_t)jk.y * nrad * ld;
  double* dstT = M + (size_t)jk.y * nrad + (size_t)jk.x * nrad * ld;
  const bool mir = mirror && jk.x != jk.y;
  for (int e = threadIdx.x; e < nrad * nrad; e += 256) {
    const int a = e % nrad, b = e / nrad;
    const double v = src[e];
    dst[a + (size_t)b * ld] = v;
    if (mir) dstT[b + (size_t)a * ld] = v;
  }
}

}  // namespace hfb
