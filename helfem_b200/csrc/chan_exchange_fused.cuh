// sm_100a kernels: the right angular contraction of the exchange operator FUSED into its radial contraction
// (stages x2 + x3 of chan_exchange.cuh in one pass; atomic/TwoDBasis.cpp:1023-1134, diatomic/basis.cpp:1604-1726).
//
// The unfused pipeline wrote R^{ab}[j,k] = fac (+-) sum_{c in r} sum_l c^b_c(k,l) W^a_c[j,l] for every output block and
// family to HBM as zero-padded element-pair sub-blocks (110 GB at BASELINE configs[3]) and read it back once in the
// radial stage.  Here the CTA that contracts an element pair builds exactly the sub-blocks it needs from W in shared
// memory and consumes them in place; R never exists in global memory.
//
// Everything about a tile of <= 8 output blocks (j, k_0..k_7) of one family that depends only on the sparsity pattern of
// P -- the union of the W blocks (c, l) its rows couple to and the dense (slot x k x b) coefficient table -- is built once
// per pattern on the host (ChanEngine::build_items) instead of by every CTA:
//   tile_u[t]    = (first slot, slot count padded to a multiple of 4)
//   slot_wblk[s] = block of the batch's W buffer holding W_c[j,l]     (padding slots repeat a valid block)
//   slot_coef[s] = [2 stack halves b][8 items] coupling coefficients  (zero for padding slots / absent couplings)
// Phase 1 (both kernels) is a DMMA GEMM with M = the (np x np) radial sub-block of the element pair (A operand: the W
// sub-blocks, fetched by bulk copies through an mbarrier pipeline), N = the tile's items, K = the slots.
//
//   chan_xf_same_kernel<NI>       same element: B tile [NI items][np^2] in shared memory, then the GEMM with the K-ordered
//                                 in-element integrals (A operand pre-swizzled to fragment order, one 256 B load / fragment)
//   chan_xf_pair_kernel<NI,NTL>   ordered pairs of different elements: both stack halves b of NI items in shared memory,
//                                 then iint^a . Rsub^{ab} . (jint^b)^T as two chained DMMA products per item
// A group = consecutive tiles of ONE family whose items fill the N dimension densely (items of consecutive tiles are
// consecutive): int4 (first item, item count, first tile, tile count).
#pragma once
#include "chan_exchange.cuh"

namespace hfb {

constexpr int CHXF_TMAX = 16;  // tiles per group (a group closes at NI items or CHXF_TMAX tiles)
constexpr int CHXF_SLOTS = 8;  // W sub-blocks per pipeline stage (two DMMA k-steps); the host pads every tile's slot count to it
#ifndef HFB_XF_STAGES
#define HFB_XF_STAGES 3
#endif
constexpr int CHXF_STAGES = HFB_XF_STAGES;

// Rows of the (np x np) sub-block a CTA fetches per sweep (32 m-tiles of 8 rows), and the stage geometry: every slot of a
// stage owns PBS doubles, PBS = 4 (mod 16) so that the DMMA A-fragment read (slot 4 ks + t4, row 8 mt + gid) is
// bank-conflict free (and 16-byte aligned, as a bulk-copy destination must be).
__host__ __device__ inline int chxf_rows(int PB) { return PB < 256 ? PB : 256; }
__host__ __device__ inline int chxf_pbs(int PB) { return (chxf_rows(PB) - 4 + 15) / 16 * 16 + 4; }
__host__ __device__ inline size_t chxf_ring_doubles(int PB) { return (size_t)CHXF_STAGES * CHXF_SLOTS * chxf_pbs(PB); }

// ---- mbarrier / bulk-copy primitives (PTX) ----
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(a), "r"(parity) : "memory");
}
// global -> shared bulk copy (the TMA engine's 1-D form), completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   (unsigned)__cvta_generic_to_shared(dst)),
               "l"(src), "r"(bytes), "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}

// Pipeline state that survives from one phase-1 call of a kernel to the next (the barriers keep their phases).
struct ChxfPipe {
  unsigned long long* full;   // [CHXF_STAGES]  bytes of the stage have landed        (1 arrival + transaction bytes)
  unsigned long long* empty;  // [CHXF_STAGES]  all 8 warps are done with the stage   (8 arrivals)
  int c_stage = 0, p_stage = 0;
  unsigned c_phase = 0, p_phase = 0;
};
__device__ __forceinline__ void chxf_pipe_init(ChxfPipe& pipe, unsigned long long* bars) {
  pipe.full = bars;
  pipe.empty = bars + CHXF_STAGES;
  if (threadIdx.x == 0) {
    for (int s = 0; s < CHXF_STAGES; s++) { mbar_init(pipe.full + s, 1); mbar_init(pipe.empty + s, 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
}

// Phase 1 of the fused kernels for the tiles [t0, t0 + ntl) of a group, rows [k0, k0 + 256) of the sub-block `pair` = ie nel + je,
// stack half a of W (Wa = W + a WB):   R[row][item] = sum_slots W^a_slot[row] coef^b[slot][item]    (b = b0 .. b0 + NB - 1)
// as a DMMA GEMM (M = rows, N = the tile's 8 items, K = slots).  The A operand arrives through a ring of CHXF_STAGES stages
// of CHXF_SLOTS sub-blocks: thread 0 issues one BULK COPY (cp.async.bulk, the TMA engine) per sub-block -- W is stored
// element pair by element pair, so a sub-block is one contiguous run -- which signals the stage's `full` mbarrier by byte
// count; every warp arrives on the stage's `empty` mbarrier when it is done with it; the producer runs CHXF_STAGES - 1
// stages ahead ACROSS tiles and no block-wide barrier is involved.  Warp w owns the m-tiles w, w + 8, w + 16, w + 24;
// `emit(tt, mi, b, c, value)` receives R[row k0 + 8 (w + 8 mi) + gid][item 2 t4 + c] of tile tt after the tile's last stage.
template <int NB, class Emit>
__device__ __forceinline__ void chxf_phase1(const ChanXParams& prm, ChxfPipe& pipe, const double* __restrict__ Wa, int t0, int ntl, int pair,
                                            int k0, int b0, double* ring, Emit emit) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, t4 = lane & 3;
  const int PB = prm.PB, PBS = chxf_pbs(PB), rows = min(256, PB - k0);  // PB is a multiple of 8: whole m-tiles, 64-byte multiples
  const size_t wstride = (size_t)prm.stack * prm.WB;
  const double* src0 = Wa + (size_t)pair * PB + k0;
  const unsigned bytes = (unsigned)rows * 8u;
  // the ring may alias storage that was written through the generic proxy (the pair kernel's staging tiles): order
  // those writes before the async-proxy writes of the bulk copies, and make sure nobody still uses that storage
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  // producer cursor over (tile, chunk of CHXF_SLOTS slots); only thread 0 issues
  int p_tt = 0, p_s = 0, p_nu = 0, p_u = 0;
  if (ntl > 0) { const int2 U = prm.tile_u[t0]; p_u = U.x; p_nu = U.y; }
  auto produce = [&]() {
    if (tid == 0 && p_tt < ntl) {
      mbar_wait(pipe.empty + pipe.p_stage, pipe.p_phase ^ 1u);  // passes at once on the first trip round the ring
      mbar_expect_tx(pipe.full + pipe.p_stage, bytes * CHXF_SLOTS);
      double* dst = ring + (size_t)(pipe.p_stage * CHXF_SLOTS) * PBS;
#pragma unroll
      for (int q = 0; q < CHXF_SLOTS; q++) {
        const long long wb = __ldg(prm.slot_wblk + p_u + p_s + q);
        bulk_g2s(dst + (size_t)q * PBS, src0 + (size_t)wb * wstride, bytes, pipe.full + pipe.p_stage);
      }
      p_s += CHXF_SLOTS;
      if (p_s >= p_nu) {
        p_tt++;
        p_s = 0;
        if (p_tt < ntl) { const int2 U = prm.tile_u[t0 + p_tt]; p_u = U.x; p_nu = U.y; }
      }
      if (++pipe.p_stage == CHXF_STAGES) { pipe.p_stage = 0; pipe.p_phase ^= 1u; }
    }
  };
  for (int q = 0; q < CHXF_STAGES - 1; q++) produce();
  for (int tt = 0; tt < ntl; tt++) {
    const int2 U = prm.tile_u[t0 + tt];
    double acc[4][NB][2];
#pragma unroll
    for (int mi = 0; mi < 4; mi++)
#pragma unroll
      for (int b = 0; b < NB; b++) { acc[mi][b][0] = 0.0; acc[mi][b][1] = 0.0; }
    for (int s0 = 0; s0 < U.y; s0 += CHXF_SLOTS) {
      produce();
      double bf[CHXF_SLOTS / 4][NB];  // coefficient fragments: requested before the wait
#pragma unroll
      for (int ks = 0; ks < CHXF_SLOTS / 4; ks++)
#pragma unroll
        for (int b = 0; b < NB; b++) bf[ks][b] = __ldg(prm.slot_coef + (size_t)(U.x + s0 + 4 * ks + t4) * 16 + (b0 + b) * 8 + gid);
      mbar_wait(pipe.full + pipe.c_stage, pipe.c_phase);
      const double* S = ring + (size_t)(pipe.c_stage * CHXF_SLOTS) * PBS + gid;
#pragma unroll
      for (int ks = 0; ks < CHXF_SLOTS / 4; ks++) {
#pragma unroll
        for (int mi = 0; mi < 4; mi++) {
          if (8 * (warp + 8 * mi) < rows) {  // warp-uniform
            const double af = S[(4 * ks + t4) * PBS + 8 * (warp + 8 * mi)];
#pragma unroll
            for (int b = 0; b < NB; b++) dmma(acc[mi][b][0], acc[mi][b][1], af, bf[ks][b]);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(pipe.empty + pipe.c_stage);
      if (++pipe.c_stage == CHXF_STAGES) { pipe.c_stage = 0; pipe.c_phase ^= 1u; }
    }
#pragma unroll
    for (int mi = 0; mi < 4; mi++)
#pragma unroll
      for (int b = 0; b < NB; b++) {
        emit(tt, mi, b, 0, acc[mi][b][0]);
        emit(tt, mi, b, 1, acc[mi][b][1]);
      }
  }
}

// grid (groups of the batch, nel), 256 threads, dynamic smem NI * LDK + chxf_ring_doubles(PB) doubles.
template <int NI>
__global__ void __launch_bounds__(256, 2) chan_xf_same_kernel(ChanXParams prm, const int4* __restrict__ groups, const int2* __restrict__ tiles) {
  extern __shared__ __align__(128) double Bt[];  // [NI][LDK]: item n, index k = i + l np of its (np x np) sub-block; padding stays zero
  constexpr int NTN = NI / 8;
  const int nel = prm.nel, nel2 = nel * nel, np2 = prm.np2, st = prm.stack, st2 = st * st;
  const int e = blockIdx.y;
  const int4 G = groups[blockIdx.x];
  const int it0 = G.x, nn = G.y, t0 = G.z, ntl = G.w;
  const int r = prm.item_fam[it0];
  const double fac = prm.fam_fac[r];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, t4 = lane & 3;
  const int MT = prm.MT, KS = prm.KS, LDK = prm.LDK;
  __shared__ unsigned long long bars[2 * CHXF_STAGES];
  double* ring = Bt + (size_t)NI * LDK;
  for (int q = tid; q < NI * LDK; q += 256) Bt[q] = 0.0;
  ChxfPipe pipe;
  chxf_pipe_init(pipe, bars);

  double acc[4][NTN][2];
#pragma unroll
  for (int mi = 0; mi < 4; mi++)
#pragma unroll
    for (int nt = 0; nt < NTN; nt++) { acc[mi][nt][0] = 0.0; acc[mi][nt][1] = 0.0; }
  double* Obase = prm.blocks + ((size_t)(it0 - prm.item0) * nel2 + e * nel + e) * np2;

  for (int mt0 = 0; mt0 < MT; mt0 += 32) {  // 32 m-tiles per sweep (4 per warp); np2 <= 256 needs one sweep
    for (int ab = 0; ab < st2; ab++) {
      const int a = ab / st, b = ab % st;
      const double sg = (a == b) ? fac : -fac;
      __syncthreads();  // the previous product is done with the tile (first pass: the zero fill is complete)
      // ---- phase 1: Bt[n][k] = sg sum_slots W^a_slot[k] coef^b[slot][n] ----
      const double* Wa = prm.W + (size_t)a * prm.WB;
      for (int k0 = 0; k0 < np2; k0 += 256) {
        chxf_phase1<1>(prm, pipe, Wa, t0, ntl, e * nel + e, k0, b, ring, [&](int tt, int mi, int, int c, double v) {
          const int2 T = tiles[t0 + tt];
          const int row = k0 + 8 * (warp + 8 * mi) + gid, kk = 2 * t4 + c;
          if (row < np2 && kk < T.y) Bt[(T.x - it0 + kk) * LDK + row] = sg * v;  // rows outside the element are zero in W
        });
      }
      __syncthreads();
      // ---- phase 2: C[np2 x NI] += teiK^{ab} [np2 x np2] . Bt ----
      const double* A = prm.teiK + ((((size_t)r * nel + e) * st2 + ab) * MT) * KS * 32 + lane;
      const double* bs = Bt + gid * LDK + t4;
      // software pipeline: the A fragments of k-step ks + 1 are requested before the DMMAs of step ks are issued
      double af[4], an[4];
#pragma unroll
      for (int mi = 0; mi < 4; mi++) {
        const int mt = mt0 + warp + 8 * mi;
        af[mi] = (mt < MT) ? __ldg(A + (size_t)mt * KS * 32) : 0.0;
      }
#pragma unroll 2
      for (int ks = 0; ks < KS; ks++) {
        double bf[NTN];
#pragma unroll
        for (int mi = 0; mi < 4; mi++) {
          const int mt = mt0 + warp + 8 * mi;
          an[mi] = (mt < MT && ks + 1 < KS) ? __ldg(A + ((size_t)mt * KS + ks + 1) * 32) : 0.0;
        }
#pragma unroll
        for (int nt = 0; nt < NTN; nt++) bf[nt] = bs[8 * nt * LDK + 4 * ks];
#pragma unroll
        for (int mi = 0; mi < 4; mi++) {
          if (mt0 + warp + 8 * mi < MT) {  // warp-uniform: a missing m-tile (np2 = 225: 29 tiles on 32 slots) costs no DMMA
#pragma unroll
            for (int nt = 0; nt < NTN; nt++) dmma(acc[mi][nt][0], acc[mi][nt][1], af[mi], bf[nt]);
          }
        }
#pragma unroll
        for (int mi = 0; mi < 4; mi++) af[mi] = an[mi];
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
}

// Shared-memory geometry of the pair kernel: every (item, b) owns an (8 NTL x 8 NTL) matrix with row stride
// LDR = 4 (mod 16) doubles (bank-conflict-free B-fragment reads in the first product), element (i, l) at i + l LDR.
template <int NTL>
struct ChxfPair {
  static constexpr int LDR = (8 * NTL + 15) / 16 * 16 + 4;
  static constexpr int RS = 8 * NTL * LDR;
  static constexpr int LDY = LDR;
  static constexpr int YS = 8 * NTL * LDY;
};

// grid (groups of the batch, nel (nel - 1)), 256 threads, dynamic smem NI * stack * RS + max(chxf_ring_doubles(PB), 8 * YS)
// doubles (the per-warp staging of the first product aliases the cp.async ring of phase 1).
template <int NI, int NTL>
__global__ void __launch_bounds__(256, (NTL <= 2) ? 2 : 1) chan_xf_pair_kernel(ChanXParams prm, const int4* __restrict__ groups, const int2* __restrict__ tiles) {
  extern __shared__ __align__(128) double sm[];
  using Geo = ChxfPair<NTL>;
  constexpr int LDR = Geo::LDR, RS = Geo::RS, LDY = Geo::LDY;
  constexpr int NIT = (NI + 7) / 8;  // items a warp carries
  const int nel = prm.nel, nel2 = nel * nel, np = prm.np, np2 = prm.np2, st = prm.stack;
  const int ie = blockIdx.y / (nel - 1);
  const int jj = blockIdx.y % (nel - 1);
  const int je = jj + (jj >= ie ? 1 : 0);
  const int4 G = groups[blockIdx.x];
  const int it0 = G.x, nn = G.y, t0 = G.z, ntl = G.w;
  const int r = prm.item_fam[it0];
  const double fac = prm.fam_fac[r];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, t4 = lane & 3;
  __shared__ unsigned long long bars[2 * CHXF_STAGES];
  double* Rt = sm;                                            // [NI][st][RS]
  double* ring = sm + (size_t)NI * st * RS;                   // phase 1: cp.async stages
  double* Y = ring + warp * Geo::YS;                          // phase 2: per-warp staging of the first product (aliases the ring)
  const bool lower = ie > je;
  for (int q = tid; q < NI * st * RS; q += 256) Rt[q] = 0.0;  // padding rows / columns stay zero for good
  ChxfPipe pipe;
  chxf_pipe_init(pipe, bars);
  // where this lane's rows of a sweep land inside an item's (i, l) matrix (no integer division in the hot loop)
  int rofs[4];  // first sweep (all of the sub-block when np <= 16)
#pragma unroll
  for (int mi = 0; mi < 4; mi++) {
    const int row = 8 * (warp + 8 * mi) + gid;
    rofs[mi] = (row % np) + (row / np) * LDR;
  }

  double kacc[NIT][NTL][NTL][2];
#pragma unroll
  for (int it = 0; it < NIT; it++)
#pragma unroll
    for (int mt = 0; mt < NTL; mt++)
#pragma unroll
      for (int qt = 0; qt < NTL; qt++) { kacc[it][mt][qt][0] = 0.0; kacc[it][mt][qt][1] = 0.0; }

  for (int a = 0; a < st; a++) {
    __syncthreads();
    // ---- phase 1: Rt[n][b](i, l) = (+-) fac sum_slots W^a_slot(i, l) coef^b[slot][n], both stack halves b at once ----
    const double* Wa = prm.W + (size_t)a * prm.WB;
    for (int k0 = 0; k0 < np2; k0 += 256) {
      auto emit = [&](int tt, int mi, int b, int c, double v) {
        const int2 T = tiles[t0 + tt];
        const int kk = 2 * t4 + c;
        const int row = k0 + 8 * (warp + 8 * mi) + gid;
        if (row < np2 && kk < T.y)
          Rt[((size_t)(T.x - it0 + kk) * st + b) * RS + (k0 == 0 ? rofs[mi] : (row % np) + (row / np) * LDR)] = ((a == b) ? fac : -fac) * v;
      };
      if (st == 2) chxf_phase1<2>(prm, pipe, Wa, t0, ntl, ie * nel + je, k0, 0, ring, emit);
      else chxf_phase1<1>(prm, pipe, Wa, t0, ntl, ie * nel + je, k0, 0, ring, emit);
    }
    __syncthreads();
    // ---- phase 2: Ksub += iint^a . Rsub^{ab} . (jint^b)^T for the warp's items ----
    const double* iint = (lower ? prm.big_f : prm.small_f) + (((size_t)r * nel + ie) * st + a) * np2;
    double af[2 * NTL][NTL];
#pragma unroll
    for (int ks = 0; ks < 2 * NTL; ks++) {
      const int kk = 4 * ks + t4;
#pragma unroll
      for (int mt = 0; mt < NTL; mt++) af[ks][mt] = (kk < np && 8 * mt + gid < np) ? __ldg(iint + (8 * mt + gid) + kk * np) : 0.0;
    }
    for (int b = 0; b < st; b++) {
      const double* jint = (lower ? prm.small_f : prm.big_f) + (((size_t)r * nel + je) * st + b) * np2;
      double jf[2 * NTL][NTL];
#pragma unroll
      for (int ks = 0; ks < 2 * NTL; ks++) {
        const int kk = 4 * ks + t4;
#pragma unroll
        for (int mt = 0; mt < NTL; mt++) jf[ks][mt] = (kk < np && 8 * mt + gid < np) ? __ldg(jint + (8 * mt + gid) + kk * np) : 0.0;
      }
#pragma unroll
      for (int it = 0; it < NIT; it++) {
        const int n = warp + 8 * it;
        if (n >= nn) break;  // warp-uniform
        const double* Rn = Rt + ((size_t)n * st + b) * RS;
        // step 1: Y = iint . Rsub   (B[k][q] = Rsub[k, q]: rows 4 ks + t4, column 8 nt + gid)
        double y[NTL][NTL][2];
#pragma unroll
        for (int mt = 0; mt < NTL; mt++)
#pragma unroll
          for (int nt = 0; nt < NTL; nt++) { y[mt][nt][0] = 0.0; y[mt][nt][1] = 0.0; }
#pragma unroll
        for (int ks = 0; ks < 2 * NTL; ks++) {
          double rf[NTL];
#pragma unroll
          for (int nt = 0; nt < NTL; nt++) rf[nt] = Rn[4 * ks + t4 + (8 * nt + gid) * LDR];
#pragma unroll
          for (int mt = 0; mt < NTL; mt++)
#pragma unroll
            for (int nt = 0; nt < NTL; nt++) dmma(y[mt][nt][0], y[mt][nt][1], af[ks][mt], rf[nt]);
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
  }
#pragma unroll
  for (int it = 0; it < NIT; it++) {
    const int n = warp + 8 * it;
    if (n >= nn) break;
    double* On = prm.blocks + ((size_t)(it0 - prm.item0 + n) * nel2 + ie * nel + je) * np2;
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

}  // namespace hfb
