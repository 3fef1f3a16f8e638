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
// Debug only (-DHFB_XF_TRACE, tools/xf_trace.py): phase timeline of the same-element kernel's CTAs on the first SMs.
#ifdef HFB_XF_TRACE
__device__ unsigned long long g_trace[148 * 4096];
__device__ int g_trace_n[148];
__device__ __forceinline__ void xf_trace(int ev, unsigned long long payload = 0) {
  if (threadIdx.x != 0) return;
  unsigned smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  if (smid >= 4) return;
  unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  if (ev == 5) t = payload;
  const int i = atomicAdd(&g_trace_n[smid], 1);
  if (i < 4096) g_trace[smid * 4096 + i] = (t << 20) | ((unsigned long long)(blockIdx.x + gridDim.x * blockIdx.y) & 0xffff) << 4 | (unsigned)ev;
}
#else
#define xf_trace(...)
#endif

constexpr int CHXF_TMAX = 16;  // tiles per group (a group closes at NI items or CHXF_TMAX tiles)
#ifndef HFB_XF_SLOTS
#define HFB_XF_SLOTS 4
#endif
#ifndef HFB_XF_STAGES
#define HFB_XF_STAGES 6
#endif
constexpr int CHXF_SLOTS = HFB_XF_SLOTS;    // W sub-blocks per pipeline stage (a multiple of 4 = one DMMA k-step); the host pads every tile's slot count to it
constexpr int CHXF_STAGES = HFB_XF_STAGES;  // stages of the bulk-copy ring
constexpr int CHXF_THREADS = 288;           // 8 consumer warps + 1 producer warp

// Rows of the (np x np) sub-block a CTA fetches per sweep (32 m-tiles of 8 rows), and the stage geometry: every slot of a
// stage owns PBS doubles, PBS = 4 (mod 16) so that the DMMA A-fragment read (slot 4 ks + t4, row 8 mt + gid) is
// bank-conflict free (and 16-byte aligned, as a bulk-copy destination must be).
__host__ __device__ inline int chxf_rows(int PB) { return PB < 256 ? PB : 256; }
__host__ __device__ inline int chxf_pbs(int PB) { return (chxf_rows(PB) - 4 + 15) / 16 * 16 + 4; }
// A stage = CHXF_SLOTS sub-blocks of PBS doubles + the stage's coupling coefficients (CHXF_SLOTS x [2 stack halves][8 items]):
// both arrive by bulk copy and complete the same `full` barrier, so no consumer ever waits on a global load of its own.
#ifndef HFB_XF_PSTAGES
#define HFB_XF_PSTAGES 5
#endif
constexpr int CHXF_PSTAGES = HFB_XF_PSTAGES;  // ring depth of the pair kernel (its shared memory also holds the unit's integral tables)
__host__ __device__ inline int chxf_stage_doubles(int PB) { return CHXF_SLOTS * (chxf_pbs(PB) + 16); }
__host__ __device__ inline size_t chxf_ring_doubles(int PB) { return (size_t)CHXF_PSTAGES * chxf_stage_doubles(PB); }
__host__ __device__ inline size_t chxf_inline_ring_doubles(int PB) { return (size_t)CHXF_STAGES * chxf_stage_doubles(PB); }

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
// the same wait for the producer warp, which has nothing else to do: a suspend-time hint keeps its polling out of the issue slots
__device__ __forceinline__ void mbar_wait_relaxed(unsigned long long* bar, unsigned parity) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(a), "r"(parity), "r"(20000u) : "memory");
}
// global -> shared bulk copy (the TMA engine's 1-D form), completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   (unsigned)__cvta_generic_to_shared(dst)),
               "l"(src), "r"(bytes), "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
// 8-byte asynchronous copy global -> shared (the integral tables of a pair unit: their rows are only 8-byte aligned)
__device__ __forceinline__ void cp_async8(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
// disjoint-factor tables of a pair unit in shared memory: [stack halves a] iint^a, then [stack halves b] jint^b, TBL doubles each
__host__ __device__ inline int chxf_pair_tbl(int np2) { return (np2 + 1) & ~1; }
__host__ __device__ inline size_t chxf_pair_table_doubles(int np2, int stack) { return (size_t)2 * stack * chxf_pair_tbl(np2); }
// barrier of the 8 consumer warps only (the producer warp never joins it)
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// The fused kernels are PERSISTENT and WARP-SPECIALISED.  A CTA walks through work units u = blockIdx.x, + gridDim.x, ...
// (unit = group of tiles x element pair).  Warp 8 is the PRODUCER: one lane issues, for every unit, stack half and tile in
// the order the consumers will need them, one BULK COPY (cp.async.bulk, the TMA engine) per W sub-block -- W is stored
// element pair by element pair, so a sub-block is one contiguous run -- into a ring of CHXF_STAGES stages of CHXF_SLOTS
// sub-blocks; a stage's `full` mbarrier completes by byte count, its `empty` mbarrier when the 8 consumer warps have
// arrived.  The producer is only ever held back by `empty`, so it runs ahead ACROSS phases and units: while the
// consumers multiply (phase 2) the first stages of their next angular contraction (phase 1) are already landing.
struct ChxfPipe {
  unsigned long long* full;   // [CHXF_PSTAGES]  bytes of the stage have landed        (1 arrival + transaction bytes)
  unsigned long long* empty;  // [CHXF_PSTAGES]  all 8 consumer warps are done with it (8 arrivals)
  int stage = 0;
  unsigned phase = 0;
  __device__ __forceinline__ void advance() {
    if (++stage == CHXF_PSTAGES) { stage = 0; phase ^= 1u; }
  }
};
__device__ __forceinline__ void chxf_pipe_init(ChxfPipe& pipe, unsigned long long* bars) {
  pipe.full = bars;
  pipe.empty = bars + CHXF_PSTAGES;
  if (threadIdx.x == 0) {
    for (int s = 0; s < CHXF_PSTAGES; s++) { mbar_init(pipe.full + s, 1); mbar_init(pipe.empty + s, 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
}

// Producer side of one phase-1 pass over the slots [s_begin, s_end) (the slots of a group's tiles are consecutive and every
// tile's count is a multiple of CHXF_SLOTS, so a stage never straddles tiles), rows [k0, k0 + 256) of sub-block `pair`, stack
// half a (Wa = W + a WB).  Called by the whole producer WARP: lane q < CHXF_SLOTS fetches slot q of every stage; the block
// number of the next stage's slot is requested before the wait on the current stage's `empty` barrier.
__device__ __forceinline__ void chxf_produce(const ChanXParams& prm, ChxfPipe& pipe, const double* __restrict__ Wa, int s_begin, int s_end,
                                             int pair, int k0, double* ring) {
  const int lane = threadIdx.x & 31;
  const int PB = prm.PB, PBS = chxf_pbs(PB), SD = chxf_stage_doubles(PB), rows = min(256, PB - k0);  // PB is a multiple of 8: 64-byte multiples
  const size_t wstride = (size_t)prm.stack * prm.WB;
  const double* src0 = Wa + (size_t)pair * PB + k0;
  const unsigned bytes = (unsigned)rows * 8u;
  long long wb = (lane < CHXF_SLOTS && s_begin < s_end) ? __ldg(prm.slot_wblk + s_begin + lane) : 0;
  for (int s0 = s_begin; s0 < s_end; s0 += CHXF_SLOTS) {
    const long long wb_next = (lane < CHXF_SLOTS && s0 + CHXF_SLOTS < s_end) ? __ldg(prm.slot_wblk + s0 + CHXF_SLOTS + lane) : 0;
    if (lane == 0) {
      mbar_wait_relaxed(pipe.empty + pipe.stage, pipe.phase ^ 1u);  // passes at once on the first trip round the ring
      mbar_expect_tx(pipe.full + pipe.stage, bytes * CHXF_SLOTS + CHXF_SLOTS * 128u);
    }
    __syncwarp();
    double* dst = ring + (size_t)pipe.stage * SD;
    if (lane < CHXF_SLOTS) bulk_g2s(dst + (size_t)lane * PBS, src0 + (size_t)wb * wstride, bytes, pipe.full + pipe.stage);
    else if (lane == CHXF_SLOTS) bulk_g2s(dst + (size_t)CHXF_SLOTS * PBS, prm.slot_coef + (size_t)s0 * 16, CHXF_SLOTS * 128u, pipe.full + pipe.stage);
    wb = wb_next;
    pipe.advance();
  }
}

// Consumer side of the same pass (all 8 consumer warps):   R[row][item] = sum_slots W^a_slot[row] coef^b[slot][item]
// (b = b0 .. b0 + NB - 1) as a DMMA GEMM (M = rows, N = the tile's 8 items, K = slots).  Warp w owns the m-tiles w, w + 8,
// w + 16, w + 24; `emit(tt, acc)` receives acc[mi][b][c] = R[row k0 + 8 (w + 8 mi) + gid][item 2 t4 + c] of tile tt after the
// tile's last stage.  sU: the tiles' (first slot, slot count).
template <int NB, class Emit>
__device__ __forceinline__ void chxf_consume(const ChanXParams& prm, ChxfPipe& pipe, const int2* sU, int ntl, int k0, int b0, const double* ring, Emit emit) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, t4 = lane & 3;
  const int PB = prm.PB, PBS = chxf_pbs(PB), SD = chxf_stage_doubles(PB), rows = min(256, PB - k0);
  for (int tt = 0; tt < ntl; tt++) {
    const int2 U = sU[tt];  // (first slot, slot count) of the tile, staged in shared memory by the caller
    double acc[4][NB][2];
#pragma unroll
    for (int mi = 0; mi < 4; mi++)
#pragma unroll
      for (int b = 0; b < NB; b++) { acc[mi][b][0] = 0.0; acc[mi][b][1] = 0.0; }
    for (int s0 = 0; s0 < U.y; s0 += CHXF_SLOTS) {
      mbar_wait(pipe.full + pipe.stage, pipe.phase);
      const double* S = ring + (size_t)pipe.stage * SD + gid;
      double bf[CHXF_SLOTS / 4][NB];  // coefficient fragments of the stage (they came with it)
#pragma unroll
      for (int ks = 0; ks < CHXF_SLOTS / 4; ks++)
#pragma unroll
        for (int b = 0; b < NB; b++) bf[ks][b] = S[CHXF_SLOTS * PBS + (4 * ks + t4) * 16 + (b0 + b) * 8];
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
      if (lane == 0) mbar_arrive(pipe.empty + pipe.stage);
      pipe.advance();
    }
    emit(tt, acc);
  }
}

// ---- same-element kernel: one CTA per (group, element), 256 threads, no producer warp (the kernel needs 128 registers per
//      thread for its 4 x 4 accumulator tiles, which a ninth warp at two CTAs per SM does not leave) ----
// A single thread needs ~0.3 us to issue one stage (mbarrier wait + expect_tx + five bulk copies, measured with
// tools/micro/bulk_copy_bench.cu: ~0.2 us fixed + ~40 ns per copy), far more than the 8 warps need to consume it, so the
// issuing is STATELESS and ROTATES over the warps: the stages a CTA consumes form one flat sequence g = 0 .. G-1 (sweeps x stack
// pairs ab x row chunks k0 x stages; the slots of a group are consecutive, so stage j of a pass is slots 4j .. 4j+3 of the
// group), everything about stage g follows from g alone, and lane 0 of warp (g mod 8) issues it CHXF_STAGES - 1 stages ahead
// of its own consumption, in any order relative to the other warps.  The sequence runs across the stack pairs: the first
// stages of the next angular contraction land during the current radial product (phase 2).
struct ChxfPipeInline {
  unsigned long long* full;   // [CHXF_STAGES]  bytes of the stage have landed        (1 arrival + transaction bytes)
  unsigned long long* empty;  // [CHXF_STAGES]  all 8 warps are done with the stage   (8 arrivals)
  int g = 0;                  // next stage this warp consumes
  int c_stage = 0;            // = g mod CHXF_STAGES
  unsigned c_phase = 0;       // = (g / CHXF_STAGES) & 1
};
__device__ __forceinline__ void chxf_pipe_inline_init(ChxfPipeInline& pipe, unsigned long long* bars) {
  pipe.full = bars;
  pipe.empty = bars + CHXF_STAGES;
  if (threadIdx.x == 0) {
    for (int s = 0; s < CHXF_STAGES; s++) { mbar_init(pipe.full + s, 1); mbar_init(pipe.empty + s, 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
}
constexpr int CHXF_SWB = 512;  // block numbers of a group's slots kept in shared memory
// Issue stage g (one thread).  nst: stages of one pass over the group's slots; passes: (sweep, ab, k0 chunk), k0 fastest.
__device__ __forceinline__ void chxf_same_issue(const ChanXParams& prm, const ChxfPipeInline& pipe, int g, int nst, const int* sWb, int slot0, int pair,
                                                double* ring) {
  const int PB = prm.PB, PBS = chxf_pbs(PB), st = prm.stack, nk0 = (prm.np2 + 255) >> 8;
  const int j = g % nst;
  int q = g / nst;
  const int k0 = (q % nk0) << 8;
  q /= nk0;
  const int a = (q % (st * st)) / st;
  const int rs = g % CHXF_STAGES;
  const unsigned use = (unsigned)(g / CHXF_STAGES);
  const unsigned bytes = (unsigned)min(256, PB - k0) * 8u;
  const size_t wstride = (size_t)st * prm.WB;
  const double* src0 = prm.W + (size_t)a * prm.WB + (size_t)pair * PB + k0;
  mbar_wait(pipe.empty + rs, (use & 1u) ^ 1u);  // the consumers are done with the previous use of the ring slot (first use: passes at once)
  mbar_expect_tx(pipe.full + rs, bytes * CHXF_SLOTS + CHXF_SLOTS * 128u);
  double* dst = ring + (size_t)rs * chxf_stage_doubles(PB);
#pragma unroll
  for (int c = 0; c < CHXF_SLOTS; c++) {
    const int rel = CHXF_SLOTS * j + c;  // the group's block numbers are staged in shared memory (sWb) as far as they fit
    const long long wb = rel < CHXF_SWB ? (long long)sWb[rel] : __ldg(prm.slot_wblk + slot0 + rel);
    bulk_g2s(dst + (size_t)c * PBS, src0 + (size_t)wb * wstride, bytes, pipe.full + rs);
  }
  bulk_g2s(dst + (size_t)CHXF_SLOTS * PBS, prm.slot_coef + (size_t)(slot0 + CHXF_SLOTS * j) * 16, CHXF_SLOTS * 128u, pipe.full + rs);
}

// Phase 1 of the same-element kernel for the tiles of a group (sU: their (first slot, slot count)), rows [k0, k0 + 256) of the
// sub-block `pair`:   R[row][item] = sum_slots W^a_slot[row] coef^b[slot][item]    (b = b0 .. b0 + NB - 1)
// as a DMMA GEMM (M = rows, N = the tile's 8 items, K = slots).  The A operand and the coefficients arrive through the ring
// (chxf_same_issue: bulk copies, the TMA engine's 1-D form; W is stored element pair by element pair, so a sub-block is one
// contiguous run) which signals a stage's `full` mbarrier by byte count; every warp arrives on the stage's `empty` mbarrier when
// it is done with it; no block-wide barrier is involved.  Warp w owns the m-tiles w, w + 8, w + 16, w + 24;
// `emit(tt, acc)` receives acc[mi][b][c] = R[row k0 + 8 (w + 8 mi) + gid][item 2 t4 + c] of tile tt.  G: stages of the CTA.
template <int NB, class Emit>
__device__ __forceinline__ void chxf_phase1_inline(const ChanXParams& prm, ChxfPipeInline& pipe, const int2* sU, const int* sWb, int slot0, int nst, int G,
                                                   int ntl, int pair, int k0, int b0, double* ring, Emit emit) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, t4 = lane & 3;
  const int PB = prm.PB, PBS = chxf_pbs(PB), SD = chxf_stage_doubles(PB), rows = min(256, PB - k0);  // PB is a multiple of 8: whole m-tiles
  for (int tt = 0; tt < ntl; tt++) {
    const int nu = sU[tt].y;
    double acc[4][NB][2];
#pragma unroll
    for (int mi = 0; mi < 4; mi++)
#pragma unroll
      for (int b = 0; b < NB; b++) { acc[mi][b][0] = 0.0; acc[mi][b][1] = 0.0; }
    for (int s0 = 0; s0 < nu; s0 += CHXF_SLOTS) {
      {
        const int gp = pipe.g + CHXF_STAGES - 1;
        if (lane == 0 && (gp & 7) == warp && gp < G) chxf_same_issue(prm, pipe, gp, nst, sWb, slot0, pair, ring);
      }
      mbar_wait(pipe.full + pipe.c_stage, pipe.c_phase);
      const double* S = ring + (size_t)pipe.c_stage * SD + gid;
      double bf[CHXF_SLOTS / 4][NB];  // coefficient fragments of the stage (they came with it)
#pragma unroll
      for (int ks = 0; ks < CHXF_SLOTS / 4; ks++)
#pragma unroll
        for (int b = 0; b < NB; b++) bf[ks][b] = S[CHXF_SLOTS * PBS + (4 * ks + t4) * 16 + (b0 + b) * 8];
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
      pipe.g++;
      if (++pipe.c_stage == CHXF_STAGES) { pipe.c_stage = 0; pipe.c_phase ^= 1u; }
    }
    emit(tt, acc);
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
  ChxfPipeInline pipe;
  __shared__ int2 sT[CHXF_TMAX], sU[CHXF_TMAX];  // the group's tiles: (first item, item count), (first slot, slot count)
  __shared__ int sWb[CHXF_SWB];                  // W block of the group's slots (the slots of a group are consecutive)
  const int2 U0 = prm.tile_u[t0], U1 = prm.tile_u[t0 + ntl - 1];
  const int slot0 = U0.x, nslot = U1.x + U1.y - U0.x;
  const int nst = nslot / CHXF_SLOTS;                                   // stages of one pass over the group's slots
  const int GS = nst * ((np2 + 255) >> 8) * st2 * ((MT + 31) / 32);    // stages this CTA consumes
  if (tid < ntl) { sT[tid] = tiles[t0 + tid]; sU[tid] = prm.tile_u[t0 + tid]; }
  for (int q = tid; q < min(nslot, CHXF_SWB); q += 256) sWb[q] = (int)__ldg(prm.slot_wblk + slot0 + q);
  xf_trace(5, (unsigned long long)(ntl * 100000 + nslot));
  chxf_pipe_inline_init(pipe, bars);  // (block-wide barrier: the tables are visible)
  if (lane == 0 && warp < CHXF_STAGES - 1 && warp < GS) chxf_same_issue(prm, pipe, warp, nst, sWb, slot0, e * nel + e, ring);

  double acc[4][NTN][2];
#pragma unroll
  for (int mi = 0; mi < 4; mi++)
#pragma unroll
    for (int nt = 0; nt < NTN; nt++) { acc[mi][nt][0] = 0.0; acc[mi][nt][1] = 0.0; }
  double* Obase = prm.blocks + ((size_t)(it0 - prm.item0) * nel2 + e * nel + e) * np2;

  for (int mt0 = 0; mt0 < MT; mt0 += 32) {  // 32 m-tiles per sweep (4 per warp); np2 <= 256 needs one sweep
    for (int ab = 0; ab < st2; ab++) {
      const int b = ab % st;
      const double sg = (ab / st == b) ? fac : -fac;
      __syncthreads();  // the previous product is done with the tile (first pass: the zero fill is complete)
      xf_trace(1);
      // ---- phase 1: Bt[n][k] = sg sum_slots W^a_slot[k] coef^b[slot][n] ----
      for (int k0 = 0; k0 < np2; k0 += 256) {
        chxf_phase1_inline<1>(prm, pipe, sU, sWb, slot0, nst, GS, ntl, e * nel + e, k0, b, ring, [&](int tt, double (&ra)[4][1][2]) {
          const int2 T = sT[tt];
          double* Bn = Bt + (T.x - it0 + 2 * t4) * LDK + k0 + 8 * warp + gid;  // rows outside the element are zero in W
          const bool ok0 = 2 * t4 < T.y, ok1 = 2 * t4 + 1 < T.y;
#pragma unroll
          for (int mi = 0; mi < 4; mi++)
            if (k0 + 8 * (warp + 8 * mi) + gid < np2) {
              if (ok0) Bn[64 * mi] = sg * ra[mi][0][0];
              if (ok1) Bn[64 * mi + LDK] = sg * ra[mi][0][1];
            }
        });
      }
      xf_trace(2);
      __syncthreads();
      xf_trace(3);
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
#ifdef HFB_XF_SKIP_P2
      const int KS_run = 1;  // timing experiment only (wrong results): what the kernel costs without its radial product
#else
      const int KS_run = KS;
#endif
#pragma unroll 2
      for (int ks = 0; ks < KS_run; ks++) {
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
      xf_trace(4);
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

// persistent grid, CHXF_THREADS threads, dynamic smem NI * stack * RS + 8 * YS + chxf_ring_doubles(PB) + chxf_pair_table_doubles doubles.
// Units: (group, ordered pair of different elements).
template <int NI, int NTL>
__global__ void __launch_bounds__(CHXF_THREADS, (NTL <= 2) ? 2 : 1) chan_xf_pair_kernel(ChanXParams prm, const int4* __restrict__ groups, int ngroups,
                                                                                        const int2* __restrict__ tiles) {
  extern __shared__ __align__(128) double sm[];
  __shared__ unsigned long long bars[2 * CHXF_PSTAGES];
  __shared__ int2 sT[CHXF_TMAX], sU[CHXF_TMAX];
  using Geo = ChxfPair<NTL>;
  constexpr int LDR = Geo::LDR, RS = Geo::RS, LDY = Geo::LDY;
  constexpr int NIT = (NI + 7) / 8;  // items a warp carries
  const int nel = prm.nel, nel2 = nel * nel, np = prm.np, np2 = prm.np2, st = prm.stack, npairs = nel * (nel - 1);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, t4 = lane & 3;
  double* Rt = sm;                                                   // [NI][st][RS]
  double* Y = sm + (size_t)NI * st * RS + (warp & 7) * Geo::YS;      // per-warp staging of the first product
  double* ring = sm + (size_t)NI * st * RS + (size_t)8 * Geo::YS;    // bulk-copy stages
  double* sI = ring + chxf_ring_doubles(prm.PB);                     // the unit's disjoint-factor tables (cp.async by the consumers)
  const int TBL = chxf_pair_tbl(np2);
  for (int q = tid; q < NI * st * RS; q += CHXF_THREADS) Rt[q] = 0.0;  // padding rows / columns stay zero for good
  ChxfPipe pipe;
  chxf_pipe_init(pipe, bars);
  const int nunits = ngroups * npairs;

  if (warp == 8) {  // ---------------- producer warp ----------------
    // slot range of a unit's group: first slot of its first tile .. end of its last tile (requested one unit ahead)
    auto slot_range = [&](int u, int& sb, int& se, int& pair) {
      const int4 G = groups[u / npairs];
      const int pq = u % npairs, ie = pq / (nel - 1), jj = pq % (nel - 1), je = jj + (jj >= ie ? 1 : 0);
      const int2 U0 = prm.tile_u[G.z], U1 = prm.tile_u[G.z + G.w - 1];
      sb = U0.x; se = U1.x + U1.y; pair = ie * nel + je;
    };
    int sb = 0, se = 0, pair = 0;
    if ((int)blockIdx.x < nunits) slot_range(blockIdx.x, sb, se, pair);
    for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
      int sb_n = 0, se_n = 0, pair_n = 0;
      if (u + (int)gridDim.x < nunits) slot_range(u + gridDim.x, sb_n, se_n, pair_n);
      for (int a = 0; a < st; a++)
        for (int k0 = 0; k0 < np2; k0 += 256) chxf_produce(prm, pipe, prm.W + (size_t)a * prm.WB, sb, se, pair, k0, ring);
      sb = sb_n; se = se_n; pair = pair_n;
    }
    return;
  }
  // ---------------- consumer warps ----------------
  // where this lane's rows of the first sweep land inside an item's (i, l) matrix (no integer division in the hot loop)
  int rofs[4];
#pragma unroll
  for (int mi = 0; mi < 4; mi++) {
    const int row = 8 * (warp + 8 * mi) + gid;
    rofs[mi] = (row % np) + (row / np) * LDR;
  }
  for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
    const int4 G = groups[u / npairs];
    const int pq = u % npairs, ie = pq / (nel - 1), jj = pq % (nel - 1), je = jj + (jj >= ie ? 1 : 0);
    const int it0 = G.x, nn = G.y, t0 = G.z, ntl = G.w;
    const int r = prm.item_fam[it0];
    const double fac = prm.fam_fac[r];
    const bool lower = ie > je;
    // the group's tile tables (item range, slot range): staged once per unit; the barrier at the top of the a loop
    // publishes them (and the previous unit's last readers are behind its phase-1 / phase-2 barrier)
    if (tid < ntl) { sT[tid] = tiles[t0 + tid]; sU[tid] = prm.tile_u[t0 + tid]; }
    double kacc[NIT][NTL][NTL][2];
#pragma unroll
    for (int it = 0; it < NIT; it++)
#pragma unroll
      for (int mt = 0; mt < NTL; mt++)
#pragma unroll
        for (int qt = 0; qt < NTL; qt++) { kacc[it][mt][qt][0] = 0.0; kacc[it][mt][qt][1] = 0.0; }

    for (int a = 0; a < st; a++) {
      consumer_sync();  // the previous products are done with Rt (and with the previous unit's tables)
      xf_trace(6);
      if (a == 0) {
        // this unit's disjoint factors iint^a (element ie), jint^b (element je): requested now, they land during phase 1
        const double* ib = (lower ? prm.big_f : prm.small_f) + ((size_t)r * nel + ie) * st * np2;
        const double* jb = (lower ? prm.small_f : prm.big_f) + ((size_t)r * nel + je) * st * np2;
        for (int q = tid; q < st * np2; q += 256) {  // both stack halves of a factor are one contiguous run
          cp_async8(sI + q, ib + q);
          cp_async8(sI + st * TBL + q, jb + q);
        }
        cp_async_commit();
      }
      // ---- phase 1: Rt[n][b](i, l) = (+-) fac sum_slots W^a_slot(i, l) coef^b[slot][n], both stack halves b at once ----
      for (int k0 = 0; k0 < np2; k0 += 256) {
        auto store = [&](int tt, int mi, int b, double v0, double v1) {
          const int2 T = sT[tt];
          const int row = k0 + 8 * (warp + 8 * mi) + gid;
          if (row >= np2) return;
          int ro = rofs[mi];
          if constexpr (NTL > 2) {  // several sweeps (np > 16): the later ones pay the integer division
            if (k0 != 0) ro = (row % np) + (row / np) * LDR;
          }
          double* Rn = Rt + ((T.x - it0 + 2 * t4) * st + b) * RS + ro;
          const double sg = (a == b) ? fac : -fac;
          if (2 * t4 < T.y) Rn[0] = sg * v0;
          if (2 * t4 + 1 < T.y) Rn[st * RS] = sg * v1;
        };
        auto emit2 = [&](int tt, double (&ra)[4][2][2]) {
#pragma unroll
          for (int mi = 0; mi < 4; mi++)
#pragma unroll
            for (int b = 0; b < 2; b++) store(tt, mi, b, ra[mi][b][0], ra[mi][b][1]);
        };
        auto emit1 = [&](int tt, double (&ra)[4][1][2]) {
#pragma unroll
          for (int mi = 0; mi < 4; mi++) store(tt, mi, 0, ra[mi][0][0], ra[mi][0][1]);
        };
        if (st == 2) chxf_consume<2>(prm, pipe, sU, ntl, k0, 0, ring, emit2);
        else chxf_consume<1>(prm, pipe, sU, ntl, k0, 0, ring, emit1);
      }
      if (a == 0) cp_async_wait_all();
      xf_trace(7);
      consumer_sync();
      xf_trace(8);
    // ---- phase 2: Ksub += iint^a . Rsub^{ab} . (jint^b)^T for the warp's items ----
    const double* iint = sI + a * np2;
    double af[2 * NTL][NTL];
#pragma unroll
    for (int ks = 0; ks < 2 * NTL; ks++) {
      const int kk = 4 * ks + t4;
#pragma unroll
      for (int mt = 0; mt < NTL; mt++) af[ks][mt] = (kk < np && 8 * mt + gid < np) ? iint[(8 * mt + gid) + kk * np] : 0.0;
    }
    for (int b = 0; b < st; b++) {
      const double* jint = sI + st * TBL + b * np2;
      double jf[2 * NTL][NTL];
#pragma unroll
      for (int ks = 0; ks < 2 * NTL; ks++) {
        const int kk = 4 * ks + t4;
#pragma unroll
        for (int mt = 0; mt < NTL; mt++) jf[ks][mt] = (kk < np && 8 * mt + gid < np) ? jint[(8 * mt + gid) + kk * np] : 0.0;
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
  }  // units
}

}  // namespace hfb
