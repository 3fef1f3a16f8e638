// sm_100a kernel for DFTGrid::eval_Fxc of the spherically averaged atom: one WARP per
// (Fock build, radial element), every dense product on the FP64 tensor pipe (DMMA.8x8x4).
//
// Fuses, per (build, element), the reference's per-element worker chain
//   compute_bf      sadatom/dftgrid.cpp:842-863   (tables are device-resident constants)
//   update_density  sadatom/dftgrid.cpp:46-120 (R), :122-227 (U)
//   compute_xc      sadatom/dftgrid.cpp:306-486   (device functionals, xc_point.cuh)
//   eval_Exc / Nel  sadatom/dftgrid.cpp:228-238,566-572
//   eval_Fxc        sadatom/dftgrid.cpp:610-680 (R), :682-805 (U)
//
// A CTA owns one element: its B / dB tables (np_pad x ldq, zero padded) are staged once in
// shared memory and shared by the CTA's warps, each of which carries its own build through
// the stages with warp-level synchronisation only:
//   0. the (spin, l) slices of the element block of P are streamed from HBM with fully coalesced
//      loads (all slices of 8 column groups in flight), summed over l and parked in a per-warp
//      shared block (row stride ldp = 4 mod 16); when the Coulomb kernel follows in the same
//      Fock build, the triangle-folded total density and the two disjoint traces it needs
//      (sadatom_coulomb_mma.cuh) are emitted from that block, so P is read from HBM ONCE;
//   1. Q^T = B^T P  (points x functions; A operand = table, B operand = density block),
//      contracted with B / dB in the accumulator layout -> rho, d rho, tau per point (one
//      quad shuffle per point tile, fixed order => deterministic);
//   2. XC on the points (lanes stride the points), potentials folded into quadrature weights;
//   3. H = B (a B + g dB)^T + dB (g B + c dB)^T  [+ Hl = B dl B^T], upper tile triangle only,
//      mirrored into the per-warp shared block and written to every l slice as H + l(l+1) Hl
//      with fully coalesced stores.
// Table row stride ldq = 4 (mod 16) doubles makes all fragment access patterns
// shared-memory-bank-conflict free.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include "xc_point.cuh"

namespace hfb {

struct SadGridParams {
  // constant tables
  const double* Bt;    // [nel][np_pad][ldq]  B_i(r_p)/r     (zero for i >= np_e, p >= nq)
  const double* dBt;   // [nel][np_pad][ldq]  d/dr
  const double* r;     // [nel][nq]
  const double* wrad;  // [nel][nq]
  int nel, nq, nq8, ldq, np_max, np_pad, nslices, nb;
  int x_func, c_func;
  double dens_thr;
  // per batch (element-band layout, see include/helfem_b200.h)
  const double* P;     // [nb][nspin][nslices][nel][np_max^2]
  double* H;           // same shape
  double* partial;     // [nb][nel][4]  Exc, Nel, tau integral, unused
  // optional by-product for the Coulomb kernel that follows (null: not wanted), sadatom_coulomb_mma.cuh
  double* pc;          // [nb][nel][jKP]  triangle-folded total density
  double* traces;      // [nb][nel][2]    Lfac tr(dL P), Lfac tr(dm1L P)
  const double* dLc;   // [nel][jMP]
  const double* dm1Lc; // [nel][jMP]
  const int2* tri;     // [ntri]
  int ntri, jKP, jMP;
  int gpc;             // build groups (of blockDim/32 builds) a CTA walks through with its tables staged once
};

// tuning knobs (tools/build_variant.py): builds (= warps) per CTA, resident CTAs per SM the register budget is cut for
#ifndef HFB_GRID_WARPS
#define HFB_GRID_WARPS 8
#endif
// default: 3 CTAs (24 warps) per SM where the kernel fits 80 registers without spilling -- the compile-time
// LDA / GGA pairs, unpolarised, up to 16 functions per element -- else 2
#ifndef HFB_GRID_MINBLOCKS
#define HFB_GRID_MINBLOCKS ((FAM < xc::FAM_MGGA && NSPIN == 1 && NT <= 2 && PAIR == 1) ? 3 : 2)
#endif

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  // not volatile: the scheduler may interleave independent accumulator chains
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ double quad_sum(double v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  return v;
}

/// row stride of the per-warp density / Fock block: >= np_pad and = 4 (mod 16)
__host__ __device__ constexpr int sad_grid_ldp(int np_pad) { return np_pad <= 20 ? 20 : 36; }

/// The functional pair a family is usually run with: kernels specialised on it (PAIR = 1) fold the functional
/// switch at compile time (a third of the code, half the registers of the XC stage); any other pair of the
/// family takes the PAIR = 0 instantiation that switches on prm.x_func / prm.c_func at run time.
template <int FAM> struct CanonicalPair {
  static constexpr int x = (FAM == xc::FAM_LDA) ? xc::XC_LDA_X : (FAM == xc::FAM_GGA ? xc::XC_GGA_X_PBE : xc::XC_MGGA_X_R2SCAN);
  static constexpr int c = (FAM == xc::FAM_LDA) ? xc::XC_LDA_C_PW : (FAM == xc::FAM_GGA ? xc::XC_GGA_C_PBE : xc::XC_MGGA_C_R2SCAN);
};

// FAM: highest functional family; NSPIN: 1 or 2; NT: 8-row tiles of the element block (np_pad = 8 NT);
// PAIR: 1 = functional ids are CanonicalPair<FAM>, 0 = read from prm.
template <int FAM, int NSPIN, int NT, int PAIR>
__global__ void __launch_bounds__(32 * HFB_GRID_WARPS, HFB_GRID_MINBLOCKS) sad_grid_mma_kernel(SadGridParams prm) {
  extern __shared__ double smem[];
  constexpr bool GGA = (FAM >= xc::FAM_GGA);
  constexpr bool MGGA = (FAM >= xc::FAM_MGGA);
  constexpr int KS = 2 * NT;                   // k-steps over the functions of an element
  constexpr int NARR = 1 + (GGA ? 1 : 0) + (MGGA ? 2 : 0);  // per-spin per-point arrays
  constexpr int LDP = sad_grid_ldp(8 * NT);
  constexpr int PBLK = 8 * NT * LDP;           // one density / Fock block
  constexpr int NBLK = MGGA ? 2 : 1;           // + the l(l+1)-weighted one
  // coalesced block transfers: LW lanes walk down a column, 32/LW columns per pass
  constexpr int LW = (NT == 1) ? 8 : (NT == 2 ? 16 : 32);
  constexpr int CPI = 32 / LW;
  constexpr int NPASS = 8 * NT / CPI;
  const int el = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  const int gid = lane >> 2, t4 = lane & 3;
  const int np = prm.np_max, np2 = np * np, nq = prm.nq, nq8 = prm.nq8, ldq = prm.ldq;
  const int tabsz = 8 * NT * ldq;

  double* sB = smem;
  double* sdB = sB + tabsz;  // only touched if GGA
  // Coulomb by-product constants of this element: disjoint_L | disjoint_m1L (triangle packed) | (i,j) of m
  double* sDL = sB + (GGA ? 2 : 1) * tabsz;
  double* sDM = sDL + prm.jKP;
  int2* sTri = reinterpret_cast<int2*>(sDM + prm.jKP);
  double* sR = sDM + 2 * prm.jKP;  // quadrature radii | weights of this element
  double* sW = sR + nq8;
  // per warp: point arrays [NSPIN][NARR][nq8] | per-point Exc, Nel, tau integrands [3][nq8] | density / Fock block(s)
#ifdef HFB_GRID_ALIAS_SPT  // experiment: integrand arrays share the density / Fock block (needs 3 nq8 <= NBLK PBLK)
  const int wstride = NSPIN * NARR * nq8 + NBLK * PBLK;
  double* wbase = sDL + 3 * prm.jKP + 2 * nq8;
  double* scratch = wbase + (size_t)warp * wstride;
  double* spt = scratch + NSPIN * NARR * nq8;
  double* pblk = spt;
#else
  const int wstride = (NSPIN * NARR + 3) * nq8 + NBLK * PBLK;
  double* wbase = sDL + 3 * prm.jKP + 2 * nq8;
  double* scratch = wbase + (size_t)warp * wstride;
  double* spt = scratch + NSPIN * NARR * nq8;
  double* pblk = spt + 3 * nq8;
#endif

#ifndef HFB_GRID_NO_ENTRY_PREFETCH
  // ---- start pulling this warp's density blocks out of HBM while the CTA stages its tables ----
  {
    const int b0 = (int)blockIdx.y * prm.gpc * nwarps + warp;
    if (b0 < prm.nb) {
      const int nblk = NSPIN * prm.nslices;
      const char* blk0 = reinterpret_cast<const char*>(prm.P + ((size_t)b0 * nblk * prm.nel + el) * np2);
      const int nline = (np2 * 8 + 127) >> 7;
      for (int k = lane; k < nblk * nline; k += 32) {
        const int blk = k / nline, ln = k - blk * nline;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(blk0 + (size_t)blk * prm.nel * np2 * 8 + (size_t)ln * 128));
      }
    }
  }
#endif
  // ---- stage the element tables (whole CTA) ----
  {
    const double* gB = prm.Bt + (size_t)el * tabsz;
    for (int k = tid; k < tabsz; k += blockDim.x) sB[k] = gB[k];
    if (GGA) {
      const double* gdB = prm.dBt + (size_t)el * tabsz;
      for (int k = tid; k < tabsz; k += blockDim.x) sdB[k] = gdB[k];
    }
    if (prm.pc != nullptr) {
      for (int m = tid; m < prm.jKP; m += blockDim.x) {
        const bool ok = m < prm.ntri;
        sDL[m] = ok ? prm.dLc[(size_t)el * prm.jMP + m] : 0.0;
        sDM[m] = ok ? prm.dm1Lc[(size_t)el * prm.jMP + m] : 0.0;
        sTri[m] = ok ? prm.tri[m] : make_int2(0, 0);
      }
    }
    for (int p = tid; p < nq; p += blockDim.x) {
      sR[p] = prm.r[(size_t)el * nq + p];
      sW[p] = prm.wrad[(size_t)el * nq + p];
    }
  }
  __syncthreads();

#ifdef HFB_GRID_PROFILE
  long long tk[8]; int tki = 0;
#define HFB_TICK() do { if (tki < 8) tk[tki++] = clock64(); } while (0)
#else
#define HFB_TICK()
#endif
  HFB_TICK();
  const size_t slice_stride = (size_t)prm.nel * np2;
  const int MT = nq8 >> 3;
  const int brow = lane & (LW - 1), bcol0 = lane / LW;  // this lane's row / first column in block transfers
  const bool brow_ok = brow < np;

  // ---- the CTA walks through prm.gpc groups of nwarps builds (one build per warp) ----
#pragma unroll 1
  for (int gi = 0; gi < prm.gpc; gi++) {
  const int bbase = ((int)blockIdx.y * prm.gpc + gi) * nwarps;
  if (bbase >= prm.nb) break;  // block-uniform
  const int b = bbase + warp;
  const bool active = b < prm.nb;  // idle warps of the last group only keep the block barriers company
  // pull the density blocks of this warp's NEXT build towards L2 while this one is worked on
#ifdef HFB_GRID_PREFETCH
  if (gi + 1 < prm.gpc && b + nwarps < prm.nb) {
    const int nblk = NSPIN * prm.nslices;
    const char* nxt = reinterpret_cast<const char*>(prm.P + ((size_t)(b + nwarps) * nblk * prm.nel + el) * np2);
    const int nline = (np2 * 8 + 127) >> 7;
    for (int k = lane; k < nblk * nline; k += 32) {
      const int blk = k / nline, ln = k - blk * nline;
      asm volatile("prefetch.global.L2 [%0];" ::"l"(nxt + (size_t)blk * slice_stride * 8 + (size_t)ln * 128));
    }
  }
#endif

  // ================= stages 0 + 1: density, gradient, tau on the points =================
#pragma unroll 1
  for (int s = 0; s < (active ? NSPIN : 0); s++) {
    const double* gP = prm.P + ((size_t)(b * NSPIN + s) * prm.nslices * prm.nel + el) * np2;
    double* aR = scratch + (size_t)s * NARR * nq8;  // rho | g | t1 | t2
    // ---- stage 0: sum the l slices of the block into shared memory ----
    {
      double acc[NPASS], accl[MGGA ? NPASS : 1];
#pragma unroll
      for (int it = 0; it < NPASS; it++) { acc[it] = 0.0; if (MGGA) accl[it] = 0.0; }
      for (int l = 0; l < prm.nslices; l++) {
        const double* src = gP + (size_t)l * slice_stride + brow;
        double v[NPASS];
#pragma unroll
        for (int it = 0; it < NPASS; it++) {
          const int col = it * CPI + bcol0;
          v[it] = (brow_ok && col < np) ? __ldg(src + col * np) : 0.0;
        }
        const double ll = (double)(l * (l + 1));
#pragma unroll
        for (int it = 0; it < NPASS; it++) {
          acc[it] += v[it];
          if (MGGA) accl[it] = fma(ll, v[it], accl[it]);
        }
      }
      __syncwarp();  // previous users of the block (stage 1 of the other spin) are done
      if (brow < 8 * NT) {
#pragma unroll
        for (int it = 0; it < NPASS; it++) {
          const int col = it * CPI + bcol0;
          pblk[col * LDP + brow] = acc[it];
          if (MGGA) pblk[PBLK + col * LDP + brow] = accl[it];
        }
      }
      __syncwarp();
    }
    if (s == 0) { HFB_TICK(); }
    // ---- by-product: folded total density and disjoint traces for the Coulomb kernel ----
    if (prm.pc != nullptr) {
      double* gpc = prm.pc + ((size_t)b * prm.nel + el) * prm.jKP;
      double ts = 0.0, tb = 0.0;
      for (int m = lane; m < prm.jKP; m += 32) {
        const int2 ij = sTri[m];
        double v = 0.0;
        if (m < prm.ntri) {
          v = pblk[ij.y * LDP + ij.x];
          if (ij.x != ij.y) v += pblk[ij.x * LDP + ij.y];
          if (NSPIN == 2 && s > 0) v += gpc[m];
        }
        ts = fma(sDL[m], v, ts);
        tb = fma(sDM[m], v, tb);
        gpc[m] = v;
      }
      if (s == NSPIN - 1) {
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
    }
    if (s == 0) { HFB_TICK(); }
    // ---- stage 1: B operand fragments Bop[k][n] = P(n, k), n = 8 ni + gid, k = 4 ks + t4 ----
    // held in registers across the point tiles when they fit, else re-read from the shared block per use
    constexpr bool PREG = (NT * KS * (MGGA ? 2 : 1) <= 16);
    double pf[PREG ? NT : 1][PREG ? KS : 1], plf[(PREG && MGGA) ? NT : 1][(PREG && MGGA) ? KS : 1];
    const double* pfs = pblk + t4 * LDP + gid;
    if (PREG) {
#pragma unroll
      for (int ni = 0; ni < NT; ni++)
#pragma unroll
        for (int ks = 0; ks < KS; ks++) {
          pf[ni][ks] = pfs[4 * ks * LDP + 8 * ni];
          if (MGGA) plf[ni][ks] = pfs[PBLK + 4 * ks * LDP + 8 * ni];
        }
    }
#pragma unroll 2
    for (int mi = 0; mi < MT; mi++) {
      const double* ab = sB + (size_t)t4 * ldq + 8 * mi + gid;
      const double* ad = sdB + (size_t)t4 * ldq + 8 * mi + gid;
      double av[KS], adv[MGGA ? KS : 1];
#pragma unroll
      for (int ks = 0; ks < KS; ks++) {
        av[ks] = ab[(size_t)4 * ks * ldq];
        if (MGGA) adv[ks] = ad[(size_t)4 * ks * ldq];
      }
      double rho = 0.0, g = 0.0, t1 = 0.0, t2 = 0.0;
#pragma unroll
      for (int ni = 0; ni < NT; ni++) {
        double q0 = 0.0, q1 = 0.0, qp0 = 0.0, qp1 = 0.0, ql0 = 0.0, ql1 = 0.0;
#pragma unroll
        for (int ks = 0; ks < KS; ks++) {
          const double pv = PREG ? pf[PREG ? ni : 0][PREG ? ks : 0] : pfs[4 * ks * LDP + 8 * ni];
          dmma(q0, q1, av[ks], pv);
          if (MGGA) {
            dmma(qp0, qp1, adv[ks], pv);
            dmma(ql0, ql1, av[ks], (PREG && MGGA) ? plf[PREG ? ni : 0][PREG ? ks : 0] : pfs[PBLK + 4 * ks * LDP + 8 * ni]);
          }
        }
        const double* mb0 = sB + (size_t)(8 * ni + 2 * t4) * ldq + gid + 8 * mi;
        const double b0 = mb0[0], b1 = mb0[ldq];
        rho = fma(q1, b1, fma(q0, b0, rho));
        if (GGA) {
          const double* md0 = sdB + (size_t)(8 * ni + 2 * t4) * ldq + gid + 8 * mi;
          const double d0 = md0[0], d1 = md0[ldq];
          g = fma(q1, d1, fma(q0, d0, g));
          if (MGGA) {
            t1 = fma(qp1, d1, fma(qp0, d0, t1));
            t2 = fma(ql1, b1, fma(ql0, b0, t2));
          }
        }
      }
      rho = quad_sum(rho);
      if (GGA) g = quad_sum(g);
      if (MGGA) { t1 = quad_sum(t1); t2 = quad_sum(t2); }
      if (t4 == 0) {
        const int p = 8 * mi + gid;
        aR[p] = rho;
        if (GGA) aR[nq8 + p] = g;
        if (MGGA) { aR[2 * nq8 + p] = t1; aR[3 * nq8 + p] = t2; }
      }
    }
  }
#ifndef HFB_GRID_NO_POOL
  __syncthreads();
#else
  __syncwarp();
#endif
  HFB_TICK();

  // ================= stage 2: XC on the points of ALL the CTA's builds, pooled over its threads =================
  // (nq is rarely a multiple of 32: per-warp passes would leave a quarter of the FP64 lanes idle in the last one)
  {
    const int xf = PAIR ? CanonicalPair<FAM>::x : prm.x_func;
    const int cf = PAIR ? CanonicalPair<FAM>::c : prm.c_func;
    const xc::XcThresholds thr = xc::make_thresholds(prm.dens_thr);
#ifdef HFB_GRID_NO_POOL  // experiment: every warp evaluates its own build's points, no block barriers
    for (int p = lane; p < (active ? nq : 0); p += 32) {
      double* a0 = scratch;
      double* a1 = a0 + (size_t)NARR * nq8;
      double* sp = spt;
      const double rr = sR[p];
      const double wr = sW[p];
      const double w4 = 4.0 * HFB_PI * wr;
      const double wtot = w4 * (rr * rr);
#else
    const int nbw = min(nwarps, prm.nb - bbase);  // builds of this group
    const int npts = nbw * nq, nthr = blockDim.x;
    // F full passes of the whole CTA, then the remainder on the LAST warps only; just the warps whose builds the
    // remainder touches (and the ones running it) wait for it, on their own named barrier
    const int F = npts / nthr, R = npts - F * nthr;
    const int wrem = nwarps - ((R + 31) >> 5);
    const int wmin = R ? min((F * nthr) / nq, wrem) : nwarps;
#pragma unroll 1
    for (int pass = 0; pass <= F; pass++) {
      int gp = pass * nthr + tid;
      if (pass == F) {
        __syncthreads();  // every full pass is done
        if (warp < wrem) break;
        gp -= wrem * 32;
        if (gp >= npts) continue;
      }
      const int w = gp / nq, p = gp - w * nq;
      double* a0 = wbase + (size_t)w * wstride;
      double* a1 = a0 + (size_t)NARR * nq8;  // spin b (NSPIN == 2)
      double* sp = a0 + (size_t)NSPIN * NARR * nq8;
      const double rr = sR[p];
      const double wr = sW[p];
      const double w4 = 4.0 * HFB_PI * wr;
      const double wtot = w4 * (rr * rr);
#endif
      if (NSPIN == 1) {
        const double rho = a0[p];
        const double gr = GGA ? 2.0 * a0[nq8 + p] : 0.0;
        const double sigma = gr * gr;
        const double tau = MGGA ? 0.5 * (a0[2 * nq8 + p] + fmax(a0[3 * nq8 + p] / (rr * rr), 0.0)) : 0.0;
        double exc = 0.0, vrho = 0.0, vsigma = 0.0, vtau = 0.0;
        xc::xc_point_unpol<FAM>(xf, thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
        xc::xc_point_unpol<FAM>(cf, thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
        sp[p] = wtot * exc * rho;
        sp[nq8 + p] = wtot * rho;
        sp[2 * nq8 + p] = wtot * tau;
        a0[p] = wtot * vrho;
        if (GGA) a0[nq8 + p] = 2.0 * wtot * vsigma * gr;
        if (MGGA) { a0[2 * nq8 + p] = wtot * 0.5 * vtau; a0[3 * nq8 + p] = vtau * (0.5 * w4); }
      } else {
        double rho[2] = {a0[p], a1[p]};
        double gr[2] = {GGA ? 2.0 * a0[nq8 + p] : 0.0, GGA ? 2.0 * a1[nq8 + p] : 0.0};
        double sigma[3] = {gr[0] * gr[0], gr[0] * gr[1], gr[1] * gr[1]};
        double tau[2] = {0.0, 0.0};
        if (MGGA) {
          tau[0] = 0.5 * (a0[2 * nq8 + p] + fmax(a0[3 * nq8 + p] / (rr * rr), 0.0));
          tau[1] = 0.5 * (a1[2 * nq8 + p] + fmax(a1[3 * nq8 + p] / (rr * rr), 0.0));
        }
        double exc = 0.0, vrho[2] = {0.0, 0.0}, vsigma[3] = {0.0, 0.0, 0.0}, vtau[2] = {0.0, 0.0};
        xc::xc_point_pol<FAM>(xf, thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
        xc::xc_point_pol<FAM>(cf, thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
        sp[p] = wtot * exc * (rho[0] + rho[1]);
        sp[nq8 + p] = wtot * (rho[0] + rho[1]);
        sp[2 * nq8 + p] = wtot * (tau[0] + tau[1]);
        a0[p] = wtot * vrho[0];
        a1[p] = wtot * vrho[1];
        if (GGA) {
          a0[nq8 + p] = wtot * (2.0 * vsigma[0] * gr[0] + vsigma[1] * gr[1]);
          a1[nq8 + p] = wtot * (2.0 * vsigma[2] * gr[1] + vsigma[1] * gr[0]);
        }
        if (MGGA) {
          a0[2 * nq8 + p] = wtot * 0.5 * vtau[0];
          a1[2 * nq8 + p] = wtot * 0.5 * vtau[1];
          a0[3 * nq8 + p] = vtau[0] * (0.5 * w4);
          a1[3 * nq8 + p] = vtau[1] * (0.5 * w4);
        }
      }
    }
#ifndef HFB_GRID_NO_POOL
    if (warp >= wmin) asm volatile("bar.sync 1, %0;" ::"r"((nwarps - wmin) * 32) : "memory");
#else
    __syncwarp();
#endif
  }
  if (active) {
  // the three scalars of this warp's build: fixed lane striding + shuffle tree => run-to-run deterministic
  {
    double my_exc = 0.0, my_nel = 0.0, my_tau = 0.0;
    for (int p = lane; p < nq; p += 32) {
      my_exc += spt[p];
      my_nel += spt[nq8 + p];
      my_tau += spt[2 * nq8 + p];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      my_exc += __shfl_xor_sync(0xffffffffu, my_exc, o);
      my_nel += __shfl_xor_sync(0xffffffffu, my_nel, o);
      my_tau += __shfl_xor_sync(0xffffffffu, my_tau, o);
    }
    if (lane == 0) {
      double* out = prm.partial + ((size_t)b * prm.nel + el) * 4;
      out[0] = my_exc; out[1] = my_nel; out[2] = my_tau; out[3] = 0.0;
    }
  }
  HFB_TICK();

  // ================= stage 3: Fock blocks =================
  constexpr int NTRI = NT * (NT + 1) / 2;
  const int KP = nq8 >> 2;
#pragma unroll 1
  for (int s = 0; s < NSPIN; s++) {
    const double* wa = scratch + (size_t)s * NARR * nq8 + t4;
    const double* tb = sB + (size_t)gid * ldq + t4;
    const double* td = sdB + (size_t)gid * ldq + t4;
    // GGA: H = a B B^T + g (B dB^T + dB B^T) = W + W^T with W = B (a/2 B + g dB)^T -- ONE product over all
    // NT^2 tiles (then mirrored through the shared block) instead of two over the upper tile triangle.
    constexpr bool WSYM = GGA && !MGGA;
    if constexpr (WSYM) {
      constexpr int KSPL = (NT <= 2) ? 2 : 1;  // K split in two accumulator sets: twice the independent DMMA chains
      double w[KSPL][NT * NT][2];
#pragma unroll
      for (int q = 0; q < KSPL; q++)
#pragma unroll
        for (int t = 0; t < NT * NT; t++) { w[q][t][0] = 0.0; w[q][t][1] = 0.0; }
#pragma unroll 1
      for (int ks0 = 0; ks0 < KP; ks0 += KSPL) {
#pragma unroll
        for (int q = 0; q < KSPL; q++) {
          const int ks = ks0 + q;  // KP = nq8/4 is even
          const double a = 0.5 * wa[4 * ks];
          const double g = wa[nq8 + 4 * ks];
          double bt[NT], X[NT];
#pragma unroll
          for (int t = 0; t < NT; t++) {
            bt[t] = tb[(size_t)8 * t * ldq + 4 * ks];
            X[t] = fma(g, td[(size_t)8 * t * ldq + 4 * ks], a * bt[t]);
          }
#pragma unroll
          for (int mi = 0; mi < NT; mi++)
#pragma unroll
            for (int ni = 0; ni < NT; ni++) dmma(w[q][mi * NT + ni][0], w[q][mi * NT + ni][1], bt[mi], X[ni]);
        }
      }
      if (s == 0) { HFB_TICK(); }
#pragma unroll
      for (int mi = 0; mi < NT; mi++)
#pragma unroll
        for (int ni = 0; ni < NT; ni++)
#pragma unroll
          for (int cc = 0; cc < 2; cc++) {
            double v = w[0][mi * NT + ni][cc];
            if (KSPL > 1) v += w[KSPL - 1][mi * NT + ni][cc];
            pblk[(8 * ni + 2 * t4 + cc) * LDP + 8 * mi + gid] = v;
          }
      __syncwarp();
    } else {
      // separate accumulators for the B X^T and dB Y^T products: twice the independent DMMA chains
      constexpr bool SPLIT = GGA && NT <= 2;
      double h[NTRI][2], h2[SPLIT ? NTRI : 1][2], hl[MGGA ? NTRI : 1][2];
  #pragma unroll
      for (int t = 0; t < NTRI; t++) {
        h[t][0] = 0.0; h[t][1] = 0.0;
        if (SPLIT) { h2[t][0] = 0.0; h2[t][1] = 0.0; }
        if (MGGA) { hl[t][0] = 0.0; hl[t][1] = 0.0; }
      }
      const double* tb = sB + (size_t)gid * ldq + t4;
      const double* td = sdB + (size_t)gid * ldq + t4;
  #pragma unroll 2
      for (int ks = 0; ks < KP; ks++) {
        const double a = wa[4 * ks];
        const double g = GGA ? wa[nq8 + 4 * ks] : 0.0;
        const double c = MGGA ? wa[2 * nq8 + 4 * ks] : 0.0;
        const double dl = MGGA ? wa[3 * nq8 + 4 * ks] : 0.0;
        double bt[NT], dbt[GGA ? NT : 1], X[NT], Y[GGA ? NT : 1], Z[MGGA ? NT : 1];
  #pragma unroll
        for (int t = 0; t < NT; t++) {
          bt[t] = tb[(size_t)8 * t * ldq + 4 * ks];
          if (GGA) {
            dbt[t] = td[(size_t)8 * t * ldq + 4 * ks];
            X[t] = fma(g, dbt[t], a * bt[t]);
            Y[t] = MGGA ? fma(c, dbt[t], g * bt[t]) : g * bt[t];
          } else {
            X[t] = a * bt[t];
          }
          if (MGGA) Z[t] = dl * bt[t];
        }
        int tt = 0;
  #pragma unroll
        for (int mi = 0; mi < NT; mi++)
  #pragma unroll
          for (int ni = mi; ni < NT; ni++) {
            dmma(h[tt][0], h[tt][1], bt[mi], X[ni]);
            if (GGA) {
              if (SPLIT) dmma(h2[tt][0], h2[tt][1], dbt[mi], Y[ni]);
              else dmma(h[tt][0], h[tt][1], dbt[mi], Y[ni]);
            }
            if (MGGA) dmma(hl[tt][0], hl[tt][1], bt[mi], Z[ni]);
            tt++;
          }
      }
      if (SPLIT) {
  #pragma unroll
        for (int t = 0; t < NTRI; t++) { h[t][0] += h2[t][0]; h[t][1] += h2[t][1]; }
      }
      if (s == 0) { HFB_TICK(); }
      // H(i,j) and its mirror image into the per-warp block, then coalesced stores to every l slice
      int tt = 0;
  #pragma unroll
      for (int mi = 0; mi < NT; mi++)
  #pragma unroll
        for (int ni = mi; ni < NT; ni++) {
          const int i = 8 * mi + gid;
  #pragma unroll
          for (int cc = 0; cc < 2; cc++) {
            const int j = 8 * ni + 2 * t4 + cc;
            pblk[j * LDP + i] = h[tt][cc];
            if (MGGA) pblk[PBLK + j * LDP + i] = hl[tt][cc];
            if (ni > mi) {
              pblk[i * LDP + j] = h[tt][cc];
              if (MGGA) pblk[PBLK + i * LDP + j] = hl[tt][cc];
            }
          }
          tt++;
        }
      __syncwarp();
    }
    double* gH = prm.H + ((size_t)(b * NSPIN + s) * prm.nslices * prm.nel + el) * np2;
    {
      double hv[NPASS], hlv[MGGA ? NPASS : 1];
#pragma unroll
      for (int it = 0; it < NPASS; it++) {
        const int col = it * CPI + bcol0;
        hv[it] = (brow < 8 * NT) ? pblk[col * LDP + brow] : 0.0;
        if (WSYM && brow < 8 * NT) hv[it] += pblk[brow * LDP + col];
        if (MGGA) hlv[it] = (brow < 8 * NT) ? pblk[PBLK + col * LDP + brow] : 0.0;
      }
      for (int l = 0; l < prm.nslices; l++) {
        double* dst = gH + (size_t)l * slice_stride + brow;
        const double ll = (double)(l * (l + 1));
#pragma unroll
        for (int it = 0; it < NPASS; it++) {
          const int col = it * CPI + bcol0;
          if (brow_ok && col < np) dst[col * np] = MGGA ? fma(ll, hlv[it], hv[it]) : hv[it];
        }
      }
    }
    __syncwarp();  // the block is reused by the next spin
  }
  }  // if (active)
#ifdef HFB_GRID_PROFILE
  HFB_TICK();
  if (gi == 0 && lane == 0 && el == 3 && (b % 97) == 0)
    printf("grid profile b=%d: load %lld  jpack %lld  dens %lld  xc %lld  fock %lld  store %lld cycles (first spin for load/fock split)\n", b,
           tk[1] - tk[0], tk[2] - tk[1], tk[3] - tk[2], tk[4] - tk[3], tk[5] - tk[4], tk[6] - tk[5]);
#endif
  }  // groups
}

inline int sad_grid_ldq(int nq) {
  int nq8 = (nq + 7) & ~7;
  int ldq = nq8;
  while ((ldq & 15) != 4) ldq++;
  return ldq;
}
inline size_t sad_grid_mma_smem_bytes(int fam, int nspin, int np_pad, int nq8, int ldq, int nwarps, int jKP) {
  const bool gga = fam >= xc::FAM_GGA, mgga = fam >= xc::FAM_MGGA;
  const int narr = 1 + (gga ? 1 : 0) + (mgga ? 2 : 0);
  const size_t blk = (size_t)(mgga ? 2 : 1) * np_pad * sad_grid_ldp(np_pad);
#ifdef HFB_GRID_ALIAS_SPT
  return ((size_t)np_pad * ldq * (gga ? 2 : 1) + 3 * (size_t)jKP + 2 * (size_t)nq8 + (size_t)nwarps * (nspin * narr * nq8 + blk)) * sizeof(double);
#else
  return ((size_t)np_pad * ldq * (gga ? 2 : 1) + 3 * (size_t)jKP + 2 * (size_t)nq8 + (size_t)nwarps * ((nspin * narr + 3) * nq8 + blk)) * sizeof(double);
#endif
}

// [nb][nel][4] partial scalars -> [nb][4], fixed order
static __global__ void sad_reduce_scalars_kernel(const double* partial, int nel, double* out) {
  const int b = blockIdx.x;
  if (threadIdx.x < 4) {
    double s = 0.0;
    for (int e = 0; e < nel; e++) s += partial[((size_t)b * nel + e) * 4 + threadIdx.x];
    out[(size_t)b * 4 + threadIdx.x] = s;
  }
}

}  // namespace hfb
