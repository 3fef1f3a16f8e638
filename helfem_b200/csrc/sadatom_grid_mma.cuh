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
// the three stages with warp-level synchronisation only:
//   1. Q^T = B^T P  (points x functions; A operand = table, B operand = density block in
//      registers), contracted with B / dB in the accumulator layout -> rho, d rho, tau per
//      point (quad shuffles, fixed order => deterministic);
//   2. XC on the points (lanes stride the points), potentials folded into quadrature weights;
//   3. H = B (a B + g dB)^T + dB (g B + c dB)^T  [+ Hl = B dl B^T], upper tile triangle only,
//      written to every l slice as H + l(l+1) Hl.
// Table row stride ldq = 4 (mod 16) doubles makes all three fragment access patterns
// shared-memory-bank-conflict free.
#pragma once
#include <cuda_runtime.h>
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
};

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

// FAM: highest functional family; NSPIN: 1 or 2; NT: 8-row tiles of the element block (np_pad = 8 NT).
template <int FAM, int NSPIN, int NT>
__global__ void __launch_bounds__(256) sad_grid_mma_kernel(SadGridParams prm) {
  extern __shared__ double smem[];
  constexpr bool GGA = (FAM >= xc::FAM_GGA);
  constexpr bool MGGA = (FAM >= xc::FAM_MGGA);
  constexpr int KS = 2 * NT;                   // k-steps over the functions of an element
  constexpr int NARR = 1 + (GGA ? 1 : 0) + (MGGA ? 2 : 0);  // per-spin per-point arrays
  const int el = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  const int gid = lane >> 2, t4 = lane & 3;
  const int np = prm.np_max, np2 = np * np, nq = prm.nq, nq8 = prm.nq8, ldq = prm.ldq;
  const int tabsz = 8 * NT * ldq;

  double* sB = smem;
  double* sdB = sB + tabsz;  // only touched if GGA
  double* scratch = sB + (GGA ? 2 : 1) * tabsz + (size_t)warp * (NSPIN * NARR * nq8);

  // ---- stage the element tables (whole CTA) ----
  {
    const double* gB = prm.Bt + (size_t)el * tabsz;
    for (int k = tid; k < tabsz; k += blockDim.x) sB[k] = gB[k];
    if (GGA) {
      const double* gdB = prm.dBt + (size_t)el * tabsz;
      for (int k = tid; k < tabsz; k += blockDim.x) sdB[k] = gdB[k];
    }
  }
  __syncthreads();
  const int b = blockIdx.y * nwarps + warp;
  if (b >= prm.nb) return;  // no block-level synchronisation below

  const size_t slice_stride = (size_t)prm.nel * np2;
  const int MT = nq8 >> 3;

  // ================= stage 1: density, gradient, tau on the points =================
  for (int k = lane; k < NSPIN * NARR * nq8; k += 32) scratch[k] = 0.0;
  __syncwarp();
#pragma unroll 1
  for (int s = 0; s < NSPIN; s++) {
    const double* gP = prm.P + ((size_t)(b * NSPIN + s) * prm.nslices * prm.nel + el) * np2;
    double* aR = scratch + (size_t)s * NARR * nq8;  // rho | g | t1 | t2
#pragma unroll 1
    for (int ni = 0; ni < NT; ni++) {
      // B operand: Bop[k][n] = P(n, k), n = 8 ni + gid, k = 4 ks + t4  (summed over l)
      double pf[KS], plf[MGGA ? KS : 1];
      const int n = 8 * ni + gid;
#pragma unroll
      for (int ks = 0; ks < KS; ks++) {
        const int k = 4 * ks + t4;
        double p = 0.0, pl = 0.0;
        if (n < np && k < np) {
          const double* src = gP + n + (size_t)k * np;
          for (int l = 0; l < prm.nslices; l++) {
            const double v = __ldg(src + (size_t)l * slice_stride);
            p += v;
            if (MGGA) pl = fma((double)(l * (l + 1)), v, pl);
          }
        }
        pf[ks] = p;
        if (MGGA) plf[ks] = pl;
      }
      const double* mb0 = sB + (size_t)(8 * ni + 2 * t4) * ldq + gid;
      const double* md0 = sdB + (size_t)(8 * ni + 2 * t4) * ldq + gid;
#pragma unroll 2
      for (int mi = 0; mi < MT; mi++) {
        const double* ab = sB + (size_t)t4 * ldq + 8 * mi + gid;
        const double* ad = sdB + (size_t)t4 * ldq + 8 * mi + gid;
        double q0 = 0.0, q1 = 0.0, qp0 = 0.0, qp1 = 0.0, ql0 = 0.0, ql1 = 0.0;
#pragma unroll
        for (int ks = 0; ks < KS; ks++) {
          const double a = ab[(size_t)4 * ks * ldq];
          dmma(q0, q1, a, pf[ks]);
          if (MGGA) {
            dmma(qp0, qp1, ad[(size_t)4 * ks * ldq], pf[ks]);
            dmma(ql0, ql1, a, plf[ks]);
          }
        }
        const double b0 = mb0[8 * mi], b1 = mb0[ldq + 8 * mi];
        double rho = fma(q1, b1, q0 * b0), g = 0.0, t1 = 0.0, t2 = 0.0;
        if (GGA) {
          const double d0 = md0[8 * mi], d1 = md0[ldq + 8 * mi];
          g = fma(q1, d1, q0 * d0);
          if (MGGA) {
            t1 = fma(qp1, d1, qp0 * d0);
            t2 = fma(ql1, b1, ql0 * b0);
          }
        }
        rho = quad_sum(rho);
        if (GGA) g = quad_sum(g);
        if (MGGA) { t1 = quad_sum(t1); t2 = quad_sum(t2); }
        if (t4 == 0) {
          const int p = 8 * mi + gid;
          aR[p] += rho;
          if (GGA) aR[nq8 + p] += g;
          if (MGGA) { aR[2 * nq8 + p] += t1; aR[3 * nq8 + p] += t2; }
        }
      }
    }
  }
  __syncwarp();

  // ================= stage 2: XC on the points =================
  double my_exc = 0.0, my_nel = 0.0, my_tau = 0.0;
  {
    const xc::XcThresholds thr = xc::make_thresholds(prm.dens_thr);
    double* a0 = scratch;
    double* a1 = scratch + (size_t)NARR * nq8;  // spin b (NSPIN == 2)
    for (int p = lane; p < nq8; p += 32) {
      if (p >= nq) {
#pragma unroll
        for (int k = 0; k < NARR; k++) { a0[k * nq8 + p] = 0.0; if (NSPIN == 2) a1[k * nq8 + p] = 0.0; }
        continue;
      }
      const double rr = prm.r[(size_t)el * nq + p];
      const double wr = prm.wrad[(size_t)el * nq + p];
      const double w4 = 4.0 * HFB_PI * wr;
      const double wtot = w4 * (rr * rr);
      if (NSPIN == 1) {
        const double rho = a0[p];
        const double gr = GGA ? 2.0 * a0[nq8 + p] : 0.0;
        const double sigma = gr * gr;
        const double tau = MGGA ? 0.5 * (a0[2 * nq8 + p] + fmax(a0[3 * nq8 + p] / (rr * rr), 0.0)) : 0.0;
        double exc = 0.0, vrho = 0.0, vsigma = 0.0, vtau = 0.0;
        xc::xc_point_unpol<FAM>(prm.x_func, thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
        xc::xc_point_unpol<FAM>(prm.c_func, thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
        my_exc += wtot * exc * rho;
        my_nel += wtot * rho;
        my_tau += wtot * tau;
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
        xc::xc_point_pol<FAM>(prm.x_func, thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
        xc::xc_point_pol<FAM>(prm.c_func, thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
        my_exc += wtot * exc * (rho[0] + rho[1]);
        my_nel += wtot * (rho[0] + rho[1]);
        my_tau += wtot * (tau[0] + tau[1]);
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
  }
  // warp reduction of the three scalars (fixed tree => run-to-run deterministic)
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
  __syncwarp();

  // ================= stage 3: Fock blocks =================
  constexpr int NTRI = NT * (NT + 1) / 2;
  const int KP = nq8 >> 2;
#pragma unroll 1
  for (int s = 0; s < NSPIN; s++) {
    const double* wa = scratch + (size_t)s * NARR * nq8 + t4;
    double h[NTRI][2], hl[MGGA ? NTRI : 1][2];
#pragma unroll
    for (int t = 0; t < NTRI; t++) { h[t][0] = 0.0; h[t][1] = 0.0; if (MGGA) { hl[t][0] = 0.0; hl[t][1] = 0.0; } }
    const double* tb = sB + (size_t)gid * ldq + t4;
    const double* td = sdB + (size_t)gid * ldq + t4;
#pragma unroll 1
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
          if (GGA) dmma(h[tt][0], h[tt][1], dbt[mi], Y[ni]);
          if (MGGA) dmma(hl[tt][0], hl[tt][1], bt[mi], Z[ni]);
          tt++;
        }
    }
    // write H(i,j) (and the mirrored element for off-diagonal tiles) to every l slice
    double* gH = prm.H + ((size_t)(b * NSPIN + s) * prm.nslices * prm.nel + el) * np2;
    int tt = 0;
#pragma unroll
    for (int mi = 0; mi < NT; mi++)
#pragma unroll
      for (int ni = mi; ni < NT; ni++) {
        const int i = 8 * mi + gid;
#pragma unroll
        for (int cc = 0; cc < 2; cc++) {
          const int j = 8 * ni + 2 * t4 + cc;
          if (i < np && j < np) {
            for (int l = 0; l < prm.nslices; l++) {
              const double val = MGGA ? fma((double)(l * (l + 1)), hl[tt][cc], h[tt][cc]) : h[tt][cc];
              gH[(size_t)l * slice_stride + i + (size_t)j * np] = val;
              if (ni > mi) gH[(size_t)l * slice_stride + j + (size_t)i * np] = val;
            }
          }
        }
        tt++;
      }
  }
}

inline int sad_grid_ldq(int nq) {
  int nq8 = (nq + 7) & ~7;
  int ldq = nq8;
  while ((ldq & 15) != 4) ldq++;
  return ldq;
}
inline size_t sad_grid_mma_smem_bytes(int fam, int nspin, int np_pad, int nq8, int ldq, int nwarps) {
  const bool gga = fam >= xc::FAM_GGA, mgga = fam >= xc::FAM_MGGA;
  const int narr = 1 + (gga ? 1 : 0) + (mgga ? 2 : 0);
  return ((size_t)np_pad * ldq * (gga ? 2 : 1) + (size_t)nwarps * nspin * narr * nq8) * sizeof(double);
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
