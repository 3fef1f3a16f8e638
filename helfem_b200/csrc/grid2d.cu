// XC Fock build on the angular x radial PRODUCT grid of the `atomic` and `diatomic` programs.
//
// Reference operators (read-only tree /root/reference):
//   atomic::dftgrid::DFTGrid::eval_Fxc     src/atomic/dftgrid.cpp:810-928    (worker: :51-232, :523-673, :710-796)
//   diatomic::dftgrid::DFTGrid::eval_Fxc   src/diatomic/dftgrid.cpp:769-856  (worker: :49-210, :499-640, :669-755)
//
// The reference forms complex basis-function matrices bf = Y_lm(theta,phi) x chi_k(r) of size
// (nang*np) x (nang_pts*nq) per element and runs dense complex GEMMs on them, O((nang np)^2 npts nq).
// The basis is a product, so is the grid: here every contraction is factorised into a RADIAL stage and an
// ANGULAR stage, O(nang^2 (np^2 nq + npts nq)) -- np^2 = 225 times fewer flops -- and nothing complex is
// ever stored: the angular dependence enters through five REAL pair tables (constants of the basis)
//     z = Y_a conj(Y_b):  W0 = Re z, W1 = Re[dY_a conj Y_b], W2 = -m_a Im z, W3 = Re[dY_a conj dY_b], W4 = m_a m_b Re z
// (dY = d/dtheta Y, TwoDBasis.cpp:1407-1419; d/dphi Y = i m Y).  Per Fock build:
//   A  g2d_radial_density   R_ab(q), R'_ab(q), R''_ab(q) = chi^T P_ab chi, chi'^T P_ab chi, chi'^T P_ab chi'
//   B  g2d_point_xc         rho, grad rho (3 components / metric), tau at every (angular point, q); XC (xc_point.cuh);
//                            seven weighted potential fields; Exc, Nel, Ekin partial sums (fixed order)
//   C  g2d_angular_fock     V_ab(q), G_ab(q), T_ab(q) = sum over angular points of fields x pair tables
//   D  g2d_radial_fock      H_ab = chi V chi^T + chi' G chi^T + chi G chi'^T + chi' T chi'^T, added into the dense
//                            matrix; even elements first, then odd ones (no two blocks of a pass overlap), so the
//                            result is run-to-run deterministic without atomics.
// With the angular rule the reference always builds -- a rule in theta times a uniform rule in phi
// (angular::angular_chebyshev, general/angular.cpp:64-71; point i nphi + j) -- the angular stages factorise once more:
// Y_a = Theta_a(theta) e^{i m_a phi}, so every pair table is (a product of Theta's) x cos / sin (Delta phi), Delta = m_a - m_b.
//   B1 g2d_theta_density    a_x(t, Delta, q) = sum over the pairs of the class Delta of (theta factor) x R_ab(q)
//   B  g2d_point_xc<PROD>   fields at (t, j) = sum over Delta of cos / sin (Delta phi_j) a_x(t, Delta, q); XC as before
//   C1 g2d_phi_fields       f_x(t, Delta, q) = sum_j F_x(t, j, q) cos / sin (Delta phi_j)
//   C2 g2d_theta_fock       V_ab, G_ab, T_ab = sum_t (theta factors) x f_x(t, Delta_ab, q)
// nphi times fewer operations in the two stages that dominate; pairs whose radial density vanishes are skipped (exact).
// setup() detects the product structure from the tables it is given (and checks Y against Theta e^{i m phi}); any other
// angular rule takes the general pair-table kernels (HFB200_G2D_GENERAL=1 forces them).
// Matrices cross PCIe as packed angular blocks: the non-zero blocks of P down, the blocks of H that can be non-zero back
// (all of them, unless every non-zero block of P has m_a = m_b: then rho and every field are independent of phi and H_ab
// vanishes identically for m_a != m_b -- the reference gets round-off noise there).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <complex>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../include/helfem_b200.h"
#include "context.h"
#include "grid2d.h"
#include "../host/blockscan.h"
#include "xc_point.cuh"

namespace hfb {

struct G2DParams {
  int kind, nang, npair, nel, nq, np_max, nrad, npts;
  int fam, x_func, c_func, nspin;
  double Rhalf, dens_thr;
  const int *first, *npe;
  const double *chi, *dchi, *r, *wr, *cth, *wang, *Wp, *Wt;
  const double* Pp;  // [nspin][nin][nrad*nrad] packed angular blocks of P
  double* Hp;        // [nspin][nout][nrad*nrad] packed angular blocks of H
  int nin, nout;
  const int *pidx, *oidx;  // [npair] block -> packed slot (-1: zero / not computed)
  const int2* outblk;      // [nout]
  double *Rq, *F, *VGT, *partial;
  unsigned char* nzR;      // [nspin][nel][npair]
  int* clsnz;              // [nspin*nel][nD] some pair of the class has a non-zero radial density (product rule; zeroed per call)
  int nblk_p;  // blocks over angular points in stage B
  // product rule
  int nth, nph, nD, dmax;
  const double *Th, *DTh, *cosT, *sinT;
  const int* mval;
  const int* cls_ptr;
  const int2* cls_pair;
  const int4* chunks;
  double *Aq, *Fh;   // [s][e][field][t][class][q]
};

// ---- A: radial density transforms --------------------------------------------------------------------------
__global__ void g2d_radial_density(G2DParams p) {
  extern __shared__ double sm[];
  const int ab = blockIdx.x, e = blockIdx.y, s = blockIdx.z;
  const int np = p.npe[e], f0 = p.first[e], nq = p.nq, npm = p.np_max;
  double* Ps = sm;
  double* chi = Ps + npm * npm;
  double* dchi = chi + npm * nq;
  const int slot = p.pidx[ab];
  double* out = p.Rq + ((size_t)(s * p.nel + e) * 3) * p.npair * nq + (size_t)ab * nq;
  const bool want_d = p.fam >= xc::FAM_GGA;
  int nz = 0;
  if (slot >= 0) {
    const double* Pg = p.Pp + ((size_t)s * p.nin + slot) * p.nrad * p.nrad + f0 + (size_t)f0 * p.nrad;
    for (int i = threadIdx.x; i < np * np; i += blockDim.x) {
      const int k = i % np, l = i / np;
      const double v = Pg[k + (size_t)l * p.nrad];
      Ps[k + l * npm] = v;
      nz |= (v != 0.0);
    }
  }
  nz = __syncthreads_or(nz);
  if (threadIdx.x == 0) {
    p.nzR[(size_t)(s * p.nel + e) * p.npair + ab] = nz ? 1 : 0;
    if (nz && p.clsnz) atomicOr(p.clsnz + (s * p.nel + e) * p.nD + (p.mval[ab / p.nang] - p.mval[ab % p.nang] + p.dmax), 1);  // (a flag: order-independent)
  }
  const int q = threadIdx.x;
  if (!nz) {  // block-uniform
    if (q < nq) {
      out[q] = 0.0;
      if (want_d) { out[(size_t)p.npair * nq + q] = 0.0; out[(size_t)2 * p.npair * nq + q] = 0.0; }
    }
    return;
  }
  for (int i = threadIdx.x; i < npm * nq; i += blockDim.x) {
    chi[i] = p.chi[(size_t)e * npm * nq + i];
    if (want_d) dchi[i] = p.dchi[(size_t)e * npm * nq + i];
  }
  __syncthreads();
  if (q >= nq) return;
  double R = 0.0, Rr = 0.0, Rrr = 0.0;
  for (int l = 0; l < np; l++) {
    double u = 0.0, up = 0.0;
    for (int k = 0; k < np; k++) {
      const double pv = Ps[k + l * npm];
      u = fma(chi[k * nq + q], pv, u);
      if (want_d) up = fma(dchi[k * nq + q], pv, up);
    }
    const double cl = chi[l * nq + q];
    R = fma(u, cl, R);
    if (want_d) {
      Rr = fma(up, cl, Rr);
      Rrr = fma(up, dchi[l * nq + q], Rrr);
    }
  }
  out[q] = R;
  if (want_d) {
    out[(size_t)p.npair * nq + q] = Rr;
    out[(size_t)2 * p.npair * nq + q] = Rrr;
  }
}

// ---- B1: theta stage of the densities (product rule) -------------------------------------------------------------
// fields per (s, e): 0 Th_a Th_b R | 1 Th_a Th_b R' | 2 Th'_a Th_b R | 3 -m_a Th_a Th_b R | 4 Th_a Th_b R'' | 5 Th'_a Th'_b R
// | 6 m_a m_b Th_a Th_b R.   grid (class, nel*nspin, theta chunks of G2D_TT), one thread per q.
constexpr int G2D_TT = 4;
constexpr int G2D_SEG = 4;  // the pairs of a class are dealt over (up to) this many thread rows: a serial loop over 6 400 pairs is latency-bound
__host__ __device__ inline int g2d_nfa(int fam) { return fam >= xc::FAM_MGGA ? 7 : (fam >= xc::FAM_GGA ? 4 : 1); }
// block (nqb, nseg <= G2D_SEG); dynamic smem (nseg - 1) * NF * G2D_TT * nqb doubles
template <int FAM>
__global__ void g2d_theta_density(G2DParams p) {
  constexpr int NF = FAM >= xc::FAM_MGGA ? 7 : (FAM >= xc::FAM_GGA ? 4 : 1);
  extern __shared__ double sred[];
  const int d = blockIdx.x, se = blockIdx.y, t0 = blockIdx.z * G2D_TT, q = threadIdx.x, seg = threadIdx.y;
  const int nq = p.nq, nth = p.nth, nqb = blockDim.x, nseg = blockDim.y;
  const bool live = q < nq;
  double acc[NF][G2D_TT];
#pragma unroll
  for (int f = 0; f < NF; f++)
#pragma unroll
    for (int t = 0; t < G2D_TT; t++) acc[f][t] = 0.0;
  const unsigned char* nz = p.nzR + (size_t)se * p.npair;
  const double* Rb = p.Rq + ((size_t)se * 3) * p.npair * nq + q;
  const size_t rs = (size_t)p.npair * nq;
  if (live && p.clsnz[se * p.nD + d])  // block-uniform: an m-diagonal density populates the Delta = 0 class only
    for (int i = p.cls_ptr[d] + seg; i < p.cls_ptr[d + 1]; i += nseg) {
      const int2 ab = p.cls_pair[i];
      const int pr = ab.x * p.nang + ab.y;
      if (!nz[pr]) continue;  // warp-uniform
      const double R = Rb[(size_t)pr * nq];
      double Rr = 0.0, Rrr = 0.0;
      if (FAM >= xc::FAM_GGA) Rr = Rb[rs + (size_t)pr * nq];
      if (FAM >= xc::FAM_MGGA) Rrr = Rb[2 * rs + (size_t)pr * nq];
      const double ma = (double)p.mval[ab.x], mb = (double)p.mval[ab.y];
#pragma unroll
      for (int t = 0; t < G2D_TT; t++) {
        if (t0 + t >= nth) break;
        const double ta = __ldg(p.Th + (size_t)ab.x * nth + t0 + t), tb = __ldg(p.Th + (size_t)ab.y * nth + t0 + t);
        const double tt = ta * tb;
        acc[0][t] = fma(tt, R, acc[0][t]);
        if (FAM >= xc::FAM_GGA) {
          const double da = __ldg(p.DTh + (size_t)ab.x * nth + t0 + t);
          acc[1][t] = fma(tt, Rr, acc[1][t]);
          acc[2][t] = fma(da * tb, R, acc[2][t]);
          acc[3][t] = fma(-ma * tt, R, acc[3][t]);
          if (FAM >= xc::FAM_MGGA) {
            const double db = __ldg(p.DTh + (size_t)ab.y * nth + t0 + t);
            acc[4][t] = fma(tt, Rrr, acc[4][t]);
            acc[5][t] = fma(da * db, R, acc[5][t]);
            acc[6][t] = fma(ma * mb * tt, R, acc[6][t]);
          }
        }
      }
    }
  // fixed-order sum over the segments: rows 1 .. SEG-1 park their sums, row 0 adds them in order
  if (seg > 0) {
#pragma unroll
    for (int f = 0; f < NF; f++)
#pragma unroll
      for (int t = 0; t < G2D_TT; t++) sred[(((size_t)(seg - 1) * NF + f) * G2D_TT + t) * nqb + q] = acc[f][t];
  }
  __syncthreads();
  if (seg != 0 || !live) return;
  // Aq[se][f][t][d][q]
#pragma unroll
  for (int f = 0; f < NF; f++)
#pragma unroll
    for (int t = 0; t < G2D_TT; t++) {
      double v = acc[f][t];
      for (int s2 = 0; s2 < nseg - 1; s2++) v += sred[(((size_t)s2 * NF + f) * G2D_TT + t) * nqb + q];
      if (t0 + t < nth) p.Aq[((((size_t)se * NF + f) * nth + t0 + t) * p.nD + d) * nq + q] = v;
    }
}

// ---- B: densities at the points, XC, weighted potential fields ------------------------------------------------
// fields per spin: 0 a = w vrho | 1..3 g_c = w (2 vs g_c [+ vs_ab g'_c]) / h_c | 4..6 t_c = w vtau / (2 h_c^2)
template <int FAM, int NSPIN, bool PROD>
__global__ void g2d_point_xc(G2DParams p) {
  const int q = threadIdx.x, ip = blockIdx.x * blockDim.y + threadIdx.y, e = blockIdx.y;
  const int nq = p.nq, npair = p.npair;
  const bool live = q < nq && ip < p.npts;
  double S0[NSPIN] = {}, S0r[NSPIN] = {}, S0rr[NSPIN] = {}, S1[NSPIN] = {}, S2[NSPIN] = {}, S3[NSPIN] = {}, S4[NSPIN] = {};
  if (live && PROD) {
    // sum over the classes Delta of cos / sin (Delta phi_j) x the theta-stage fields at (t, Delta, q)
    constexpr int NF = FAM >= xc::FAM_MGGA ? 7 : (FAM >= xc::FAM_GGA ? 4 : 1);
    const int t = ip / p.nph, j = ip % p.nph, nD = p.nD, nth = p.nth;
#pragma unroll
    for (int s = 0; s < NSPIN; s++) {
      const double* A = p.Aq + ((((size_t)(s * p.nel + e) * NF) * nth + t) * nD) * nq + q;
      const size_t fs = (size_t)nth * nD * nq;
      for (int d = 0; d < nD; d++) {
        const double c = __ldg(p.cosT + d * p.nph + j);
        const double* Ad = A + (size_t)d * nq;
        S0[s] = fma(c, Ad[0], S0[s]);
        if (FAM >= xc::FAM_GGA) {
          const double sn = __ldg(p.sinT + d * p.nph + j);
          S0r[s] = fma(c, Ad[fs], S0r[s]);
          S1[s] = fma(c, Ad[2 * fs], S1[s]);
          S2[s] = fma(sn, Ad[3 * fs], S2[s]);
        }
        if (FAM >= xc::FAM_MGGA) {
          S0rr[s] = fma(c, Ad[4 * fs], S0rr[s]);
          S3[s] = fma(c, Ad[5 * fs], S3[s]);
          S4[s] = fma(c, Ad[6 * fs], S4[s]);
        }
      }
    }
  }
  if (live && !PROD) {
    const double* W = p.Wp + (size_t)ip * 5 * npair;
    for (int ab = 0; ab < npair; ab++) {
      const double w0 = W[ab];
      double w1 = 0, w2 = 0, w3 = 0, w4 = 0;
      if (FAM >= xc::FAM_GGA) { w1 = W[npair + ab]; w2 = W[2 * npair + ab]; }
      if (FAM >= xc::FAM_MGGA) { w3 = W[3 * npair + ab]; w4 = W[4 * npair + ab]; }
#pragma unroll
      for (int s = 0; s < NSPIN; s++) {
        const double* Rp = p.Rq + ((size_t)(s * p.nel + e) * 3) * npair * nq + (size_t)ab * nq + q;
        const double R = Rp[0];
        S0[s] = fma(w0, R, S0[s]);
        if (FAM >= xc::FAM_GGA) {
          S0r[s] = fma(w0, Rp[(size_t)npair * nq], S0r[s]);
          S1[s] = fma(w1, R, S1[s]);
          S2[s] = fma(w2, R, S2[s]);
        }
        if (FAM >= xc::FAM_MGGA) {
          S0rr[s] = fma(w0, Rp[(size_t)2 * npair * nq], S0rr[s]);
          S3[s] = fma(w3, R, S3[s]);
          S4[s] = fma(w4, R, S4[s]);
        }
      }
    }
  }
  double part[3] = {0.0, 0.0, 0.0};  // Exc, Nel, Ekin
  if (live) {
    // metric and weight (atomic/dftgrid.cpp:724-743, diatomic/dftgrid.cpp:690-710)
    const double cth = p.cth[ip], sth = sqrt(1.0 - cth * cth), rr = p.r[e * nq + q], wr = p.wr[e * nq + q];
    double ihr, iht, ihp, w;
    if (p.kind == 0) {
      ihr = 1.0; iht = 1.0 / rr; ihp = 1.0 / (rr * sth);
      w = p.wang[ip] * wr * rr * rr;
    } else {
      const double sh = sinh(rr), h = p.Rhalf * sqrt(sh * sh + sth * sth);
      ihr = iht = 1.0 / h; ihp = 1.0 / (p.Rhalf * sh * sth);
      w = p.wang[ip] * wr * (p.Rhalf * p.Rhalf * p.Rhalf) * sh * (sh * sh + sth * sth);
    }
    double rho[NSPIN], g[NSPIN][3], tau[NSPIN];
#pragma unroll
    for (int s = 0; s < NSPIN; s++) {
      rho[s] = S0[s];
      g[s][0] = 2.0 * S0r[s] * ihr; g[s][1] = 2.0 * S1[s] * iht; g[s][2] = 2.0 * S2[s] * ihp;
      tau[s] = 0.5 * (S0rr[s] * ihr * ihr + S3[s] * iht * iht + S4[s] * ihp * ihp);
    }
    const xc::XcThresholds thr = xc::make_thresholds(p.dens_thr);
    double exc = 0.0, vrho[2] = {0, 0}, vsig[3] = {0, 0, 0}, vtau[2] = {0, 0};
    if (NSPIN == 1) {
      const double sigma = g[0][0] * g[0][0] + g[0][1] * g[0][1] + g[0][2] * g[0][2];
      xc::xc_point_unpol<FAM>(p.x_func, thr, rho[0], sigma, tau[0], exc, vrho[0], vsig[0], vtau[0]);
      xc::xc_point_unpol<FAM>(p.c_func, thr, rho[0], sigma, tau[0], exc, vrho[0], vsig[0], vtau[0]);
    } else {
      const int o = NSPIN - 1;
      const double sg[3] = {g[0][0] * g[0][0] + g[0][1] * g[0][1] + g[0][2] * g[0][2],
                            g[0][0] * g[o][0] + g[0][1] * g[o][1] + g[0][2] * g[o][2],
                            g[o][0] * g[o][0] + g[o][1] * g[o][1] + g[o][2] * g[o][2]};
      const double rh[2] = {rho[0], rho[o]}, ta[2] = {tau[0], tau[o]};
      xc::xc_point_pol<FAM>(p.x_func, thr, rh, sg, ta, exc, vrho, vsig, vtau);
      xc::xc_point_pol<FAM>(p.c_func, thr, rh, sg, ta, exc, vrho, vsig, vtau);
    }
    double dens = 0.0, ttot = 0.0;
#pragma unroll
    for (int s = 0; s < NSPIN; s++) { dens += rho[s]; ttot += tau[s]; }
    part[0] = w * exc * dens;
    part[1] = w * dens;
    part[2] = FAM >= xc::FAM_MGGA ? w * ttot : 0.0;
    const double ih[3] = {ihr, iht, ihp};
#pragma unroll
    for (int s = 0; s < NSPIN; s++) {
      double* F = p.F + ((size_t)(s * p.nel + e) * 7) * p.npts * nq + (size_t)ip * nq + q;
      const size_t fs = (size_t)p.npts * nq;
      F[0] = w * vrho[s];
      if (FAM >= xc::FAM_GGA) {
#pragma unroll
        for (int c = 0; c < 3; c++) {
          double v;
          if (NSPIN == 1) v = 2.0 * w * vsig[0] * g[0][c];
          else v = w * (2.0 * vsig[2 * s] * g[s][c] + vsig[1] * g[(NSPIN - 1) - s][c]);
          F[(1 + c) * fs] = v * ih[c];
        }
      }
      if (FAM >= xc::FAM_MGGA) {
        const double t = 0.5 * w * vtau[s];
#pragma unroll
        for (int c = 0; c < 3; c++) F[(4 + c) * fs] = t * ih[c] * ih[c];
      }
    }
  }
  // block partial sums in a fixed order (warp shuffle tree, then warp 0 over the warps)
  __shared__ double red[3][32];
  const int tid = threadIdx.y * blockDim.x + threadIdx.x, nthr = blockDim.x * blockDim.y;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    double v = part[i];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((tid & 31) == 0) red[i][tid >> 5] = v;
  }
  __syncthreads();
  if (tid < 3) {
    double v = 0.0;
    for (int wq = 0; wq < (nthr + 31) / 32; wq++) v += red[tid][wq];
    p.partial[((size_t)e * p.nblk_p + blockIdx.x) * 3 + tid] = v;
  }
}

__global__ void __launch_bounds__(256) g2d_reduce_scalars(const double* partial, int n, double* out) {
  // fixed order (run-to-run deterministic): thread t sums the entries t, t + 256, ...; then a shared-memory tree
  __shared__ double red[3][256];
  const int t = threadIdx.x;
  double v[3] = {0.0, 0.0, 0.0};
  for (int k = t; k < n; k += 256)
#pragma unroll
    for (int i = 0; i < 3; i++) v[i] += partial[(size_t)k * 3 + i];
#pragma unroll
  for (int i = 0; i < 3; i++) red[i][t] = v[i];
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if (t < w)
#pragma unroll
      for (int i = 0; i < 3; i++) red[i][t] += red[i][t + w];
    __syncthreads();
  }
  if (t < 3) out[t] = red[t][0];
}

// ---- C1: phi stage of the fields (product rule) -------------------------------------------------------------------
// Fh[se][x][t][class][q]:  x = 0  sum_j F0 cos - Delta F3 sin  (V, with Th_a Th_b) | 1  sum_j F1 cos  (G) | 2  sum_j F2 cos
// (V, with Th'_a Th_b + Th_a Th'_b) | 3  sum_j F4 cos  (T) | 4  sum_j F5 cos  (V, Th'_a Th'_b) | 5  sum_j F6 cos (V, m_a m_b Th_a Th_b)
// grid (theta point, nel * spins, class), one thread per q.
__host__ __device__ inline int g2d_nfh(int fam) { return fam >= xc::FAM_MGGA ? 6 : (fam >= xc::FAM_GGA ? 3 : 1); }
__global__ void g2d_phi_fields(G2DParams p) {
  const int t = blockIdx.x, se = blockIdx.y, d = blockIdx.z, q = threadIdx.x;
  const int nq = p.nq, nph = p.nph, nfh = g2d_nfh(p.fam);
  if (q >= nq) return;
  const double* F = p.F + ((size_t)se * 7) * p.npts * nq + (size_t)(t * nph) * nq + q;
  const size_t fs = (size_t)p.npts * nq;
  const double delta = (double)(d - p.dmax);
  double x0 = 0.0, x1 = 0.0, x2 = 0.0, x3 = 0.0, x4 = 0.0, x5 = 0.0;
  for (int j = 0; j < nph; j++) {
    const double c = __ldg(p.cosT + d * nph + j);
    const double* Fj = F + (size_t)j * nq;
    x0 = fma(Fj[0], c, x0);
    if (p.fam >= xc::FAM_GGA) {
      const double sn = __ldg(p.sinT + d * nph + j);
      x1 = fma(Fj[fs], c, x1);
      x2 = fma(Fj[2 * fs], c, x2);
      x0 = fma(Fj[3 * fs], -delta * sn, x0);
    }
    if (p.fam >= xc::FAM_MGGA) {
      x3 = fma(Fj[4 * fs], c, x3);
      x4 = fma(Fj[5 * fs], c, x4);
      x5 = fma(Fj[6 * fs], c, x5);
    }
  }
  double* out = p.Fh + ((((size_t)se * nfh) * p.nth + t) * p.nD + d) * nq + q;
  const size_t os = (size_t)p.nth * p.nD * nq;
  out[0] = x0;
  if (p.fam >= xc::FAM_GGA) { out[os] = x1; out[2 * os] = x2; }
  if (p.fam >= xc::FAM_MGGA) { out[3 * os] = x3; out[4 * os] = x4; out[5 * os] = x5; }
}

// ---- C2: theta stage of the Fock fields (product rule) -------------------------------------------------------------
// grid (chunks of <= G2D_NPB output blocks of one class, nel, spins), one thread per q.
constexpr int G2D_NPB = 8;
template <int FAM>
__global__ void g2d_theta_fock(G2DParams p) {
  const int4 ch = p.chunks[blockIdx.x];
  const int d = ch.x, slot0 = ch.y, cnt = ch.z, e = blockIdx.y, s = blockIdx.z, q = threadIdx.x;
  const int nq = p.nq, nth = p.nth;
  constexpr int NFH = FAM >= xc::FAM_MGGA ? 6 : (FAM >= xc::FAM_GGA ? 3 : 1);
  __shared__ int sa[G2D_NPB], sb[G2D_NPB];
  if (threadIdx.x < G2D_NPB) {
    const int2 ab = p.outblk[slot0 + min((int)threadIdx.x, cnt - 1)];
    sa[threadIdx.x] = ab.x; sb[threadIdx.x] = ab.y;
  }
  __syncthreads();
  if (q >= nq) return;
  double V[G2D_NPB], G[G2D_NPB], T[G2D_NPB], mm[G2D_NPB];
#pragma unroll
  for (int k = 0; k < G2D_NPB; k++) { V[k] = G[k] = T[k] = 0.0; mm[k] = (double)(p.mval[sa[k]] * p.mval[sb[k]]); }
  const int se = s * p.nel + e;
  const double* Fh = p.Fh + (((size_t)se * NFH) * nth * p.nD + d) * nq + q;
  const size_t ts = (size_t)p.nD * nq, os = (size_t)nth * p.nD * nq;
  for (int t = 0; t < nth; t++) {
    const double* Ft = Fh + (size_t)t * ts;
    const double x0 = Ft[0];
    double x1 = 0, x2 = 0, x3 = 0, x4 = 0, x5 = 0;
    if (FAM >= xc::FAM_GGA) { x1 = Ft[os]; x2 = Ft[2 * os]; }
    if (FAM >= xc::FAM_MGGA) { x3 = Ft[3 * os]; x4 = Ft[4 * os]; x5 = Ft[5 * os]; }
#pragma unroll
    for (int k = 0; k < G2D_NPB; k++) {
      const double ta = __ldg(p.Th + (size_t)sa[k] * nth + t), tb = __ldg(p.Th + (size_t)sb[k] * nth + t);
      const double tt = ta * tb;
      if (FAM == xc::FAM_LDA) V[k] = fma(tt, x0, V[k]);
      if (FAM >= xc::FAM_GGA) {
        const double da = __ldg(p.DTh + (size_t)sa[k] * nth + t), db = __ldg(p.DTh + (size_t)sb[k] * nth + t);
        G[k] = fma(tt, x1, G[k]);
        if (FAM == xc::FAM_GGA) V[k] = fma(tt, x0, V[k]);
        else V[k] = fma(tt, fma(mm[k], x5, x0), V[k]);
        V[k] = fma(da * tb + ta * db, x2, V[k]);
        if (FAM >= xc::FAM_MGGA) {
          T[k] = fma(tt, x3, T[k]);
          V[k] = fma(da * db, x4, V[k]);
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < G2D_NPB; k++) {
    if (k >= cnt) break;
    const int ab = sa[k] * p.nang + sb[k];
    double* out = p.VGT + ((size_t)se * 3) * p.npair * nq + (size_t)ab * nq + q;
    out[0] = V[k];
    out[(size_t)p.npair * nq] = G[k];
    out[(size_t)2 * p.npair * nq] = T[k];
  }
}

// ---- C: angular contraction of the fields ----------------------------------------------------------------------
__global__ void g2d_angular_fock(G2DParams p) {
  const int2 blk = p.outblk[blockIdx.x];
  const int ab = blk.x * p.nang + blk.y, e = blockIdx.y, s = blockIdx.z, q = threadIdx.x;
  const int nq = p.nq, npts = p.npts;
  if (q >= nq) return;
  const double* F = p.F + ((size_t)(s * p.nel + e) * 7) * npts * nq + q;
  const size_t fs = (size_t)npts * nq;
  const double* W = p.Wt + (size_t)ab * npts;
  const size_t ws = (size_t)p.npair * npts;
  double V = 0.0, G = 0.0, T = 0.0;
  for (int ip = 0; ip < npts; ip++) {
    const double w0 = W[ip];
    const double* Fp = F + (size_t)ip * nq;
    V = fma(Fp[0], w0, V);
    if (p.fam >= xc::FAM_GGA) {
      G = fma(Fp[fs], w0, G);
      V = fma(Fp[2 * fs], W[ws + ip], V);
      V = fma(Fp[3 * fs], W[2 * ws + ip], V);
    }
    if (p.fam >= xc::FAM_MGGA) {
      T = fma(Fp[4 * fs], w0, T);
      V = fma(Fp[5 * fs], W[3 * ws + ip], V);
      V = fma(Fp[6 * fs], W[4 * ws + ip], V);
    }
  }
  double* out = p.VGT + ((size_t)(s * p.nel + e) * 3) * p.npair * nq + (size_t)ab * nq + q;
  out[0] = V;
  out[(size_t)p.npair * nq] = G;
  out[(size_t)2 * p.npair * nq] = T;
}

// ---- D: radial back-transform and scatter ---------------------------------------------------------------------
__global__ void g2d_radial_fock(G2DParams p, int parity) {
  extern __shared__ double sm[];
  const int2 blk = p.outblk[blockIdx.x];
  const int ab = blk.x * p.nang + blk.y, e = 2 * blockIdx.y + parity, s = blockIdx.z;
  if (e >= p.nel) return;
  const int np = p.npe[e], f0 = p.first[e], nq = p.nq, npm = p.np_max;
  double* chi = sm;
  double* dchi = chi + npm * nq;
  double* V = dchi + npm * nq;
  double* G = V + nq;
  double* T = G + nq;
  const bool want_d = p.fam >= xc::FAM_GGA;
  for (int i = threadIdx.x; i < npm * nq; i += blockDim.x) {
    chi[i] = p.chi[(size_t)e * npm * nq + i];
    dchi[i] = want_d ? p.dchi[(size_t)e * npm * nq + i] : 0.0;
  }
  const double* vin = p.VGT + ((size_t)(s * p.nel + e) * 3) * p.npair * nq + (size_t)ab * nq;
  for (int i = threadIdx.x; i < nq; i += blockDim.x) {
    V[i] = vin[i];
    G[i] = vin[(size_t)p.npair * nq + i];
    T[i] = vin[(size_t)2 * p.npair * nq + i];
  }
  __syncthreads();
  double* Hg = p.Hp + ((size_t)s * p.nout + blockIdx.x) * p.nrad * p.nrad + f0 + (size_t)f0 * p.nrad;
  for (int i = threadIdx.x; i < np * np; i += blockDim.x) {
    const int k = i % np, l = i / np;
    double h = 0.0;
    for (int q = 0; q < nq; q++) {
      const double ck = chi[k * nq + q], cl = chi[l * nq + q];
      h = fma(ck * cl, V[q], h);
      if (want_d) {
        const double dk = dchi[k * nq + q], dl = dchi[l * nq + q];
        h = fma(dk * cl + ck * dl, G[q], h);
        h = fma(dk * dl, T[q], h);
      }
    }
    Hg[k + (size_t)l * p.nrad] += h;
  }
}

// ---- host side -----------------------------------------------------------------------------------------------
// packed copies of angular blocks (list order) <-> nothing dense: the kernels address the packed blocks directly
void Grid2D::release() {
  for (double** q : {&dChi, &dDChi, &dR, &dWr, &dCth, &dWang, &dWp, &dWt, &dTh, &dDTh, &dCosT, &dSinT})
    if (*q) { cudaFree(*q); *q = nullptr; }
  for (int** q : {&dFirst, &dNpe, &dM, &dClsPtr, &dPidx, &dOidx})
    if (*q) { cudaFree(*q); *q = nullptr; }
  if (dClsPair) { cudaFree(dClsPair); dClsPair = nullptr; }
  if (dOutBlk) { cudaFree(dOutBlk); dOutBlk = nullptr; }
  if (dChunks) { cudaFree(dChunks); dChunks = nullptr; }
  if (dNzR) { cudaFree(dNzR); dNzR = nullptr; }
  if (dClsNz) { cudaFree(dClsNz); dClsNz = nullptr; }
  clsnz_cap = 0;
  if (ev0) { cudaEventDestroy(ev0); cudaEventDestroy(ev1); ev0 = ev1 = nullptr; }
  nzr_cap = 0;
  Pp.release(); Hp.release(); Rq.release(); F.release(); VGT.release(); partial.release(); scal.release(); Aq.release(); Fh.release();
  if (hX.p) { cudaFreeHost(hX.p); hX.p = nullptr; hX.cap = 0; }
  ready = false;
}

template <class T>
static int up(T** dst, const std::vector<T>& src) {
  HFB_CUDA_OK(cudaMalloc(dst, std::max<size_t>(src.size(), 1) * sizeof(T)));
  HFB_CUDA_OK(cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}

int Grid2D::setup(int kind_, double Rhalf_, const std::vector<int>& lval, const std::vector<int>& mval_, int nel_,
                  const std::vector<int>& first_, const std::vector<int>& last_, int nq_, const double* r, const double* wrad,
                  const double* bf, const double* df, int npts_, const double* cth, const double* wang, const double* Y,
                  const double* dY) {
  release();
  kind = kind_; Rhalf = Rhalf_; nang = (int)lval.size(); npair = nang * nang; nel = nel_; nq = nq_; npts = npts_;
  first = first_; mval = mval_;
  npe.resize(nel);
  np_max = 0;
  for (int e = 0; e < nel; e++) { npe[e] = last_[e] - first_[e] + 1; np_max = std::max(np_max, npe[e]); }
  nrad = last_[nel - 1] + 1;
  // radial tables, [e][k][q] q-fastest, zero padded to np_max rows
  std::vector<double> chi((size_t)nel * np_max * nq, 0.0), dchi(chi.size(), 0.0);
  size_t off = 0;
  for (int e = 0; e < nel; e++) {
    for (int k = 0; k < npe[e]; k++)
      for (int q = 0; q < nq; q++) {
        chi[((size_t)e * np_max + k) * nq + q] = bf[off + q + (size_t)k * nq];
        dchi[((size_t)e * np_max + k) * nq + q] = df[off + q + (size_t)k * nq];
      }
    off += (size_t)npe[e] * nq;
  }
  typedef std::complex<double> cx;
  auto Yat = [&](const double* T, int a, int ip) { return cx(T[2 * (a + (size_t)nang * ip)], T[2 * (a + (size_t)nang * ip) + 1]); };
  // ---- is the rule (theta rule) x (uniform phi rule), point t nphi + j, and Y = Theta e^{i m phi}? ----
  product = false; nth = nph = 0;
  if (!getenv("HFB200_G2D_GENERAL")) {
    int n1 = 1;
    while (n1 < npts && cth[n1] == cth[0]) n1++;
    bool ok = npts % n1 == 0;
    const int n0 = ok ? npts / n1 : 0;
    for (int t = 0; ok && t < n0; t++)
      for (int j = 1; j < n1; j++)
        ok = ok && cth[t * n1 + j] == cth[t * n1] && wang[t * n1 + j] == wang[t * n1];
    double worst = 0.0, scale = 0.0;
    for (int t = 0; ok && t < n0; t++)
      for (int a = 0; a < nang; a++) {
        const cx y0 = Yat(Y, a, t * n1), d0 = Yat(dY, a, t * n1);
        scale = std::max(scale, std::abs(y0));
        worst = std::max(worst, std::max(std::fabs(y0.imag()), std::fabs(d0.imag())));
        for (int j = 1; j < n1; j++) {
          const long k = (((long)mval[a] * j) % n1 + n1) % n1;
          const cx ph = std::polar(1.0, 2.0 * M_PI * (double)k / n1);
          worst = std::max(worst, std::max(std::abs(Yat(Y, a, t * n1 + j) - y0.real() * ph), std::abs(Yat(dY, a, t * n1 + j) - d0.real() * ph)));
        }
      }
    if (ok && worst <= 1e-12 * std::max(scale, 1.0)) { product = true; nth = n0; nph = n1; }
  }
  int rc;
  if ((rc = up(&dChi, chi)) || (rc = up(&dDChi, dchi)) || (rc = up(&dR, std::vector<double>(r, r + (size_t)nel * nq))) ||
      (rc = up(&dWr, std::vector<double>(wrad, wrad + (size_t)nel * nq))) || (rc = up(&dCth, std::vector<double>(cth, cth + npts))) ||
      (rc = up(&dWang, std::vector<double>(wang, wang + npts))) || (rc = up(&dFirst, first)) || (rc = up(&dNpe, npe)) || (rc = up(&dM, mval)))
    return rc;
  if (product) {
    int mlo = 0, mhi = 0;
    for (int a = 0; a < nang; a++) { mlo = std::min(mlo, mval[a]); mhi = std::max(mhi, mval[a]); }
    dmax = mhi - mlo; nD = 2 * dmax + 1;
    std::vector<double> Th((size_t)nang * nth), DTh(Th.size()), cT((size_t)nD * nph), sT(cT.size());
    for (int a = 0; a < nang; a++)
      for (int t = 0; t < nth; t++) {
        Th[(size_t)a * nth + t] = Yat(Y, a, t * nph).real();
        DTh[(size_t)a * nth + t] = Yat(dY, a, t * nph).real();
      }
    for (int d = 0; d < nD; d++)
      for (int j = 0; j < nph; j++) {
        const long k = (((long)(d - dmax) * j) % nph + nph) % nph;  // exact argument reduction
        cT[(size_t)d * nph + j] = std::cos(2.0 * M_PI * (double)k / nph);
        sT[(size_t)d * nph + j] = std::sin(2.0 * M_PI * (double)k / nph);
      }
    cls_ptr.assign(nD + 1, 0);
    for (int a = 0; a < nang; a++)
      for (int b = 0; b < nang; b++) cls_ptr[mval[a] - mval[b] + dmax + 1]++;
    for (int d = 0; d < nD; d++) cls_ptr[d + 1] += cls_ptr[d];
    cls_pair.resize((size_t)npair);
    std::vector<int> fill(cls_ptr.begin(), cls_ptr.end() - 1);
    for (int a = 0; a < nang; a++)
      for (int b = 0; b < nang; b++) cls_pair[fill[mval[a] - mval[b] + dmax]++] = make_int2(a, b);
    if ((rc = up(&dTh, Th)) || (rc = up(&dDTh, DTh)) || (rc = up(&dCosT, cT)) || (rc = up(&dSinT, sT)) || (rc = up(&dClsPtr, cls_ptr)) ||
        (rc = up(&dClsPair, cls_pair)))
      return rc;
  } else {
    // general rule: real angular pair tables over all points
    std::vector<double> Wp((size_t)npts * 5 * npair), Wt((size_t)5 * npair * npts);
#pragma omp parallel for schedule(static)
    for (int ip = 0; ip < npts; ip++) {
      std::vector<double> w1((size_t)npair), w2((size_t)npair);
      double* W = &Wp[(size_t)ip * 5 * npair];
      for (int a = 0; a < nang; a++)
        for (int b = 0; b < nang; b++) {
          const cx ya = Yat(Y, a, ip), yb = Yat(Y, b, ip), da = Yat(dY, a, ip), db = Yat(dY, b, ip);
          const cx z = ya * std::conj(yb);
          const int ab = a * nang + b;
          W[ab] = z.real();
          W[npair + ab] = w1[ab] = (da * std::conj(yb)).real();
          W[2 * npair + ab] = w2[ab] = -mval[a] * z.imag();
          W[3 * npair + ab] = (da * std::conj(db)).real();
          W[4 * npair + ab] = (double)(mval[a] * mval[b]) * z.real();
        }
      for (int a = 0; a < nang; a++)
        for (int b = 0; b < nang; b++) {
          const int ab = a * nang + b, ba = b * nang + a;
          Wt[((size_t)0 * npair + ab) * npts + ip] = W[ab];
          Wt[((size_t)1 * npair + ab) * npts + ip] = w1[ab] + w1[ba];
          Wt[((size_t)2 * npair + ab) * npts + ip] = w2[ab] + w2[ba];
          Wt[((size_t)3 * npair + ab) * npts + ip] = W[3 * npair + ab];
          Wt[((size_t)4 * npair + ab) * npts + ip] = W[4 * npair + ab];
        }
    }
    if ((rc = up(&dWp, Wp)) || (rc = up(&dWt, Wt))) return rc;
  }
  HFB_CUDA_OK(cudaMalloc(&dPidx, (size_t)npair * sizeof(int)));
  HFB_CUDA_OK(cudaMalloc(&dOidx, (size_t)npair * sizeof(int)));
  HFB_CUDA_OK(cudaMalloc(&dOutBlk, (size_t)npair * sizeof(int2)));
  HFB_CUDA_OK(cudaMalloc(&dChunks, (size_t)npair * sizeof(int4)));
  ready = true;
  return 0;
}

template <int FAM, bool PROD>
static void launch_point(int nspin, dim3 grid, dim3 block, cudaStream_t st, const G2DParams& p) {
  if (nspin == 1) g2d_point_xc<FAM, 1, PROD><<<grid, block, 0, st>>>(p);
  else g2d_point_xc<FAM, 2, PROD><<<grid, block, 0, st>>>(p);
}

int Grid2D::eval_fxc(int fam, int x_func, int c_func, int nspin, const double* Pa, const double* Pb, int ldp, double dens_thr,
                     double* Ha, double* Hb, int ldh, int beta, double* scalars3, cudaStream_t st, const int* skip_in) {
  const size_t nrad2 = (size_t)nrad * nrad;
  const int nqb = (nq + 31) / 32 * 32;
  if (nqb > 1024) { set_error("grid2d: more than 1024 radial quadrature points per element are not supported"); return HFB200_EINVAL; }
  const int PB = std::max(1, 256 / nqb);
  const int nblk_p = (npts + PB - 1) / PB;
  const int nsp_out = (nspin == 2 && !beta) ? 1 : nspin;
  const bool verbose = getenv("HFB200_X_VERBOSE") != nullptr;
  auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  const double t_start = now();
  launches = 0;
  int rc;
  // rows of angular function i in the caller's matrices: r0[i] .. r0[i] + cnt[i] - 1  <->  radial index skip[i] ..
  std::vector<int> r0(nang), skip(nang), cnt(nang);
  int nout_rows = 0;
  for (int i = 0; i < nang; i++) {
    skip[i] = skip_in ? skip_in[i] : 0;
    cnt[i] = nrad - skip[i];
    r0[i] = nout_rows;
    nout_rows += cnt[i];
  }
  // ---- non-zero angular blocks of P (either spin) ----
  std::vector<unsigned char> nz((size_t)npair, 0);
  {
    const double* mats[2] = {Pa, Pb};
    scan_nonzero_blocks(mats, nspin, ldp, nang, r0.data(), cnt.data(), nz.data());
  }
  std::vector<int2> inb, outb;
  std::vector<int> pidx((size_t)npair, -1), oidx((size_t)npair, -1);
  bool mdiag = true;
  for (int i = 0; i < nang; i++)
    for (int l = 0; l < nang; l++)
      if (nz[(size_t)i * nang + l]) {
        pidx[(size_t)i * nang + l] = (int)inb.size();
        inb.push_back(make_int2(i, l));
        mdiag = mdiag && mval[i] == mval[l];
      }
  // blocks of H that can be non-zero, sorted by class (product rule) so that stage C2 shares the field loads
  std::vector<int4> chunks;
  if (product) {
    for (int d = 0; d < nD; d++) {
      if (mdiag && d != dmax) continue;
      const int s0 = (int)outb.size();
      for (int i = cls_ptr[d]; i < cls_ptr[d + 1]; i++) outb.push_back(cls_pair[i]);
      for (int s = s0; s < (int)outb.size(); s += G2D_NPB) chunks.push_back(make_int4(d, s, std::min(G2D_NPB, (int)outb.size() - s), 0));
    }
  } else {
    for (int a = 0; a < nang; a++)
      for (int b = 0; b < nang; b++)
        if (!mdiag || mval[a] == mval[b]) outb.push_back(make_int2(a, b));
  }
  for (size_t q = 0; q < outb.size(); q++) oidx[(size_t)outb[q].x * nang + outb[q].y] = (int)q;
  const size_t nin = inb.size(), nout = outb.size();
  const double t_scan = now() - t_start;
  const int nfa = g2d_nfa(fam), nfh = g2d_nfh(fam);
  if ((rc = Pp.reserve(std::max<size_t>(nin, 1) * nrad2 * nspin)) || (rc = Hp.reserve(std::max<size_t>(nout, 1) * nrad2 * nsp_out)) ||
      (rc = hX.reserve(std::max<size_t>(std::max(nin * nspin, nout * nsp_out), 1) * nrad2)) ||
      (rc = Rq.reserve((size_t)nspin * nel * 3 * npair * nq)) || (rc = F.reserve((size_t)nspin * nel * 7 * npts * nq)) ||
      (rc = VGT.reserve((size_t)nspin * nel * 3 * npair * nq)) || (rc = partial.reserve((size_t)nel * nblk_p * 3)) || (rc = scal.reserve(4)))
    return rc;
  if (product && ((rc = Aq.reserve((size_t)nspin * nel * nfa * nth * nD * nq)) || (rc = Fh.reserve((size_t)nspin * nel * nfh * nth * nD * nq))))
    return rc;
  if (!dNzR || nzr_cap < (size_t)nspin * nel * npair) {
    if (dNzR) cudaFree(dNzR);
    dNzR = nullptr;
    HFB_CUDA_OK(cudaMalloc(&dNzR, (size_t)nspin * nel * npair));
    nzr_cap = (size_t)nspin * nel * npair;
  }
  // ---- pack the non-zero blocks (full-space layout: zero rows / columns for the absent functions) and upload ----
#pragma omp parallel for collapse(2) schedule(dynamic, 4)
  for (int s = 0; s < nspin; s++)
    for (long q = 0; q < (long)nin; q++) {
      const double* P = s ? Pb : Pa;
      const int i = inb[q].x, l = inb[q].y;
      double* dst = hX.p + ((size_t)s * nin + q) * nrad2;
      for (int b = 0; b < nrad; b++) {
        double* dc = dst + (size_t)b * nrad;
        if (b < skip[l]) { memset(dc, 0, sizeof(double) * nrad); continue; }
        for (int a = 0; a < skip[i]; a++) dc[a] = 0.0;
        memcpy(dc + skip[i], P + (size_t)r0[i] + (size_t)(r0[l] + b - skip[l]) * ldp, sizeof(double) * cnt[i]);
      }
    }
  HFB_CUDA_OK(cudaMemcpyAsync(Pp.p, hX.p, nin * nspin * nrad2 * sizeof(double), cudaMemcpyHostToDevice, st));
  HFB_CUDA_OK(cudaMemcpyAsync(dPidx, pidx.data(), (size_t)npair * sizeof(int), cudaMemcpyHostToDevice, st));
  HFB_CUDA_OK(cudaMemcpyAsync(dOidx, oidx.data(), (size_t)npair * sizeof(int), cudaMemcpyHostToDevice, st));
  if (nout) HFB_CUDA_OK(cudaMemcpyAsync(dOutBlk, outb.data(), nout * sizeof(int2), cudaMemcpyHostToDevice, st));
  if (!chunks.empty()) HFB_CUDA_OK(cudaMemcpyAsync(dChunks, chunks.data(), chunks.size() * sizeof(int4), cudaMemcpyHostToDevice, st));
  HFB_CUDA_OK(cudaMemsetAsync(Hp.p, 0, std::max<size_t>(nout, 1) * nrad2 * nsp_out * sizeof(double), st));
  HFB_CUDA_OK(cudaStreamSynchronize(st));  // the pageable lists and hX are reused below
  const double t_pack = now() - t_start - t_scan;
  if (!ev0) { HFB_CUDA_OK(cudaEventCreate(&ev0)); HFB_CUDA_OK(cudaEventCreate(&ev1)); }
  HFB_CUDA_OK(cudaEventRecord(ev0, st));

  G2DParams p;
  memset(&p, 0, sizeof(p));
  p.kind = kind; p.nang = nang; p.npair = npair; p.nel = nel; p.nq = nq; p.np_max = np_max; p.nrad = nrad; p.npts = npts;
  p.fam = fam; p.x_func = x_func; p.c_func = c_func; p.nspin = nspin; p.Rhalf = Rhalf; p.dens_thr = dens_thr;
  p.first = dFirst; p.npe = dNpe; p.chi = dChi; p.dchi = dDChi; p.r = dR; p.wr = dWr; p.cth = dCth; p.wang = dWang; p.Wp = dWp; p.Wt = dWt;
  p.Pp = Pp.p; p.Hp = Hp.p; p.nin = (int)nin; p.nout = (int)nout; p.pidx = dPidx; p.oidx = dOidx; p.outblk = dOutBlk;
  p.Rq = Rq.p; p.F = F.p; p.VGT = VGT.p; p.partial = partial.p; p.nzR = dNzR; p.nblk_p = nblk_p;
  p.nth = nth; p.nph = nph; p.nD = nD; p.dmax = dmax; p.Th = dTh; p.DTh = dDTh; p.cosT = dCosT; p.sinT = dSinT; p.mval = dM;
  p.cls_ptr = dClsPtr; p.cls_pair = dClsPair; p.chunks = dChunks; p.Aq = Aq.p; p.Fh = Fh.p;
  if (product) {
    if (!dClsNz || clsnz_cap < (size_t)nspin * nel * nD) {
      if (dClsNz) cudaFree(dClsNz);
      dClsNz = nullptr;
      HFB_CUDA_OK(cudaMalloc(&dClsNz, (size_t)nspin * nel * nD * sizeof(int)));
      clsnz_cap = (size_t)nspin * nel * nD;
    }
    HFB_CUDA_OK(cudaMemsetAsync(dClsNz, 0, (size_t)nspin * nel * nD * sizeof(int), st));
  }
  p.clsnz = product ? dClsNz : nullptr;

  const size_t smA = ((size_t)np_max * np_max + 2 * (size_t)np_max * nq) * sizeof(double);
  const size_t smD = (2 * (size_t)np_max * nq + 3 * (size_t)nq) * sizeof(double);
  if (smA > 227 * 1024 || smD > 227 * 1024) { set_error("grid2d: element tables exceed shared memory"); return HFB200_EINVAL; }
  HFB_CUDA_OK(cudaFuncSetAttribute(g2d_radial_density, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smA));
  HFB_CUDA_OK(cudaFuncSetAttribute(g2d_radial_fock, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smD));
  g2d_radial_density<<<dim3(npair, nel, nspin), nqb, smA, st>>>(p);
  HFB_CUDA_OK(cudaGetLastError());
  launches++;
  const dim3 gB(nblk_p, nel), bB(nqb, PB);
  if (product) {
    const dim3 g1(nD, nel * nspin, (nth + G2D_TT - 1) / G2D_TT);
    const dim3 b1(nqb, std::max(1, std::min(G2D_SEG, 1024 / nqb)));
    const size_t sm1 = (size_t)(b1.y - 1) * nfa * G2D_TT * nqb * sizeof(double);
    if (fam == xc::FAM_LDA) {
      HFB_CUDA_OK(cudaFuncSetAttribute(g2d_theta_density<xc::FAM_LDA>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(sm1, 1024)));
      g2d_theta_density<xc::FAM_LDA><<<g1, b1, sm1, st>>>(p);
    } else if (fam == xc::FAM_GGA) {
      HFB_CUDA_OK(cudaFuncSetAttribute(g2d_theta_density<xc::FAM_GGA>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(sm1, 1024)));
      g2d_theta_density<xc::FAM_GGA><<<g1, b1, sm1, st>>>(p);
    } else {
      HFB_CUDA_OK(cudaFuncSetAttribute(g2d_theta_density<xc::FAM_MGGA>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(sm1, 1024)));
      g2d_theta_density<xc::FAM_MGGA><<<g1, b1, sm1, st>>>(p);
    }
    HFB_CUDA_OK(cudaGetLastError());
    launches++;
    if (fam == xc::FAM_LDA) launch_point<xc::FAM_LDA, true>(nspin, gB, bB, st, p);
    else if (fam == xc::FAM_GGA) launch_point<xc::FAM_GGA, true>(nspin, gB, bB, st, p);
    else launch_point<xc::FAM_MGGA, true>(nspin, gB, bB, st, p);
  } else {
    if (fam == xc::FAM_LDA) launch_point<xc::FAM_LDA, false>(nspin, gB, bB, st, p);
    else if (fam == xc::FAM_GGA) launch_point<xc::FAM_GGA, false>(nspin, gB, bB, st, p);
    else launch_point<xc::FAM_MGGA, false>(nspin, gB, bB, st, p);
  }
  HFB_CUDA_OK(cudaGetLastError());
  g2d_reduce_scalars<<<1, 256, 0, st>>>(partial.p, nel * nblk_p, scal.p);
  HFB_CUDA_OK(cudaGetLastError());
  launches += 2;
  if (nout) {
    if (product) {
      g2d_phi_fields<<<dim3(nth, nel * nsp_out, nD), nqb, 0, st>>>(p);
      HFB_CUDA_OK(cudaGetLastError());
      const dim3 g2((unsigned)chunks.size(), nel, nsp_out);
      if (fam == xc::FAM_LDA) g2d_theta_fock<xc::FAM_LDA><<<g2, nqb, 0, st>>>(p);
      else if (fam == xc::FAM_GGA) g2d_theta_fock<xc::FAM_GGA><<<g2, nqb, 0, st>>>(p);
      else g2d_theta_fock<xc::FAM_MGGA><<<g2, nqb, 0, st>>>(p);
      HFB_CUDA_OK(cudaGetLastError());
      launches += 2;
    } else {
      g2d_angular_fock<<<dim3((unsigned)nout, nel, nsp_out), nqb, 0, st>>>(p);
      HFB_CUDA_OK(cudaGetLastError());
      launches++;
    }
    for (int parity = 0; parity < 2; parity++) {
      const int ne = (nel - parity + 1) / 2;
      if (ne < 1) continue;
      g2d_radial_fock<<<dim3((unsigned)nout, ne, nsp_out), 256, smD, st>>>(p, parity);
      HFB_CUDA_OK(cudaGetLastError());
      launches++;
    }
  }
  HFB_CUDA_OK(cudaEventRecord(ev1, st));
  if (nout) HFB_CUDA_OK(cudaMemcpyAsync(hX.p, Hp.p, nout * nsp_out * nrad2 * sizeof(double), cudaMemcpyDeviceToHost, st));
  HFB_CUDA_OK(cudaMemcpyAsync(scalars3, scal.p, 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
  const double t_launch = now() - t_start - t_scan - t_pack;
  // zero-fill the caller's matrices while the device works (H.zeros in the reference)
  for (int s = 0; s < nsp_out; s++) zero_fill(s ? Hb : Ha, ldh, nout_rows);
  const double t_zero = now() - t_start - t_scan - t_pack - t_launch;
  HFB_CUDA_OK(cudaStreamSynchronize(st));
  const double t_wait = now() - t_start - t_scan - t_pack - t_launch - t_zero;
  HFB_CUDA_OK(cudaEventElapsedTime(&device_ms, ev0, ev1));
  h2d_bytes = (double)(nin * nspin * nrad2 * sizeof(double));
  d2h_bytes = (double)(nout * nsp_out * nrad2 * sizeof(double));
#pragma omp parallel for collapse(2) schedule(dynamic, 4)
  for (int s = 0; s < nsp_out; s++)
    for (long q = 0; q < (long)nout; q++) {
      double* H = s ? Hb : Ha;
      const int j = outb[q].x, k = outb[q].y;
      const double* src = hX.p + ((size_t)s * nout + q) * nrad2;
      for (int b = skip[k]; b < nrad; b++)
        memcpy(H + (size_t)r0[j] + (size_t)(r0[k] + b - skip[k]) * ldh, src + (size_t)b * nrad + skip[j], sizeof(double) * cnt[j]);
    }
  if (verbose)
    fprintf(stderr, "[hfb200] eval_Fxc host call (%s rule%s): scan %.1f ms, pack+upload %.1f ms (%zu blocks), launches %.1f ms, zero-fill %.1f ms, wait %.1f ms, total %.1f ms (%zu blocks back)\n",
            product ? "product" : "general", mdiag ? ", m-diagonal density" : "", 1e3 * t_scan, 1e3 * t_pack, nin, 1e3 * t_launch, 1e3 * t_zero, 1e3 * t_wait,
            1e3 * (now() - t_start), nout);
  return 0;
}

}  // namespace hfb
