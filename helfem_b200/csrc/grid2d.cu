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
// This is the first correct version of these rows: FP64 FMA kernels with
// coalesced q-fastest layouts; DMMA tiling of stages A/D and the Delta-m factorisation of B/C are the next step.
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstring>
#include <vector>
#include "../../include/helfem_b200.h"
#include "context.h"
#include "grid2d.h"
#include "xc_point.cuh"

namespace hfb {

struct G2DParams {
  int kind, nang, npair, nel, nq, np_max, nrad, npts, ntot;
  int fam, x_func, c_func;
  double Rhalf, dens_thr;
  const int *first, *npe;
  const double *chi, *dchi, *r, *wr, *cth, *wang, *Wp, *Wt;
  const double* P;  // [nspin][ntot*ntot]
  double* H;        // [nspin][ntot*ntot]
  double *Rq, *F, *VGT, *partial;
  int nblk_p;  // blocks over angular points in stage B
};

// ---- A: radial density transforms --------------------------------------------------------------------------
__global__ void g2d_radial_density(G2DParams p) {
  extern __shared__ double sm[];
  const int ab = blockIdx.x, e = blockIdx.y, s = blockIdx.z;
  const int a = ab / p.nang, b = ab % p.nang;
  const int np = p.npe[e], f0 = p.first[e], nq = p.nq, npm = p.np_max;
  double* Ps = sm;
  double* chi = Ps + npm * npm;
  double* dchi = chi + npm * nq;
  const double* Pg = p.P + (size_t)s * p.ntot * p.ntot + ((size_t)a * p.nrad + f0) + ((size_t)b * p.nrad + f0) * p.ntot;
  int nz = 0;
  for (int i = threadIdx.x; i < np * np; i += blockDim.x) {
    const int k = i % np, l = i / np;
    const double v = Pg[k + (size_t)l * p.ntot];
    Ps[k + l * npm] = v;
    nz |= (v != 0.0);
  }
  const bool want_d = p.fam >= xc::FAM_GGA;
  for (int i = threadIdx.x; i < npm * nq; i += blockDim.x) {
    chi[i] = p.chi[(size_t)e * npm * nq + i];
    if (want_d) dchi[i] = p.dchi[(size_t)e * npm * nq + i];
  }
  nz = __syncthreads_or(nz);
  const int q = threadIdx.x;
  if (q >= nq) return;
  double R = 0.0, Rr = 0.0, Rrr = 0.0;
  if (nz) {
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
  }
  double* out = p.Rq + ((size_t)(s * p.nel + e) * 3) * p.npair * nq + (size_t)ab * nq + q;
  out[0] = R;
  if (want_d) {
    out[(size_t)p.npair * nq] = Rr;
    out[(size_t)2 * p.npair * nq] = Rrr;
  }
}

// ---- B: densities at the points, XC, weighted potential fields ------------------------------------------------
// fields per spin: 0 a = w vrho | 1..3 g_c = w (2 vs g_c [+ vs_ab g'_c]) / h_c | 4..6 t_c = w vtau / (2 h_c^2)
template <int FAM, int NSPIN>
__global__ void g2d_point_xc(G2DParams p) {
  const int q = threadIdx.x, ip = blockIdx.x * blockDim.y + threadIdx.y, e = blockIdx.y;
  const int nq = p.nq, npair = p.npair;
  const bool live = q < nq && ip < p.npts;
  double S0[NSPIN] = {}, S0r[NSPIN] = {}, S0rr[NSPIN] = {}, S1[NSPIN] = {}, S2[NSPIN] = {}, S3[NSPIN] = {}, S4[NSPIN] = {};
  if (live) {
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

__global__ void g2d_reduce_scalars(const double* partial, int n, double* out) {
  // one thread per scalar, fixed order
  const int i = threadIdx.x;
  if (i < 3) {
    double v = 0.0;
    for (int k = 0; k < n; k++) v += partial[(size_t)k * 3 + i];
    out[i] = v;
  }
}

// ---- C: angular contraction of the fields ----------------------------------------------------------------------
__global__ void g2d_angular_fock(G2DParams p) {
  const int ab = blockIdx.x, e = blockIdx.y, s = blockIdx.z, q = threadIdx.x;
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
  const int ab = blockIdx.x, e = 2 * blockIdx.y + parity, s = blockIdx.z;
  if (e >= p.nel) return;
  const int a = ab / p.nang, b = ab % p.nang;
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
  double* Hg = p.H + (size_t)s * p.ntot * p.ntot + ((size_t)a * p.nrad + f0) + ((size_t)b * p.nrad + f0) * p.ntot;
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
    Hg[k + (size_t)l * p.ntot] += h;
  }
}

// ---- host side -----------------------------------------------------------------------------------------------
void Grid2D::release() {
  for (double** q : {&dChi, &dDChi, &dR, &dWr, &dCth, &dWang, &dWp, &dWt})
    if (*q) { cudaFree(*q); *q = nullptr; }
  for (int** q : {&dFirst, &dNpe})
    if (*q) { cudaFree(*q); *q = nullptr; }
  P.release(); H.release(); Rq.release(); F.release(); VGT.release(); partial.release(); scal.release();
}

template <class T>
static int up(T** dst, const std::vector<T>& src) {
  HFB_CUDA_OK(cudaMalloc(dst, std::max<size_t>(src.size(), 1) * sizeof(T)));
  HFB_CUDA_OK(cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}

int Grid2D::setup(int kind_, double Rhalf_, const std::vector<int>& lval, const std::vector<int>& mval, int nel_,
                  const std::vector<int>& first_, const std::vector<int>& last_, int nq_, const double* r, const double* wrad,
                  const double* bf, const double* df, int npts_, const double* cth, const double* wang, const double* Y,
                  const double* dY) {
  release();
  kind = kind_; Rhalf = Rhalf_; nang = (int)lval.size(); npair = nang * nang; nel = nel_; nq = nq_; npts = npts_;
  first = first_;
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
  // angular pair tables
  typedef std::complex<double> cx;
  std::vector<double> Wp((size_t)npts * 5 * npair), Wt((size_t)5 * npair * npts);
#pragma omp parallel for schedule(static)
  for (int ip = 0; ip < npts; ip++) {
    std::vector<double> w1((size_t)npair), w2((size_t)npair);
    for (int a = 0; a < nang; a++)
      for (int b = 0; b < nang; b++) {
        const cx ya(Y[2 * (a + (size_t)nang * ip)], Y[2 * (a + (size_t)nang * ip) + 1]);
        const cx yb(Y[2 * (b + (size_t)nang * ip)], Y[2 * (b + (size_t)nang * ip) + 1]);
        const cx da(dY[2 * (a + (size_t)nang * ip)], dY[2 * (a + (size_t)nang * ip) + 1]);
        const cx db(dY[2 * (b + (size_t)nang * ip)], dY[2 * (b + (size_t)nang * ip) + 1]);
        const cx z = ya * std::conj(yb);
        const int ab = a * nang + b;
        double* W = &Wp[(size_t)ip * 5 * npair];
        W[ab] = z.real();
        W[npair + ab] = w1[ab] = (da * std::conj(yb)).real();
        W[2 * npair + ab] = w2[ab] = -mval[a] * z.imag();
        W[3 * npair + ab] = (da * std::conj(db)).real();
        W[4 * npair + ab] = (double)(mval[a] * mval[b]) * z.real();
      }
    for (int a = 0; a < nang; a++)
      for (int b = 0; b < nang; b++) {
        const int ab = a * nang + b, ba = b * nang + a;
        const double* W = &Wp[(size_t)ip * 5 * npair];
        Wt[((size_t)0 * npair + ab) * npts + ip] = W[ab];
        Wt[((size_t)1 * npair + ab) * npts + ip] = w1[ab] + w1[ba];
        Wt[((size_t)2 * npair + ab) * npts + ip] = w2[ab] + w2[ba];
        Wt[((size_t)3 * npair + ab) * npts + ip] = W[3 * npair + ab];
        Wt[((size_t)4 * npair + ab) * npts + ip] = W[4 * npair + ab];
      }
  }
  int rc;
  if ((rc = up(&dChi, chi)) || (rc = up(&dDChi, dchi)) || (rc = up(&dR, std::vector<double>(r, r + (size_t)nel * nq))) ||
      (rc = up(&dWr, std::vector<double>(wrad, wrad + (size_t)nel * nq))) || (rc = up(&dCth, std::vector<double>(cth, cth + npts))) ||
      (rc = up(&dWang, std::vector<double>(wang, wang + npts))) || (rc = up(&dWp, Wp)) || (rc = up(&dWt, Wt)) ||
      (rc = up(&dFirst, first)) || (rc = up(&dNpe, npe)))
    return rc;
  ready = true;
  return 0;
}

template <int FAM>
static void launch_point(int nspin, dim3 grid, dim3 block, cudaStream_t st, const G2DParams& p) {
  if (nspin == 1) g2d_point_xc<FAM, 1><<<grid, block, 0, st>>>(p);
  else g2d_point_xc<FAM, 2><<<grid, block, 0, st>>>(p);
}

// Pa/Pb, Ha/Hb: HOST dense (ntot x ntot) matrices in the full product space with leading dimensions ldp / ldh.
int Grid2D::eval_fxc(int fam, int x_func, int c_func, int nspin, const double* Pa, const double* Pb, int ldp, double dens_thr,
                     double* Ha, double* Hb, int ldh, int beta, double* scalars3, cudaStream_t st) {
  const int ntot = nang * nrad;
  const size_t n2 = (size_t)ntot * ntot;
  const int nqb = (nq + 31) / 32 * 32;
  if (nqb > 1024) { set_error("grid2d: more than 1024 radial quadrature points per element are not supported"); return HFB200_EINVAL; }
  const int PB = std::max(1, 256 / nqb);
  const int nblk_p = (npts + PB - 1) / PB;
  launches = 0;
  int rc;
  if ((rc = P.reserve(n2 * nspin)) || (rc = H.reserve(n2 * nspin)) || (rc = Rq.reserve((size_t)nspin * nel * 3 * npair * nq)) ||
      (rc = F.reserve((size_t)nspin * nel * 7 * npts * nq)) || (rc = VGT.reserve((size_t)nspin * nel * 3 * npair * nq)) ||
      (rc = partial.reserve((size_t)nel * nblk_p * 3)) || (rc = scal.reserve(4)))
    return rc;
  for (int s = 0; s < nspin; s++)
    HFB_CUDA_OK(cudaMemcpy2DAsync(P.p + s * n2, (size_t)ntot * sizeof(double), s ? Pb : Pa, (size_t)ldp * sizeof(double),
                                  (size_t)ntot * sizeof(double), ntot, cudaMemcpyHostToDevice, st));
  HFB_CUDA_OK(cudaMemsetAsync(H.p, 0, n2 * nspin * sizeof(double), st));
  G2DParams p;
  p.kind = kind; p.nang = nang; p.npair = npair; p.nel = nel; p.nq = nq; p.np_max = np_max; p.nrad = nrad; p.npts = npts; p.ntot = ntot;
  p.fam = fam; p.x_func = x_func; p.c_func = c_func; p.Rhalf = Rhalf; p.dens_thr = dens_thr;
  p.first = dFirst; p.npe = dNpe; p.chi = dChi; p.dchi = dDChi; p.r = dR; p.wr = dWr; p.cth = dCth; p.wang = dWang; p.Wp = dWp; p.Wt = dWt;
  p.P = P.p; p.H = H.p; p.Rq = Rq.p; p.F = F.p; p.VGT = VGT.p; p.partial = partial.p; p.nblk_p = nblk_p;

  const size_t smA = ((size_t)np_max * np_max + 2 * (size_t)np_max * nq) * sizeof(double);
  const size_t smD = (2 * (size_t)np_max * nq + 3 * (size_t)nq) * sizeof(double);
  if (smA > 227 * 1024 || smD > 227 * 1024) { set_error("grid2d: element tables exceed shared memory"); return HFB200_EINVAL; }
  HFB_CUDA_OK(cudaFuncSetAttribute(g2d_radial_density, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smA));
  HFB_CUDA_OK(cudaFuncSetAttribute(g2d_radial_fock, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smD));
  g2d_radial_density<<<dim3(npair, nel, nspin), nqb, smA, st>>>(p);
  HFB_CUDA_OK(cudaGetLastError());
  const dim3 gB(nblk_p, nel), bB(nqb, PB);
  if (fam == xc::FAM_LDA) launch_point<xc::FAM_LDA>(nspin, gB, bB, st, p);
  else if (fam == xc::FAM_GGA) launch_point<xc::FAM_GGA>(nspin, gB, bB, st, p);
  else launch_point<xc::FAM_MGGA>(nspin, gB, bB, st, p);
  HFB_CUDA_OK(cudaGetLastError());
  g2d_reduce_scalars<<<1, 32, 0, st>>>(partial.p, nel * nblk_p, scal.p);
  HFB_CUDA_OK(cudaGetLastError());
  const int nsp_out = (nspin == 2 && !beta) ? 1 : nspin;
  g2d_angular_fock<<<dim3(npair, nel, nsp_out), nqb, 0, st>>>(p);
  HFB_CUDA_OK(cudaGetLastError());
  for (int parity = 0; parity < 2; parity++) {
    const int ne = (nel - parity + 1) / 2;
    if (ne < 1) continue;
    g2d_radial_fock<<<dim3(npair, ne, nsp_out), 256, smD, st>>>(p, parity);
    HFB_CUDA_OK(cudaGetLastError());
    launches++;
  }
  launches += 4;
  for (int s = 0; s < nsp_out; s++)
    HFB_CUDA_OK(cudaMemcpy2DAsync(s ? Hb : Ha, (size_t)ldh * sizeof(double), H.p + s * n2, (size_t)ntot * sizeof(double),
                                  (size_t)ntot * sizeof(double), ntot, cudaMemcpyDeviceToHost, st));
  HFB_CUDA_OK(cudaMemcpyAsync(scalars3, scal.p, 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
  HFB_CUDA_OK(cudaStreamSynchronize(st));
  return 0;
}

}  // namespace hfb
