// ORACLE (test infrastructure, not product code): CPU restatement of the reference's XC Fock build on the
// angular x radial product grid of the `atomic` and `diatomic` programs, following the reference's own
// loop structure (complex basis-function matrices, Pv = P conj(bf), per-point dots, BLAS-style
// increments).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this.
//
// Functions restated (read-only tree /root/reference):
//   atomic::dftgrid::DFTGrid::eval_Fxc          src/atomic/dftgrid.cpp:810-928   (R and U)
//     DFTGridWorker::compute_bf                 src/atomic/dftgrid.cpp:710-796   (metric 1, r, r sin(theta))
//     DFTGridWorker::update_density             src/atomic/dftgrid.cpp:51-232
//     compute_Nel / compute_Ekin / eval_Exc     src/atomic/dftgrid.cpp:234-275,495-501
//     DFTGridWorker::eval_Fxc                   src/atomic/dftgrid.cpp:523-673, increment_lda/gga dftgrid.h:190-250
//     TwoDBasis::eval_bf / eval_df / bf_list    src/atomic/TwoDBasis.cpp:1368-1421,1450-1470
//   diatomic::dftgrid::DFTGrid::eval_Fxc        src/diatomic/dftgrid.cpp:769-856 with compute_bf :669-755
//     (prolate metric h_mu = h_nu = Rh sqrt(sinh^2 mu + sin^2 nu), h_phi = Rh sinh mu sin nu; the reference
//      visits one mu shell at a time to save memory, here a whole element at once: the same sums)
// libxc calls are replaced by oracle/xc_oracle.hpp (hotpath_oracle.cpp's oracle_xc_unpol / _pol).
// Spherical harmonics Y_l^m and Y_l^{m+1} at the angular points are INPUTS (the tests take them from
// scipy.special.sph_harm_y, independent of the product's recursion); the theta derivative is formed
// here with the reference's formula (TwoDBasis.cpp:1407-1419).
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstddef>
#include <vector>

typedef std::complex<double> cx;
typedef std::vector<double> vec;
typedef std::vector<cx> cvec;

extern "C" {
void oracle_xc_unpol(int func, int n, double thr, const double* rho, const double* sigma, const double* tau,
                     double* exc, double* vrho, double* vsigma, double* vtau);
void oracle_xc_pol(int func, int n, double thr, const double* rho, const double* sigma, const double* tau,
                   double* exc, double* vrho, double* vsigma, double* vtau);
int oracle_xc_family(int func);
}

namespace {

// H += Re[(f diag(v)) f^H]   (increment_lda, atomic/dftgrid.h:190-208)
void increment_lda(vec& H, int n, int npts, const vec& v, const cvec& f) {
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
  for (int j = 0; j < n; j++)
    for (int i = 0; i < n; i++) {
      double s = 0.0;
      for (int p = 0; p < npts; p++) s += v[p] * std::real(f[i + (size_t)p * n] * std::conj(f[j + (size_t)p * n]));
      H[i + (size_t)j * n] += s;
    }
}

// H += Re[gamma f^H + f gamma^H], gamma = sum_c f_c diag(g_c)   (increment_gga, atomic/dftgrid.h:211-250)
void increment_gga(vec& H, int n, int npts, const vec g[3], const cvec& f, const cvec& fr, const cvec& ft, const cvec& fp) {
  cvec gamma((size_t)n * npts);
  for (int p = 0; p < npts; p++)
    for (int i = 0; i < n; i++) {
      const size_t k = i + (size_t)p * n;
      gamma[k] = fr[k] * g[0][p] + ft[k] * g[1][p] + fp[k] * g[2][p];
    }
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
  for (int j = 0; j < n; j++)
    for (int i = 0; i < n; i++) {
      double s = 0.0;
      for (int p = 0; p < npts; p++)
        s += std::real(gamma[i + (size_t)p * n] * std::conj(f[j + (size_t)p * n]) + f[i + (size_t)p * n] * std::conj(gamma[j + (size_t)p * n]));
      H[i + (size_t)j * n] += s;
    }
}

// Pv = P conj(f)   (n x npts)
void pv_product(cvec& Pv, int n, int npts, const vec& P, const cvec& f) {
  Pv.assign((size_t)n * npts, cx(0.0));
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
  for (int p = 0; p < npts; p++)
    for (int j = 0; j < n; j++) {
      const cx c = std::conj(f[j + (size_t)p * n]);
      for (int i = 0; i < n; i++) Pv[i + (size_t)p * n] += P[i + (size_t)j * n] * c;
    }
}

// arma::dot(a.col(p), b.col(p)) -- no conjugation
inline cx dotu(const cvec& a, const cvec& b, int n, int p) {
  cx s(0.0);
  for (int i = 0; i < n; i++) s += a[i + (size_t)p * n] * b[i + (size_t)p * n];
  return s;
}

}  // namespace

extern "C" {

// kind: 0 atomic (spherical polar metric), 1 diatomic (prolate spheroidal metric, Rhalf).
// bf/df: per element (nq x np_e) column-major, concatenated: radial.get_bf(iel) / get_df(iel).
// Y, Y1: (nang x nangpts) complex as (re,im) pairs, [a + nang*ia]: Y_{l_a}^{m_a} and Y_{l_a}^{m_a+1} at (cth,phi)(ia).
// P*, H*: (nang*nrad)^2 column-major in the full product ("dummy") space; H* are overwritten.
int oracle_grid2d_eval_fxc(int kind, double Rhalf, int nang, const int* lval, const int* mval, int nel, int nq,
                           const int* first, const int* last, int nrad, const double* r, const double* wrad,
                           const double* bfr, const double* dfr, int nangpts, const double* cth, const double* phi,
                           const double* wang, const double* Y, const double* Y1, int x_func, int c_func, int nspin,
                           const double* Pa, const double* Pb, double thr, double* Ha, double* Hb, int beta, double* Exc,
                           double* Nel, double* Ekin) {
  const size_t ntot = (size_t)nang * nrad;
  std::fill(Ha, Ha + ntot * ntot, 0.0);
  if (nspin == 2 && Hb) std::fill(Hb, Hb + ntot * ntot, 0.0);
  const int famx = oracle_xc_family(x_func), famc = oracle_xc_family(c_func);
  if ((x_func && !famx) || (c_func && !famc)) return -3;
  const int fam = std::max(famx, famc);
  const bool do_grad = fam >= 2, do_tau = fam >= 4;
  std::vector<size_t> off(nel + 1, 0);
  for (int e = 0; e < nel; e++) off[e + 1] = off[e] + (size_t)(last[e] - first[e] + 1) * nq;
  double exc = 0.0, nelsum = 0.0, ekin = 0.0;

  for (int iel = 0; iel < nel; iel++) {
    const int f0 = first[iel], np = last[iel] - f0 + 1, n = nang * np, npts = nangpts * nq;
    const double* B = bfr + off[iel];
    const double* dB = dfr + off[iel];
    const double* rr = r + (size_t)iel * nq;
    const double* wr = wrad + (size_t)iel * nq;
    // ---- compute_bf: weights, scale factors, basis functions (point index ia*nq + ir) ----
    vec wtot(npts), sr(npts), st(npts), sp(npts);
    for (int ia = 0; ia < nangpts; ia++) {
      const double sth = std::sqrt(1.0 - cth[ia] * cth[ia]);
      for (int ir = 0; ir < nq; ir++) {
        const int p = ia * nq + ir;
        if (kind == 0) {
          sr[p] = 1.0; st[p] = rr[ir]; sp[p] = rr[ir] * sth;
          wtot[p] = wang[ia] * wr[ir] * rr[ir] * rr[ir];
        } else {
          const double sh = std::sinh(rr[ir]);
          sr[p] = st[p] = Rhalf * std::sqrt(sh * sh + sth * sth);
          sp[p] = Rhalf * sh * sth;
          wtot[p] = wang[ia] * wr[ir] * std::pow(Rhalf, 3) * sh * (sh * sh + sth * sth);
        }
      }
    }
    cvec bf((size_t)n * npts), bf_r, bf_t, bf_p;
    if (do_grad) { bf_r.resize(bf.size()); bf_t.resize(bf.size()); bf_p.resize(bf.size()); }
    for (int ia = 0; ia < nangpts; ia++) {
      const double cotth = cth[ia] / std::sqrt(1.0 - cth[ia] * cth[ia]);
      for (int a = 0; a < nang; a++) {
        const cx sph(Y[2 * (a + (size_t)nang * ia)], Y[2 * (a + (size_t)nang * ia) + 1]);
        const int l = lval[a], m = mval[a];
        cx angfac = (double)m * cotth * sph;
        if (m < l) {
          const cx sph1(Y1[2 * (a + (size_t)nang * ia)], Y1[2 * (a + (size_t)nang * ia) + 1]);
          angfac += std::sqrt((double)((l - m) * (l + m + 1))) * std::exp(cx(0.0, -phi[ia])) * sph1;
        }
        for (int k = 0; k < np; k++)
          for (int ir = 0; ir < nq; ir++) {
            const size_t idx = (size_t)(a * np + k) + (size_t)(ia * nq + ir) * n;
            bf[idx] = sph * B[ir + (size_t)k * nq];
            if (do_grad) {
              bf_r[idx] = sph * dB[ir + (size_t)k * nq];
              bf_p[idx] = cx(0.0, (double)m) * sph * B[ir + (size_t)k * nq];
              bf_t[idx] = angfac * B[ir + (size_t)k * nq];
            }
          }
      }
    }
    // ---- update_density ----
    const int nsig = nspin == 1 ? 1 : 3;
    vec rho((size_t)nspin * npts, 0.0), grho((size_t)3 * nspin * npts, 0.0), sigma((size_t)nsig * npts, 0.0), tau((size_t)nspin * npts, 0.0);
    for (int s = 0; s < nspin; s++) {
      const double* Pfull = s ? Pb : Pa;
      vec P((size_t)n * n);
      for (int b = 0; b < nang; b++)
        for (int l = 0; l < np; l++)
          for (int a = 0; a < nang; a++)
            for (int k = 0; k < np; k++)
              P[(a * np + k) + (size_t)(b * np + l) * n] = Pfull[((size_t)a * nrad + f0 + k) + ((size_t)b * nrad + f0 + l) * ntot];
      cvec Pv, Pv_r, Pv_t, Pv_p;
      pv_product(Pv, n, npts, P, bf);
      for (int p = 0; p < npts; p++) rho[(size_t)nspin * p + s] = std::real(dotu(Pv, bf, n, p));
      if (do_grad)
        for (int p = 0; p < npts; p++) {
          grho[(size_t)(3 * nspin) * p + 3 * s + 0] = 2.0 * std::real(dotu(Pv, bf_r, n, p)) / sr[p];
          grho[(size_t)(3 * nspin) * p + 3 * s + 1] = 2.0 * std::real(dotu(Pv, bf_t, n, p)) / st[p];
          grho[(size_t)(3 * nspin) * p + 3 * s + 2] = 2.0 * std::real(dotu(Pv, bf_p, n, p)) / sp[p];
        }
      if (do_tau) {
        pv_product(Pv_r, n, npts, P, bf_r);
        pv_product(Pv_t, n, npts, P, bf_t);
        pv_product(Pv_p, n, npts, P, bf_p);
        for (int p = 0; p < npts; p++) {
          const double kr = std::real(dotu(Pv_r, bf_r, n, p)) / (sr[p] * sr[p]);
          const double kt = std::real(dotu(Pv_t, bf_t, n, p)) / (st[p] * st[p]);
          const double kp = std::real(dotu(Pv_p, bf_p, n, p)) / (sp[p] * sp[p]);
          tau[(size_t)nspin * p + s] = 0.5 * (kr + kt + kp);
        }
      }
    }
    if (do_grad)
      for (int p = 0; p < npts; p++) {
        const double* g = &grho[(size_t)(3 * nspin) * p];
        if (nspin == 1) sigma[p] = g[0] * g[0] + g[1] * g[1] + g[2] * g[2];
        else {
          sigma[3 * (size_t)p + 0] = g[0] * g[0] + g[1] * g[1] + g[2] * g[2];
          sigma[3 * (size_t)p + 1] = g[0] * g[3] + g[1] * g[4] + g[2] * g[5];
          sigma[3 * (size_t)p + 2] = g[3] * g[3] + g[4] * g[4] + g[5] * g[5];
        }
      }
    for (int p = 0; p < npts; p++) {
      const double d = nspin == 1 ? rho[p] : rho[2 * (size_t)p] + rho[2 * (size_t)p + 1];
      nelsum += wtot[p] * d;
      if (do_tau) ekin += wtot[p] * (nspin == 1 ? tau[p] : tau[2 * (size_t)p] + tau[2 * (size_t)p + 1]);
    }
    // ---- compute_xc (x then c, results added) ----
    vec e(npts, 0.0), vxc((size_t)nspin * npts, 0.0), vsigma((size_t)nsig * npts, 0.0), vtau((size_t)nspin * npts, 0.0);
    for (int which = 0; which < 2; which++) {
      const int id = which ? c_func : x_func;
      if (id <= 0) continue;
      if (nspin == 1) oracle_xc_unpol(id, npts, thr, rho.data(), sigma.data(), tau.data(), e.data(), vxc.data(), vsigma.data(), vtau.data());
      else oracle_xc_pol(id, npts, thr, rho.data(), sigma.data(), tau.data(), e.data(), vxc.data(), vsigma.data(), vtau.data());
    }
    for (int p = 0; p < npts; p++) exc += wtot[p] * e[p] * (nspin == 1 ? rho[p] : rho[2 * (size_t)p] + rho[2 * (size_t)p + 1]);
    // ---- eval_Fxc ----
    for (int s = 0; s < nspin; s++) {
      if (s == 1 && !beta) continue;
      vec H((size_t)n * n, 0.0), v(npts);
      for (int p = 0; p < npts; p++) v[p] = vxc[(size_t)nspin * p + s] * wtot[p];
      increment_lda(H, n, npts, v, bf);
      if (do_grad) {
        vec g[3] = {vec(npts), vec(npts), vec(npts)};
        for (int p = 0; p < npts; p++) {
          const double* gr = &grho[(size_t)(3 * nspin) * p];
          const double sc[3] = {sr[p], st[p], sp[p]};
          for (int c = 0; c < 3; c++) {
            if (nspin == 1) g[c][p] = gr[c] * 2.0 * wtot[p] * vsigma[p] / sc[c];
            else {
              const int o = 1 - s;
              g[c][p] = wtot[p] * (2.0 * vsigma[3 * (size_t)p + 2 * s] * gr[3 * s + c] + vsigma[3 * (size_t)p + 1] * gr[3 * o + c]) / sc[c];
            }
          }
        }
        increment_gga(H, n, npts, g, bf, bf_r, bf_t, bf_p);
      }
      if (do_tau) {
        vec vr(npts), vt(npts), vp(npts);
        for (int p = 0; p < npts; p++) {
          const double vtl = 0.5 * vtau[(size_t)nspin * p + s] * wtot[p];
          vr[p] = vtl / (sr[p] * sr[p]);
          vt[p] = vtl / (st[p] * st[p]);
          vp[p] = vtl / (sp[p] * sp[p]);
        }
        increment_lda(H, n, npts, vr, bf_r);
        increment_lda(H, n, npts, vt, bf_t);
        increment_lda(H, n, npts, vp, bf_p);
      }
      double* Ho = s ? Hb : Ha;
      for (int b = 0; b < nang; b++)
        for (int l = 0; l < np; l++)
          for (int a = 0; a < nang; a++)
            for (int k = 0; k < np; k++)
              Ho[((size_t)a * nrad + f0 + k) + ((size_t)b * nrad + f0 + l) * ntot] += H[(a * np + k) + (size_t)(b * np + l) * n];
    }
  }
  *Exc = exc;
  *Nel = nelsum;
  *Ekin = ekin;
  return 0;
}

}  // extern "C"
