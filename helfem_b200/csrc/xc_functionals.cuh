// Device XC functionals: Slater exchange, PW92 correlation, PBE x+c, r2SCAN x+c.
//
// Replaces, for these functionals only, the libxc calls the reference makes in
//   /root/reference/src/sadatom/dftgrid.cpp:306-486   (DFTGridWorker::compute_xc)
//   /root/reference/src/atomic/dftgrid.cpp:378-493
//   /root/reference/src/diatomic/dftgrid.cpp:343-458
// libxc itself is not vendored in the reference tree (version unpinned upstream), so the
// formulas below restate the published functional definitions with libxc's conventions:
// inputs (rho, sigma=|grad rho|^2, tau) spin-fastest, outputs exc per particle and the
// partial derivatives of the energy density per volume.
//
// Every functional is one energy expression templated on the scalar type T (double for the
// energy only, hfb::xc::Dual<N> for energy + derivatives).  There is no CPU fallback in the
// product: this header is compiled for the host only by the unit tests, to compare the
// same source against the oracle and the high-precision golden vectors without a GPU.
#pragma once
#include "dual.cuh"

namespace hfb {
namespace xc {

// Functional identifiers (the numeric values are those of libxc's registry as commonly
// shipped; drivers should map by NAME through hfb200_xc_id(), see include/helfem_b200.h).
enum XcId {
  XC_NONE = 0,
  XC_LDA_X = 1,
  XC_LDA_C_PW = 12,
  XC_LDA_C_PW_MOD = 13,
  XC_GGA_X_PBE = 101,
  XC_GGA_C_PBE = 130,
  XC_MGGA_X_R2SCAN = 497,
  XC_MGGA_C_R2SCAN = 498
};

enum XcFamily { FAM_NONE = 0, FAM_LDA = 1, FAM_GGA = 2, FAM_MGGA = 4 };

HFB_HD int xc_family(int id) {
  switch (id) {
    case XC_LDA_X: case XC_LDA_C_PW: case XC_LDA_C_PW_MOD: return FAM_LDA;
    case XC_GGA_X_PBE: case XC_GGA_C_PBE: return FAM_GGA;
    case XC_MGGA_X_R2SCAN: case XC_MGGA_C_R2SCAN: return FAM_MGGA;
    default: return FAM_NONE;
  }
}
HFB_HD bool xc_is_exchange(int id) { return id == XC_LDA_X || id == XC_GGA_X_PBE || id == XC_MGGA_X_R2SCAN; }

struct XcThresholds {
  double dens;   // xc_func_set_dens_threshold value (reference: --dftthr, default 1e-12)
  double zeta;   // DBL_EPSILON
  double sigma;  // floor on sigma
  double tau;    // floor on tau
};

HFB_HD XcThresholds make_thresholds(double dens_thr) {
  XcThresholds t;
  t.dens = dens_thr;
  t.zeta = 2.220446049250313e-16;
  t.sigma = ::pow(dens_thr, 4.0 / 3.0);
  t.sigma = t.sigma * t.sigma;
  t.tau = 1e-20;
  return t;
}

// ---- constants -----------------------------------------------------------
#define HFB_PI 3.14159265358979323846
#define HFB_CBRT_3PI2 3.0936677262801359310   /* (3 pi^2)^(1/3) */
#define HFB_RS_FACTOR 0.62035049089940001667  /* (3/(4 pi))^(1/3) */
#define HFB_GAMMA_C 0.031090690869654895035   /* (1-ln 2)/pi^2 */
#define HFB_BETA_PBE 0.06672455060314922
#define HFB_KAPPA_PBE 0.8040
#define HFB_MU_PBE 0.21951497276451709   /* beta*pi^2/3 */
#define HFB_TWO43M2 0.51984209978974632953 /* 2^(4/3)-2 */

// (1+z)^p with libxc's zeta threshold: below the threshold the value is frozen.
template <class T> HFB_HD T opz_pow(const T& opz, double p, double zthr) {
  if (val(opz) <= zthr) return constant_like(opz, ::pow(zthr, p));
  return pow(opz, p);
}

// ---- PW92 ---------------------------------------------------------------
struct Pw92Par { double a[3], fz20; };
HFB_HD Pw92Par pw92_par(bool mod) {
  Pw92Par p;
  if (mod) { p.a[0] = 0.0310907; p.a[1] = 0.01554535; p.a[2] = 0.0168869; p.fz20 = 1.709920934161365617563962776245; }
  else     { p.a[0] = 0.031091;  p.a[1] = 0.015545;   p.a[2] = 0.016887;  p.fz20 = 1.709921; }
  return p;
}

// G(rs) = -2A(1+a1 rs) ln(1 + 1/(2A(b1 rs^1/2 + b2 rs + b3 rs^3/2 + b4 rs^2))) and dG/drs
template <class T> HFB_HD void pw92_g(int k, double A, const T& rs, const T& srs, T& g, T* dg) {
  const double a1[3] = {0.21370, 0.20548, 0.11125};
  const double b1[3] = {7.5957, 14.1189, 10.357};
  const double b2[3] = {3.5876, 6.1977, 3.6231};
  const double b3[3] = {1.6382, 3.3662, 0.88026};
  const double b4[3] = {0.49294, 0.62517, 0.49671};
  T q1 = 2.0 * A * (srs * (b1[k] + srs * (b2[k] + srs * (b3[k] + srs * b4[k]))));
  T lg = log1p(1.0 / q1);
  T pre = -2.0 * A * (1.0 + a1[k] * rs);
  g = pre * lg;
  if (dg) {
    T q1p = 2.0 * A * (0.5 * b1[k] / srs + b2[k] + 1.5 * b3[k] * srs + 2.0 * b4[k] * rs);
    *dg = (-2.0 * A * a1[k]) * lg - pre * q1p / (q1 * q1 + q1);
  }
}

template <class T> HFB_HD T f_zeta(const T& zeta, double zthr) {
  return (opz_pow(1.0 + zeta, 4.0 / 3.0, zthr) + opz_pow(1.0 - zeta, 4.0 / 3.0, zthr) - 2.0) * (1.0 / HFB_TWO43M2);
}

// eps_c^{PW92}(rs, zeta) per particle; optionally its explicit rs-derivative.
template <class T> HFB_HD T pw92_eps(const Pw92Par& par, const T& rs, const T& zeta, bool pol, double zthr, T* deps_drs) {
  T srs = sqrt(rs);
  T g0, dg0;
  pw92_g(0, par.a[0], rs, srs, g0, deps_drs ? &dg0 : (T*)0);
  if (!pol) {
    if (deps_drs) *deps_drs = dg0;
    return g0;
  }
  T g1, g2, dg1, dg2;
  pw92_g(1, par.a[1], rs, srs, g1, deps_drs ? &dg1 : (T*)0);
  pw92_g(2, par.a[2], rs, srs, g2, deps_drs ? &dg2 : (T*)0);
  T fz = f_zeta(zeta, zthr);
  T z2 = zeta * zeta;
  T z4 = z2 * z2;
  const double ifz20 = 1.0 / par.fz20;
  T e = g0 + z4 * fz * (g1 - g0 + g2 * ifz20) - fz * g2 * ifz20;
  if (deps_drs) *deps_drs = dg0 + z4 * fz * (dg1 - dg0 + dg2 * ifz20) - fz * dg2 * ifz20;
  return e;
}

// ---- r2SCAN switching functions -----------------------------------------
template <class T> HFB_HD T r2scan_f_alpha(const T& a, double c1, double c2, double d, const double* cp) {
  if (val(a) <= 0.0) return exp(-c1 * a / (1.0 - a));
  if (val(a) <= 2.5) {
    T r = constant_like(a, cp[7]);
#pragma unroll
    for (int i = 6; i >= 0; i--) r = r * a + cp[i];
    return r;
  }
  return -d * exp(c2 / (1.0 - a));
}

// ---- exchange: one spin channel, energy per volume -----------------------
// rho_s, sigma_ss, tau_s are the spin-channel quantities; the spin-scaling relation
// E_x[rho_a,rho_b] = (E_x[2 rho_a] + E_x[2 rho_b])/2 is applied here.
template <class T> HFB_HD T exchange_spin(int id, const T& rho_s, const T& sigma_ss, const T& tau_s) {
  T n = 2.0 * rho_s;
  T n13 = cbrt(n);
  T kf = HFB_CBRT_3PI2 * n13;
  T exunif = (-3.0 / (4.0 * HFB_PI)) * kf * n;  // energy per volume of the unpolarised gas
  if (id == XC_LDA_X) return 0.5 * exunif;

  T kfn = kf * n;
  T p = sigma_ss / (kfn * kfn);  // s^2 of the doubled density: |grad n|^2 = 4 sigma_ss
  if (id == XC_GGA_X_PBE) {
    T fx = 1.0 + HFB_KAPPA_PBE - (HFB_KAPPA_PBE * HFB_KAPPA_PBE) / (HFB_KAPPA_PBE + HFB_MU_PBE * p);
    return 0.5 * exunif * fx;
  }
  // r2SCAN exchange (Furness, Kaplan, Ning, Perdew, Sun, J. Phys. Chem. Lett. 11, 8208 (2020))
  const double eta = 0.001, k1 = 0.065, dp2 = 0.361, a1 = 4.9479, h0x = 1.174;
  const double c1x = 0.667, c2x = 0.8, dx = 1.24, mu_ak = 10.0 / 81.0;
  const double cx[8] = {1.0, -0.667, -0.4445555, -0.663086601049, 1.451297044490,
                        -0.887998041597, 0.234528941479, -0.023185843322};
  double c2 = 0.0;
  for (int i = 1; i < 8; i++) c2 += i * cx[i];
  c2 *= (h0x - 1.0);  // C_2x = f_x'(1) (h0x - 1) ~ -0.162742
  const double ceta = 20.0 / 27.0 + eta * 5.0 / 3.0;
  const double dp24 = dp2 * dp2 * dp2 * dp2;

  T tauw = sigma_ss / (2.0 * n);                                         // |grad n|^2/(8n)
  T tauunif = (0.3 * HFB_CBRT_3PI2 * HFB_CBRT_3PI2) * n13 * n13 * n;     // (3/10)(3pi^2)^(2/3) n^(5/3)
  T alpha = (2.0 * tau_s - tauw) / (tauunif + eta * tauw);
  T x = (ceta * c2 * exp(-(p * p) / dp24) + mu_ak) * p;
  T h1x = 1.0 + k1 - k1 / (1.0 + x / k1);
  T fa = r2scan_f_alpha(alpha, c1x, c2x, dx, cx);
  T gx;
  if (val(p) < 1e-60) gx = constant_like(p, 1.0);
  else gx = 1.0 - exp(-a1 / sqrt(sqrt(p)));
  T fx = (h1x + fa * (h0x - h1x)) * gx;
  return 0.5 * exunif * fx;
}

// ---- correlation: energy per volume as a function of (n, zeta, sigma, tau) ----
template <class T>
HFB_HD T correlation(int id, const T& n, const T& zeta, const T& sigma, const T& tau, bool pol, double zthr) {
  T n13 = cbrt(n);
  T rs = HFB_RS_FACTOR / n13;
  if (id == XC_LDA_C_PW || id == XC_LDA_C_PW_MOD) {
    Pw92Par par = pw92_par(id == XC_LDA_C_PW_MOD);
    return n * pw92_eps(par, rs, zeta, pol, zthr, (T*)0);
  }
  const Pw92Par par = pw92_par(true);
  T kf = HFB_CBRT_3PI2 * n13;
  T phi, phi3;
  if (pol) {
    phi = 0.5 * (opz_pow(1.0 + zeta, 2.0 / 3.0, zthr) + opz_pow(1.0 - zeta, 2.0 / 3.0, zthr));
    phi3 = phi * phi * phi;
  } else {
    phi = constant_like(n, 1.0);
    phi3 = phi;
  }
  if (id == XC_GGA_C_PBE) {
    T eps = pw92_eps(par, rs, zeta, pol, zthr, (T*)0);
    // t^2 = sigma / (4 phi^2 ks^2 n^2),  ks^2 = 4 kF/pi
    T t2 = sigma * (HFB_PI / 16.0) / (phi * phi * kf * n * n);
    const double bg = HFB_BETA_PBE / HFB_GAMMA_C;
    T A = bg / expm1(-eps / (HFB_GAMMA_C * phi3));
    T At2 = A * t2;
    T H = (HFB_GAMMA_C * phi3) * log1p(bg * t2 * (1.0 + At2) / (1.0 + At2 + At2 * At2));
    return n * (eps + H);
  }
  // r2SCAN correlation
  const double eta = 0.001, dp2 = 0.361;
  const double b1c = 0.0285764, b2c = 0.0889, b3c = 0.125541, chi_inf = 0.12802585262625815;
  const double c1c = 0.64, c2c = 1.5, dc = 0.7, gcnst = 2.3631;
  const double cc[8] = {1.0, -0.64, -0.4352, -1.535685604549, 3.061560252175,
                        -1.915710236206, 0.516884468372, -0.051848879792};
  double dfc2 = 0.0;
  for (int i = 1; i < 8; i++) dfc2 += i * cc[i];
  const double dp24 = dp2 * dp2 * dp2 * dp2;

  T kfn = kf * n;
  T p = sigma / (4.0 * kfn * kfn);  // s^2 of the total density
  T ds, gc;
  if (pol) {
    ds = 0.5 * (opz_pow(1.0 + zeta, 5.0 / 3.0, zthr) + opz_pow(1.0 - zeta, 5.0 / 3.0, zthr));
    T dxz = 0.5 * (opz_pow(1.0 + zeta, 4.0 / 3.0, zthr) + opz_pow(1.0 - zeta, 4.0 / 3.0, zthr));
    T z2 = zeta * zeta; T z4 = z2 * z2; T z12 = z4 * z4 * z4;
    gc = (1.0 - gcnst * (dxz - 1.0)) * (1.0 - z12);
  } else {
    ds = constant_like(n, 1.0);
    gc = constant_like(n, 1.0);
  }
  T tauw = sigma / (8.0 * n);
  T tauunif = (0.3 * HFB_CBRT_3PI2 * HFB_CBRT_3PI2) * n13 * n13 * n * ds;
  T alpha = (tau - tauw) / (tauunif + eta * tauw);

  // eps_c^1: PBE-like with rs-dependent beta and the r2SCAN gradient correction
  T deps1;
  T eps1lsda = pw92_eps(par, rs, zeta, pol, zthr, &deps1);
  T gphi3 = HFB_GAMMA_C * phi3;
  T w1 = expm1(-eps1lsda / gphi3);
  T beta = 0.066725 * (1.0 + 0.1 * rs) / (1.0 + 0.1778 * rs);
  // t^2 = sigma / (4 phi^2 ks^2 n^2) = (3 pi^2/16)^(2/3) s^2/(phi^2 rs)
  T t2 = sigma * (HFB_PI / 16.0) / (phi * phi * kf * n * n);
  T y = beta * t2 / (HFB_GAMMA_C * w1);
  // eps_c^0 pieces (needed for delta_y too)
  T srs = sqrt(rs);
  T den0 = 1.0 + b2c * srs + b3c * rs;
  T eps0lda = -b1c / den0;
  T deps0lda = b1c * (0.5 * b2c / srs + b3c) / (den0 * den0);
  T dy = (dfc2 / (27.0 * HFB_GAMMA_C)) / (ds * phi3 * w1) *
         (20.0 * rs * (deps0lda * gc - deps1) - 45.0 * eta * (eps0lda * gc - eps1lsda)) * p * exp(-(p * p) / dp24);
  T g1 = 1.0 / sqrt(sqrt(1.0 + 4.0 * (y - dy)));
  T H1 = gphi3 * log1p(w1 * (1.0 - g1));
  T eps1 = eps1lsda + H1;

  T w0 = expm1(-eps0lda / b1c);
  T ginf = 1.0 / sqrt(sqrt(1.0 + 4.0 * chi_inf * p));
  T H0 = b1c * log1p(w0 * (1.0 - ginf));
  T eps0 = (eps0lda + H0) * gc;

  T fa = r2scan_f_alpha(alpha, c1c, c2c, dc, cc);
  return n * (eps1 + fa * (eps0 - eps1));
}

}  // namespace xc
}  // namespace hfb
