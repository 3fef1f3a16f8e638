// ORACLE (test infrastructure, not product code): CPU restatement of the three XC
// functional pairs the hot path supports, with libxc's calling conventions.
//
// The reference obtains these numbers from libxc, a third-party dependency that is NOT in
// /root/reference (found by find_library, /root/reference/LibXcConfig.cmake:1-2; version
// unpinned; API use implies libxc >= 5).  Call sites:
// /root/reference/src/sadatom/dftgrid.cpp:349-398, src/atomic/dftgrid.cpp:422-492,
// src/diatomic/dftgrid.cpp:343-458.  None of the reference's tests pins an XC value:
// PARITY UNPINNED upstream.  This file restates the published functional forms
//   Slater/Dirac exchange;  Perdew-Wang 1992 correlation (lda_c_pw / lda_c_pw_mod constants);
//   PBE exchange and correlation (PRL 77, 3865);  r2SCAN (JPCL 11, 8208 and erratum)
// independently of the device code: plain energy-density expressions on a generic scalar,
// differentiated by the COMPLEX-STEP method (f'(x) = Im f(x+ih)/h, h=1e-40: no subtractive
// cancellation), whereas the device code uses forward-mode dual numbers.
#pragma once
#include <cmath>
#include <complex>

namespace oracle_xc {

typedef std::complex<double> cplx;

// ---- scalar helpers valid for double and complex (branching on the real part) ----
inline double re(double x) { return x; }
inline double re(const cplx& x) { return x.real(); }
inline double c_log1p(double x) { return std::log1p(x); }
inline cplx c_log1p(const cplx& z) {
  const double a = z.real(), b = z.imag();
  return cplx(0.5 * std::log1p(2 * a + a * a + b * b), std::atan2(b, 1.0 + a));
}
inline double c_expm1(double x) { return std::expm1(x); }
inline cplx c_expm1(const cplx& z) {
  const double a = z.real(), b = z.imag();
  const double s = std::sin(0.5 * b);
  return cplx(std::expm1(a) * std::cos(b) - 2.0 * s * s, std::exp(a) * std::sin(b));
}
template <class S> S c_pow(const S& x, double p) { return std::pow(x, p); }
template <class S> S c_exp(const S& x) { return std::exp(x); }
template <class S> S c_sqrt(const S& x) { return std::sqrt(x); }

enum { ID_LDA_X = 1, ID_LDA_C_PW = 12, ID_LDA_C_PW_MOD = 13, ID_GGA_X_PBE = 101, ID_GGA_C_PBE = 130,
       ID_MGGA_X_R2SCAN = 497, ID_MGGA_C_R2SCAN = 498 };
inline int family(int id) {
  if (id == ID_LDA_X || id == ID_LDA_C_PW || id == ID_LDA_C_PW_MOD) return 1;
  if (id == ID_GGA_X_PBE || id == ID_GGA_C_PBE) return 2;
  if (id == ID_MGGA_X_R2SCAN || id == ID_MGGA_C_R2SCAN) return 4;
  return 0;
}
inline bool is_x(int id) { return id == ID_LDA_X || id == ID_GGA_X_PBE || id == ID_MGGA_X_R2SCAN; }

static const double PI = 3.14159265358979323846;

// (1 +- zeta)^p with the zeta threshold of libxc
template <class S> S opz(const S& x, double p, double zthr) {
  if (re(x) <= zthr) return S(std::pow(zthr, p));
  return c_pow(x, p);
}

// ---- Perdew-Wang 92 ----
template <class S> S pw_G(const S& rs, double A, double a1, double b1, double b2, double b3, double b4) {
  S den = 2.0 * A * (b1 * c_sqrt(rs) + b2 * rs + b3 * rs * c_sqrt(rs) + b4 * rs * rs);
  return -2.0 * A * (1.0 + a1 * rs) * c_log1p(1.0 / den);
}
template <class S> S pw_dG(const S& rs, double A, double a1, double b1, double b2, double b3, double b4) {
  S q = 2.0 * A * (b1 * c_sqrt(rs) + b2 * rs + b3 * rs * c_sqrt(rs) + b4 * rs * rs);
  S qp = 2.0 * A * (0.5 * b1 / c_sqrt(rs) + b2 + 1.5 * b3 * c_sqrt(rs) + 2.0 * b4 * rs);
  return -2.0 * A * a1 * c_log1p(1.0 / q) + 2.0 * A * (1.0 + a1 * rs) * qp / (q * q + q);
}
// eps_c(rs, zeta); deriv=true returns d eps_c / d rs instead
template <class S> S pw_eps(const S& rs, const S& zeta, bool mod, bool deriv, double zthr) {
  const double A0 = mod ? 0.0310907 : 0.031091, A1 = mod ? 0.01554535 : 0.015545, A2 = mod ? 0.0168869 : 0.016887;
  const double fz20 = mod ? 1.709920934161365617563962776245 : 1.709921;
  S e0 = deriv ? pw_dG(rs, A0, 0.21370, 7.5957, 3.5876, 1.6382, 0.49294) : pw_G(rs, A0, 0.21370, 7.5957, 3.5876, 1.6382, 0.49294);
  S e1 = deriv ? pw_dG(rs, A1, 0.20548, 14.1189, 6.1977, 3.3662, 0.62517) : pw_G(rs, A1, 0.20548, 14.1189, 6.1977, 3.3662, 0.62517);
  S mac = deriv ? pw_dG(rs, A2, 0.11125, 10.357, 3.6231, 0.88026, 0.49671) : pw_G(rs, A2, 0.11125, 10.357, 3.6231, 0.88026, 0.49671);  // = -alpha_c
  S f = (opz(1.0 + zeta, 4.0 / 3.0, zthr) + opz(1.0 - zeta, 4.0 / 3.0, zthr) - 2.0) / (std::pow(2.0, 4.0 / 3.0) - 2.0);
  S z4 = zeta * zeta * zeta * zeta;
  // eps = e0 + alpha_c f/f''(0) (1 - z^4) + (e1 - e0) f z^4, alpha_c = -mac
  return e0 - mac * f / fz20 * (1.0 - z4) + (e1 - e0) * f * z4;
}

// ---- r2SCAN interpolation function ----
template <class S> S f_alpha(const S& a, double c1, double c2, double d, const double* c) {
  if (re(a) <= 0.0) return c_exp(-c1 * a / (1.0 - a));
  if (re(a) <= 2.5) {
    S s(0.0), ap(1.0);
    for (int i = 0; i < 8; i++) { s += c[i] * ap; ap *= a; }
    return s;
  }
  return -d * c_exp(c2 / (1.0 - a));
}
static const double R2_CX[8] = {1.0, -0.667, -0.4445555, -0.663086601049, 1.451297044490, -0.887998041597, 0.234528941479, -0.023185843322};
static const double R2_CC[8] = {1.0, -0.64, -0.4352, -1.535685604549, 3.061560252175, -1.915710236206, 0.516884468372, -0.051848879792};

// ---- exchange energy per volume of a spin-unpolarised density n with |grad n|^2 = g2, tau ----
template <class S> S ex_unpol(int id, const S& n, const S& g2, const S& tau) {
  const double kfc = std::cbrt(3.0 * PI * PI);
  S kf = kfc * c_pow(n, 1.0 / 3.0);
  S ex_lda = -3.0 / (4.0 * PI) * kf * n;
  if (id == ID_LDA_X) return ex_lda;
  S s2 = g2 / (4.0 * kf * kf * n * n);
  if (id == ID_GGA_X_PBE) {
    const double kappa = 0.804, mu = 0.06672455060314922 * PI * PI / 3.0;
    return ex_lda * (1.0 + kappa - kappa / (1.0 + mu * s2 / kappa));
  }
  // r2SCAN
  const double eta = 0.001, k1 = 0.065, dp2 = 0.361, a1 = 4.9479, h0x = 1.174, c1 = 0.667, c2 = 0.8, d = 1.24;
  double fp1 = 0.0;  // f_x'(alpha=1) from the polynomial
  for (int i = 1; i < 8; i++) fp1 += i * R2_CX[i];
  const double C2x = fp1 * (h0x - 1.0);
  const double Ceta = 20.0 / 27.0 + eta * 5.0 / 3.0;
  S tW = g2 / (8.0 * n);
  S tU = 0.3 * kfc * kfc * c_pow(n, 5.0 / 3.0);
  S al = (tau - tW) / (tU + eta * tW);
  S x = (Ceta * C2x * c_exp(-s2 * s2 / std::pow(dp2, 4)) + 10.0 / 81.0) * s2;
  S h1 = 1.0 + k1 - k1 / (1.0 + x / k1);
  S gx = (re(s2) < 1e-60) ? S(1.0) : S(1.0 - c_exp(-a1 / c_pow(s2, 0.25)));
  return ex_lda * (h1 + f_alpha(al, c1, c2, d, R2_CX) * (h0x - h1)) * gx;
}

// ---- correlation energy per volume, variables (n, zeta, g2=|grad n|^2, tau) ----
template <class S> S ec_general(int id, const S& n, const S& zeta, const S& g2, const S& tau, bool pol, double zthr) {
  const double gamma = (1.0 - std::log(2.0)) / (PI * PI);
  S rs = std::cbrt(3.0 / (4.0 * PI)) / c_pow(n, 1.0 / 3.0);
  S z = pol ? zeta : S(0.0);
  if (id == ID_LDA_C_PW || id == ID_LDA_C_PW_MOD) return n * pw_eps(rs, z, id == ID_LDA_C_PW_MOD, false, zthr);
  S phi = pol ? S(0.5 * (opz(1.0 + z, 2.0 / 3.0, zthr) + opz(1.0 - z, 2.0 / 3.0, zthr))) : S(1.0);
  S kf = std::cbrt(3.0 * PI * PI) * c_pow(n, 1.0 / 3.0);
  S ks2 = 4.0 * kf / PI;
  S t2 = g2 / (4.0 * phi * phi * ks2 * n * n);
  S gp3 = gamma * phi * phi * phi;
  if (id == ID_GGA_C_PBE) {
    const double beta = 0.06672455060314922;
    S eps = pw_eps(rs, z, true, false, zthr);
    S A = beta / gamma / c_expm1(-eps / gp3);
    S H = gp3 * c_log1p(beta / gamma * t2 * (1.0 + A * t2) / (1.0 + A * t2 + A * A * t2 * t2));
    return n * (eps + H);
  }
  // r2SCAN correlation
  const double eta = 0.001, dp2 = 0.361, b1c = 0.0285764, b2c = 0.0889, b3c = 0.125541, chi = 0.12802585262625815;
  const double c1 = 0.64, c2 = 1.5, d = 0.7;
  double dfc2 = 0.0;
  for (int i = 1; i < 8; i++) dfc2 += i * R2_CC[i];
  S s2 = g2 / (4.0 * kf * kf * n * n);
  S ds = pol ? S(0.5 * (opz(1.0 + z, 5.0 / 3.0, zthr) + opz(1.0 - z, 5.0 / 3.0, zthr))) : S(1.0);
  S dx = pol ? S(0.5 * (opz(1.0 + z, 4.0 / 3.0, zthr) + opz(1.0 - z, 4.0 / 3.0, zthr))) : S(1.0);
  S z2 = z * z, z4 = z2 * z2;
  S z12 = z4 * z4 * z4;
  S Gc = pol ? S((1.0 - 2.3631 * (dx - 1.0)) * (1.0 - z12)) : S(1.0);
  S tW = g2 / (8.0 * n);
  S tU = 0.3 * std::pow(3.0 * PI * PI, 2.0 / 3.0) * c_pow(n, 5.0 / 3.0) * ds;
  S al = (tau - tW) / (tU + eta * tW);
  // alpha = 1 limit
  S e1l = pw_eps(rs, z, true, false, zthr);
  S de1l = pw_eps(rs, z, true, true, zthr);
  S w1 = c_expm1(-e1l / gp3);
  S beta = 0.066725 * (1.0 + 0.1 * rs) / (1.0 + 0.1778 * rs);
  S y = beta / (gamma * w1) * t2;
  S e0l = -b1c / (1.0 + b2c * c_sqrt(rs) + b3c * rs);
  S de0l = b1c * (0.5 * b2c / c_sqrt(rs) + b3c) / ((1.0 + b2c * c_sqrt(rs) + b3c * rs) * (1.0 + b2c * c_sqrt(rs) + b3c * rs));
  S dy = dfc2 / (27.0 * gamma * ds * phi * phi * phi * w1) *
         (20.0 * rs * (de0l * Gc - de1l) - 45.0 * eta * (e0l * Gc - e1l)) * s2 * c_exp(-s2 * s2 / std::pow(dp2, 4));
  S H1 = gp3 * c_log1p(w1 * (1.0 - c_pow(1.0 + 4.0 * (y - dy), -0.25)));
  S e1 = e1l + H1;
  // alpha = 0 limit
  S w0 = c_expm1(-e0l / b1c);
  S H0 = b1c * c_log1p(w0 * (1.0 - c_pow(1.0 + 4.0 * chi * s2, -0.25)));
  S e0 = (e0l + H0) * Gc;
  return n * (e1 + f_alpha(al, c1, c2, d, R2_CC) * (e0 - e1));
}

struct Thr { double dens, zeta, sigma, tau; };
inline Thr make_thr(double dens) {
  Thr t; t.dens = dens; t.zeta = 2.220446049250313e-16;
  double s = std::pow(dens, 4.0 / 3.0); t.sigma = s * s; t.tau = 1e-20; return t;
}

// Energy per volume as a function of the 7 libxc inputs (polarised) after clamping.
template <class S>
S energy_pol(int id, const S* v /* ra rb saa sab sbb ta tb */, const Thr& thr, const bool* screened) {
  if (is_x(id)) {
    S e(0.0);
    for (int s = 0; s < 2; s++) {
      if (screened[s]) continue;
      S n = 2.0 * v[s], g2 = 4.0 * v[2 + 2 * s], tau = 2.0 * v[5 + s];
      e += 0.5 * ex_unpol(id, n, g2, tau);
    }
    return e;
  }
  S n = v[0] + v[1];
  S zeta = (v[0] - v[1]) / n;
  S g2 = v[2] + 2.0 * v[3] + v[4];
  S tau = v[5] + v[6];
  return ec_general(id, n, zeta, g2, tau, true, thr.zeta);
}

/// libxc-style polarised evaluation, accumulating into the outputs.
inline void eval_pol(int id, double dens_thr, const double* rho, const double* sigma, const double* tau,
                     double* exc, double* vrho, double* vsigma, double* vtau) {
  Thr thr = make_thr(dens_thr);
  if (id == 0 || rho[0] + rho[1] < thr.dens) return;
  const int fam = family(id);
  double v[7] = {std::max(rho[0], thr.dens), std::max(rho[1], thr.dens), 0, 0, 0, 0, 0};
  if (fam >= 2) { v[2] = std::max(sigma[0], thr.sigma); v[4] = std::max(sigma[2], thr.sigma); }
  if (fam >= 4) {
    v[5] = std::max(tau[0], thr.tau); v[6] = std::max(tau[1], thr.tau);
    v[2] = std::min(v[2], 8.0 * v[0] * v[5]); v[4] = std::min(v[4], 8.0 * v[1] * v[6]);
  }
  if (fam >= 2) { double sa = 0.5 * (v[2] + v[4]); v[3] = std::min(std::max(sigma[1], -sa), sa); }
  bool scr[2] = {v[0] <= thr.dens, v[1] <= thr.dens};
  const double h = 1e-40;
  double e = energy_pol<double>(id, v, thr, scr);
  *exc += e / (v[0] + v[1]);
  const int nv = (fam == 1) ? 2 : (fam == 2 ? 5 : 7);
  double grad[7] = {0, 0, 0, 0, 0, 0, 0};
  for (int k = 0; k < nv; k++) {
    cplx z[7];
    for (int j = 0; j < 7; j++) z[j] = cplx(v[j], j == k ? h : 0.0);
    grad[k] = energy_pol<cplx>(id, z, thr, scr).imag() / h;
  }
  vrho[0] += grad[0]; vrho[1] += grad[1];
  if (fam >= 2) { vsigma[0] += grad[2]; vsigma[1] += grad[3]; vsigma[2] += grad[4]; }
  if (fam >= 4) { vtau[0] += grad[5]; vtau[1] += grad[6]; }
}

template <class S> S energy_unpol(int id, const S& n, const S& g2, const S& tau, const Thr& thr) {
  if (is_x(id)) return ex_unpol(id, n, g2, tau);
  return ec_general(id, n, S(0.0), g2, tau, false, thr.zeta);
}

/// libxc-style unpolarised evaluation, accumulating into the outputs.
inline void eval_unpol(int id, double dens_thr, double rho, double sigma, double tau,
                       double* exc, double* vrho, double* vsigma, double* vtau) {
  Thr thr = make_thr(dens_thr);
  if (id == 0 || rho < thr.dens) return;
  const int fam = family(id);
  rho = std::max(rho, thr.dens);
  if (fam >= 2) sigma = std::max(sigma, thr.sigma); else sigma = 0.0;
  if (fam >= 4) { tau = std::max(tau, thr.tau); sigma = std::min(sigma, 8.0 * rho * tau); } else tau = 0.0;
  if (is_x(id) && 0.5 * rho <= thr.dens) return;
  const double h = 1e-40;
  *exc += energy_unpol<double>(id, rho, sigma, tau, thr) / rho;
  *vrho += energy_unpol<cplx>(id, cplx(rho, h), cplx(sigma, 0), cplx(tau, 0), thr).imag() / h;
  if (fam >= 2) *vsigma += energy_unpol<cplx>(id, cplx(rho, 0), cplx(sigma, h), cplx(tau, 0), thr).imag() / h;
  if (fam >= 4) *vtau += energy_unpol<cplx>(id, cplx(rho, 0), cplx(sigma, 0), cplx(tau, h), thr).imag() / h;
}

}  // namespace oracle_xc
