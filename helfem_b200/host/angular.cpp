#include "angular.h"
#include <atomic>
#include <cmath>
#include <mutex>
#include "fem.h"

namespace hfb {
namespace angular {

void sph_plm_array(int lmax, int m, double x, double* out) {
  // S_mm
  double smm = sqrt(1.0 / (4.0 * M_PI));
  const double sx = sqrt((1.0 - x) * (1.0 + x));
  for (int k = 1; k <= m; k++) smm *= -sqrt((2.0 * k + 1.0) / (2.0 * k)) * sx;
  out[0] = smm;
  if (lmax == m) return;
  double sm1 = x * sqrt(2.0 * m + 3.0) * smm;
  out[1] = sm1;
  double a = smm, b = sm1;
  for (int l = m + 2; l <= lmax; l++) {
    const double c1 = sqrt((4.0 * l * l - 1.0) / ((double)l * l - (double)m * m));
    const double c2 = sqrt((((double)l - 1.0) * (l - 1.0) - (double)m * m) / (4.0 * (l - 1.0) * (l - 1.0) - 1.0));
    const double s = c1 * (x * b - c2 * a);
    out[l - m] = s;
    a = b;
    b = s;
  }
}

double sph_plm(int l, int m, double x) {
  std::vector<double> tmp(l - m + 1);
  sph_plm_array(l, m, x, tmp.data());
  return tmp[l - m];
}

std::complex<double> spherical_harmonics(int l, int m, double cth, double phi) {
  if (m < 0) {
    std::complex<double> y = spherical_harmonics(l, -m, cth, phi);
    return std::conj(((-m) % 2 ? -1.0 : 1.0) * y);
  }
  return std::polar(1.0, m * phi) * sph_plm(l, m, cth);
}

void gauss_legendre(int n, Vec& x, Vec& w) {
  x.assign(n, 0.0);
  w.assign(n, 0.0);
  for (int i = 0; i < (n + 1) / 2; i++) {
    long double z = cosl(M_PIl * (i + 0.75L) / (n + 0.5L));
    long double pp = 1.0L;
    for (int it = 0; it < 100; it++) {
      long double p1 = 1.0L, p2 = 0.0L;
      for (int j = 1; j <= n; j++) {
        long double p3 = p2;
        p2 = p1;
        p1 = ((2.0L * j - 1.0L) * z * p2 - (j - 1.0L) * p3) / j;
      }
      pp = n * (z * p1 - p2) / (z * z - 1.0L);
      long double z1 = z;
      z = z1 - p1 / pp;
      if (fabsl(z - z1) < 1e-19L) break;
    }
    x[i] = (double)(-z);
    x[n - 1 - i] = (double)z;
    w[i] = w[n - 1 - i] = (double)(2.0L / ((1.0L - z * z) * pp * pp));
  }
}

// Gauss-Legendre rule + normalised associated Legendre functions on its nodes, shared by every Gaunt evaluation:
// one rule that is exact for all products up to 3 lmax replaces a rule (and a Newton solve) per coefficient.
namespace {
struct GauntTable {
  int lmax = -1, mmax = -1, n = 0;
  Vec w;                    // 2 pi w_i
  std::vector<double> S;    // [m][l][i], m >= 0, zero for l < m
  const double* row(int l, int mabs) const { return S.data() + ((size_t)mabs * (lmax + 1) + l) * n; }
};
std::atomic<const GauntTable*> g_table{nullptr};
std::mutex g_table_mutex;

const GauntTable* gaunt_table(int lneed, int mneed) {
  const GauntTable* t = g_table.load(std::memory_order_acquire);
  if (t && t->lmax >= lneed && t->mmax >= mneed) return t;
  std::lock_guard<std::mutex> lock(g_table_mutex);
  t = g_table.load(std::memory_order_acquire);
  if (t && t->lmax >= lneed && t->mmax >= mneed) return t;
  GauntTable* nt = new GauntTable();
  nt->lmax = std::max(lneed, t ? t->lmax : 8);
  nt->mmax = std::min(nt->lmax, std::max(mneed, t ? t->mmax : 4));
  nt->n = (3 * nt->lmax) / 2 + 2;
  Vec x;
  gauss_legendre(nt->n, x, nt->w);
  for (double& v : nt->w) v *= 2.0 * M_PI;
  nt->S.assign((size_t)(nt->mmax + 1) * (nt->lmax + 1) * nt->n, 0.0);
  std::vector<double> tmp(nt->lmax + 1);
  for (int m = 0; m <= nt->mmax; m++)
    for (int i = 0; i < nt->n; i++) {
      sph_plm_array(nt->lmax, m, x[i], tmp.data());
      for (int l = m; l <= nt->lmax; l++) nt->S[((size_t)m * (nt->lmax + 1) + l) * nt->n + i] = tmp[l - m];
    }
  g_table.store(nt, std::memory_order_release);  // the superseded table stays alive for concurrent readers
  return nt;
}
}  // namespace

double gaunt_coefficient(int L, int M, int l, int m, int lp, int mp) {
  if (M != m + mp) return 0.0;
  if (L < std::abs(l - lp) || L > l + lp) return 0.0;
  if (std::abs(M) > L || std::abs(m) > l || std::abs(mp) > lp) return 0.0;
  if ((L + l + lp) % 2) return 0.0;  // parity: (L l lp; 0 0 0) vanishes
  // The theta integrand is a polynomial of degree L+l+lp in x (total power of sin is even).
  const int aM = std::abs(M), am = std::abs(m), amp = std::abs(mp);
  const GauntTable* t = gaunt_table(std::max(L, std::max(l, lp)), std::max(aM, std::max(am, amp)));
  const double *s1 = t->row(L, aM), *s2 = t->row(l, am), *s3 = t->row(lp, amp);
  long double s = 0.0L;
  for (int i = 0; i < t->n; i++) s += (long double)(t->w[i] * s1[i]) * (long double)(s2[i] * s3[i]);
  // S_{l,-m} = (-1)^m S_{l,m}
  int sign = 0;
  if (M < 0) sign += aM;
  if (m < 0) sign += am;
  if (mp < 0) sign += amp;
  return (sign % 2) ? (double)(-s) : (double)s;
}

double modified_gaunt_coefficient(int lj, int mj, int L, int M, int li, int mi) {
  static const double const0(2.0 / 3.0 * sqrt(M_PI));
  static const double const2(4.0 / 15.0 * sqrt(5.0 * M_PI));
  double cpl0 = gaunt_coefficient(L, M, 0, 0, L, M) * gaunt_coefficient(lj, mj, li, mi, L, M);
  double cpl2 = 0.0;
  for (int Lp = std::max(std::max(L - 2, 0), std::abs(M)); Lp <= L + 2; Lp++)
    cpl2 += gaunt_coefficient(Lp, M, 2, 0, L, M) * gaunt_coefficient(lj, mj, li, mi, Lp, M);
  return const0 * cpl0 + const2 * cpl2;
}

void angular_chebyshev(int l, int m, Vec& cth, Vec& phi, Vec& w) {
  Vec xl, wl;
  fem::chebyshev(l, xl, wl);
  cth.assign((size_t)l * m, 0.0);
  phi.assign((size_t)l * m, 0.0);
  w.assign((size_t)l * m, 0.0);
  const double dphi = 2.0 * M_PI / m;
  for (int i = 0; i < l; i++)
    for (int j = 0; j < m; j++) {
      const size_t idx = (size_t)i * m + j;
      cth[idx] = xl[i];
      phi[idx] = j * dphi;
      w[idx] = wl[i] * dphi;
    }
}

void angular_basis(int lmax, int mmax, std::vector<int>& lval, std::vector<int>& mval) {
  lval.clear();
  mval.clear();
  for (int mabs = 0; mabs <= mmax; mabs++)
    for (int l = mabs; l <= lmax; l++) {
      lval.push_back(l);
      mval.push_back(mabs);
      if (mabs > 0) {
        lval.push_back(l);
        mval.push_back(-mabs);
      }
    }
}

}  // namespace angular
}  // namespace hfb
