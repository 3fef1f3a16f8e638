// Associated Legendre functions for xi > 1 (see legendre.h).
#include "legendre.h"

#include <algorithm>
#include <cmath>
#include <vector>

namespace hfb {
namespace legendre {

void plm_array(int Lmax, int Mmax, double mu, double* out) {
  const int ld = Lmax + 1;
  std::fill(out, out + (size_t)(Mmax + 1) * ld, 0.0);
  const long double x = coshl((long double)mu), s = sinhl((long double)mu);  // s = sqrt(x^2-1)
  long double pmm = 1.0L;  // P_M^M = (2M-1)!! s^M
  for (int M = 0; M <= Mmax && M <= Lmax; M++) {
    if (M > 0) pmm *= (2 * M - 1) * s;
    long double p0 = pmm, p1 = x * (2 * M + 1) * pmm;
    out[(size_t)M * ld + M] = (double)p0;
    if (M + 1 <= Lmax) out[(size_t)M * ld + M + 1] = (double)p1;
    for (int L = M + 1; L < Lmax; L++) {
      // (L-M+1) P_{L+1}^M = (2L+1) x P_L^M - (L+M) P_{L-1}^M
      const long double p2 = ((2 * L + 1) * x * p1 - (L + M) * p0) / (L - M + 1);
      out[(size_t)M * ld + L + 1] = (double)p2;
      p0 = p1;
      p1 = p2;
    }
  }
}

void qlm_array(int Lmax, int Mmax, double mu, double* out) {
  const int ld = Lmax + 1;
  std::fill(out, out + (size_t)(Mmax + 1) * ld, 0.0);
  const long double x = coshl((long double)mu), s = sinhl((long double)mu);
  // Q_0 = 1/2 ln((x+1)/(x-1)) = ln(coth(mu/2))
  const long double q00 = -logl(tanhl(0.5L * (long double)mu));
  std::vector<long double> q(Lmax + 2, 0.0L);
  // Q_L / Q_{L-1} -> exp(-mu): the forward recursion amplifies errors like exp(2 mu L)
  if (2.0 * mu * Lmax < 3.0) {
    q[0] = q00;
    if (Lmax >= 1) q[1] = x * q00 - 1.0L;
    for (int L = 1; L < Lmax; L++) q[L + 1] = ((2 * L + 1) * x * q[L] - L * q[L - 1]) / (L + 1);
  } else {
    // Miller: backward from Ls where the start-value error has decayed below 1e-20
    const int pad = (int)std::ceil(50.0 / (2.0 * mu)) + 10;
    const int Ls = Lmax + pad;
    long double f2 = 0.0L, f1 = 1e-300L;
    std::vector<long double> t(Lmax + 1, 0.0L);
    for (int L = Ls; L >= 1; L--) {
      // L f_{L-1} = (2L+1) x f_L - (L+1) f_{L+1}
      long double f0 = ((2 * L + 1) * x * f1 - (L + 1) * f2) / L;
      if (L - 1 <= Lmax) t[L - 1] = f0;
      f2 = f1;
      f1 = f0;
      if (fabsl(f1) > 1e3000L) {  // rescale
        f1 *= 1e-3000L;
        f2 *= 1e-3000L;
        for (int k = L - 1; k <= Lmax; k++) t[k] *= 1e-3000L;
      }
    }
    const long double scale = q00 / t[0];
    for (int L = 0; L <= Lmax; L++) q[L] = t[L] * scale;
  }
  // upward in M:  s Q_L^{M+1} = (L-M) x Q_L^M - (L+M) Q_{L-1}^M      (L >= M+1)
  std::vector<long double> cur(q.begin(), q.begin() + Lmax + 1), nxt(Lmax + 1, 0.0L);
  for (int M = 0; M <= Mmax && M <= Lmax; M++) {
    for (int L = M; L <= Lmax; L++) out[(size_t)M * ld + L] = (double)cur[L];
    if (M == Mmax) break;
    for (int L = M + 1; L <= Lmax; L++) nxt[L] = ((L - M) * x * cur[L] - (L + M) * cur[L - 1]) / s;
    cur = nxt;
  }
}

}  // namespace legendre
}  // namespace hfb
