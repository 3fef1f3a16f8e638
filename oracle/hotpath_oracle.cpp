// ORACLE (test infrastructure, not product code): CPU restatement of the reference's Fock-build
// operators for the spherically averaged atom, following the reference's own loop structure.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library; the product (libhelfem_b200.so) never does.
//
// Functions restated (read-only tree /root/reference):
//   oracle_sadatom_coulomb    src/sadatom/basis.cpp:296-358   TwoDBasis::coulomb
//   oracle_sadatom_exchange   src/sadatom/basis.cpp:360-526   TwoDBasis::exchange
//   oracle_sadatom_eval_fxc   src/sadatom/dftgrid.cpp:874-993 DFTGrid::eval_Fxc (R and U), with
//       compute_bf :842-863, update_density :46-227, compute_xc :306-486 (libxc calls replaced by
//       oracle/xc_oracle.hpp), eval_Exc :566-572, eval_Fxc :610-805, increment_lda/gga dftgrid.h:190-239
//   oracle_gaunt              src/general/gaunt.cpp:35-53 (3j symbols by the Racah formula)
// The tables (basis values, weights, stored integrals) are INPUTS, exactly what the C ABI of the
// product receives; tests feed the oracle tables from oracle/fem_oracle.py and the product tables
// from its own host library, so the two chains are independent end to end.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "xc_oracle.hpp"

namespace {

typedef std::vector<double> vec;

// ---- 3j symbols / Gaunt (Racah formula, long double; exact to rounding for l <~ 20) ----
long double lfact(int n) {
  long double f = 1.0L;
  for (int i = 2; i <= n; i++) f *= i;
  return f;
}
long double wigner3j(int j1, int j2, int j3, int m1, int m2, int m3) {
  if (m1 + m2 + m3 != 0) return 0.0L;
  if (j3 < std::abs(j1 - j2) || j3 > j1 + j2) return 0.0L;
  if (std::abs(m1) > j1 || std::abs(m2) > j2 || std::abs(m3) > j3) return 0.0L;
  const int t1 = j2 - m1 - j3, t2 = j1 + m2 - j3, t3 = j1 + j2 - j3, t4 = j1 - m1, t5 = j2 + m2;
  const int tmin = std::max(0, std::max(t1, t2)), tmax = std::min(t3, std::min(t4, t5));
  long double sum = 0.0L;
  for (int t = tmin; t <= tmax; t++) {
    long double den = lfact(t) * lfact(t - t1) * lfact(t - t2) * lfact(t3 - t) * lfact(t4 - t) * lfact(t5 - t);
    sum += ((t % 2) ? -1.0L : 1.0L) / den;
  }
  long double tri = lfact(j1 + j2 - j3) * lfact(j1 - j2 + j3) * lfact(-j1 + j2 + j3) / lfact(j1 + j2 + j3 + 1);
  long double pre = sqrtl(tri * lfact(j1 + m1) * lfact(j1 - m1) * lfact(j2 + m2) * lfact(j2 - m2) * lfact(j3 + m3) * lfact(j3 - m3));
  const int ph = j1 - j2 - m3;
  return ((std::abs(ph) % 2) ? -1.0L : 1.0L) * pre * sum;
}
double gaunt(int L, int M, int l, int m, int lp, int mp) {
  if (M != m + mp) return 0.0;
  if (L < std::abs(l - lp) || L > l + lp) return 0.0;
  long double res = sqrtl((2 * L + 1) * (2 * l + 1) * (2 * lp + 1) / (4.0L * M_PIl));
  res *= wigner3j(L, l, lp, 0, 0, 0);
  res *= wigner3j(L, l, lp, -M, m, mp);
  if (std::abs(M) % 2) res = -res;
  return (double)res;
}

struct Tables {
  int nel, nq;
  const int *first, *last;
  int np(int e) const { return last[e] - first[e] + 1; }
};

}  // namespace

extern "C" {

double oracle_gaunt(int L, int M, int l, int m, int lp, int mp) { return gaunt(L, M, l, m, lp, mp); }
int oracle_xc_family(int func) { return oracle_xc::family(func); }

void oracle_xc_unpol(int func, int n, double thr, const double* rho, const double* sigma, const double* tau,
                     double* exc, double* vrho, double* vsigma, double* vtau) {
  for (int i = 0; i < n; i++) {
    double e = 0, vr = 0, vs = 0, vt = 0;
    oracle_xc::eval_unpol(func, thr, rho[i], sigma ? sigma[i] : 0.0, tau ? tau[i] : 0.0, &e, &vr, &vs, &vt);
    exc[i] += e; vrho[i] += vr;
    if (vsigma) vsigma[i] += vs;
    if (vtau) vtau[i] += vt;
  }
}
void oracle_xc_pol(int func, int n, double thr, const double* rho, const double* sigma, const double* tau,
                   double* exc, double* vrho, double* vsigma, double* vtau) {
  for (int i = 0; i < n; i++) {
    double e = 0, vr[2] = {0, 0}, vs[3] = {0, 0, 0}, vt[2] = {0, 0};
    double s3[3] = {0, 0, 0}, t2[2] = {0, 0};
    if (sigma) { s3[0] = sigma[3 * i]; s3[1] = sigma[3 * i + 1]; s3[2] = sigma[3 * i + 2]; }
    if (tau) { t2[0] = tau[2 * i]; t2[1] = tau[2 * i + 1]; }
    oracle_xc::eval_pol(func, thr, rho + 2 * i, s3, t2, &e, vr, vs, vt);
    exc[i] += e; vrho[2 * i] += vr[0]; vrho[2 * i + 1] += vr[1];
    if (vsigma) { vsigma[3 * i] += vs[0]; vsigma[3 * i + 1] += vs[1]; vsigma[3 * i + 2] += vs[2]; }
    if (vtau) { vtau[2 * i] += vt[0]; vtau[2 * i + 1] += vt[1]; }
  }
}

// J = coulomb(P) -- sadatom/basis.cpp:296-358 (L = 0 only, :306).  dL/dm1L/tei: arrays of nel pointers.
int oracle_sadatom_coulomb(int nel, const int* first, const int* last, const double* const* dL,
                           const double* const* dm1L, const double* const* tei, const double* P, int nrad, double* J) {
  std::fill(J, J + (size_t)nrad * nrad, 0.0);
  const double Lfac = 4.0 * M_PI;
  for (int jel = 0; jel < nel; jel++) {
    const int jf = first[jel], Nj = last[jel] - jf + 1;
    vec Psub((size_t)Nj * Nj);
    for (int b = 0; b < Nj; b++)
      for (int a = 0; a < Nj; a++) Psub[a + b * Nj] = P[(jf + a) + (size_t)(jf + b) * nrad];
    double ts = 0.0, tb = 0.0;  // trace(D*P) = sum_ab D(a,b) P(b,a)
    for (int a = 0; a < Nj; a++)
      for (int b = 0; b < Nj; b++) {
        ts += dL[jel][a + b * Nj] * Psub[b + a * Nj];
        tb += dm1L[jel][a + b * Nj] * Psub[b + a * Nj];
      }
    const double jsmall = Lfac * ts, jbig = Lfac * tb;
    for (int iel = 0; iel < jel; iel++) {
      const int f = first[iel], Ni = last[iel] - f + 1;
      for (int b = 0; b < Ni; b++)
        for (int a = 0; a < Ni; a++) J[(f + a) + (size_t)(f + b) * nrad] += jbig * dL[iel][a + b * Ni];
    }
    for (int iel = jel + 1; iel < nel; iel++) {
      const int f = first[iel], Ni = last[iel] - f + 1;
      for (int b = 0; b < Ni; b++)
        for (int a = 0; a < Ni; a++) J[(f + a) + (size_t)(f + b) * nrad] += jsmall * dm1L[iel][a + b * Ni];
    }
    const size_t n2 = (size_t)Nj * Nj;
    vec Jsub(n2, 0.0);
    for (size_t c = 0; c < n2; c++) {
      const double p = Psub[c];
      const double* col = tei[jel] + c * n2;
      for (size_t r = 0; r < n2; r++) Jsub[r] += col[r] * p;
    }
    for (int b = 0; b < Nj; b++)
      for (int a = 0; a < Nj; a++) J[(jf + a) + (size_t)(jf + b) * nrad] += Lfac * Jsub[a + b * Nj];
  }
  return 0;
}

// -K = exchange(P) -- sadatom/basis.cpp:360-526.  dL/dm1L/ktei: arrays of (2 lmax+1)*nel pointers, [L*nel+iel].
int oracle_sadatom_exchange(int lmax, int nel, const int* first, const int* last, const double* const* dL,
                            const double* const* dm1L, const double* const* ktei, const double* P, int nrad, double* K) {
  const int nsl = lmax + 1, nL = 2 * lmax + 1;
  const size_t n2 = (size_t)nrad * nrad;
  std::fill(K, K + n2 * nsl, 0.0);
#ifdef _OPENMP
#pragma omp parallel for
#endif
  for (int lout = 0; lout <= lmax; lout++) {
    std::vector<vec> Prad(nL, vec(n2, 0.0));
    std::vector<bool> coupling(nL, false);
    for (int lin = 0; lin <= lmax; lin++) {
      double nrm = 0.0;
      for (size_t k = 0; k < n2; k++) nrm += P[lin * n2 + k] * P[lin * n2 + k];
      if (nrm == 0.0) continue;
      const int Lmin = std::abs(lin - lout), Lmax = lin + lout;
      vec totcoup(Lmax + 1, 0.0);
      for (int mout = -lout; mout <= lout; mout++)
        for (int min = -lin; min <= lin; min++) {
          const int M = mout - min;
          for (int L = Lmin; L <= Lmax; L++) {
            const double g = gaunt(lout, mout, L, M, lin, min);
            totcoup[L] += g * g;
          }
        }
      for (int L = Lmin; L <= Lmax; L++) {
        totcoup[L] /= 2 * lout + 1;
        if (totcoup[L] == 0.0) continue;
        const double Lfac = 4.0 * M_PI / (2 * L + 1);
        for (size_t k = 0; k < n2; k++) Prad[L][k] += (Lfac * totcoup[L]) * P[lin * n2 + k];
        coupling[L] = true;
      }
    }
    double* Kl = K + (size_t)lout * n2;
    for (int L = 0; L < nL; L++) {
      if (!coupling[L]) continue;
      const vec& PL = Prad[L];
      for (int iel = 0; iel < nel; iel++)
        for (int jel = 0; jel < nel; jel++) {
          const int fi = first[iel], fj = first[jel], Ni = last[iel] - fi + 1, Nj = last[jel] - fj + 1;
          if (iel == jel) {
            const size_t m2 = (size_t)Ni * Nj;
            vec pv(m2), kv(m2, 0.0);
            for (int b = 0; b < Nj; b++)
              for (int a = 0; a < Ni; a++) pv[a + b * Ni] = PL[(fi + a) + (size_t)(fj + b) * nrad];
            const double* T = ktei[L * nel + iel];
            for (size_t c = 0; c < m2; c++)
              for (size_t r = 0; r < m2; r++) kv[r] += T[r + c * m2] * pv[c];
            for (int b = 0; b < Nj; b++)
              for (int a = 0; a < Ni; a++) Kl[(fi + a) + (size_t)(fj + b) * nrad] -= kv[a + b * Ni];
          } else {
            const double* iint = (iel > jel) ? dm1L[L * nel + iel] : dL[L * nel + iel];
            const double* jint = (iel > jel) ? dL[L * nel + jel] : dm1L[L * nel + jel];
            vec T((size_t)Ni * Nj, 0.0);
            for (int k = 0; k < Nj; k++)
              for (int b = 0; b < Nj; b++) {
                const double jv = jint[k + b * Nj];  // jint^T(b,k)
                for (int a = 0; a < Ni; a++) T[a + k * Ni] += PL[(fi + a) + (size_t)(fj + b) * nrad] * jv;
              }
            for (int k = 0; k < Nj; k++)
              for (int c = 0; c < Ni; c++) {
                const double tv = T[c + k * Ni];
                for (int a = 0; a < Ni; a++) Kl[(fi + a) + (size_t)(fj + k) * nrad] -= iint[a + c * Ni] * tv;
              }
          }
        }
    }
  }
  return 0;
}

// eval_Fxc -- sadatom/dftgrid.cpp:874-993.  bf/df: per element (nq x np_e) column-major, concatenated
// (what TwoDBasis::eval_bf/eval_df return).  P*, H*: nrad x nrad x nslices cubes.
int oracle_sadatom_eval_fxc(int nel, int nq, const int* first, const int* last, const double* r, const double* wrad,
                            const double* bf, const double* df, int x_func, int c_func, int nspin, const double* Pa,
                            const double* Pb, int nrad, int nslices, double thr, double* Ha, double* Hb, int beta,
                            double* Exc, double* Nel) {
  const size_t n2 = (size_t)nrad * nrad;
  std::fill(Ha, Ha + n2 * nslices, 0.0);
  if (nspin == 2 && Hb) std::fill(Hb, Hb + n2 * nslices, 0.0);
  const int famx = oracle_xc::family(x_func), famc = oracle_xc::family(c_func);
  if ((x_func && !famx) || (c_func && !famc)) return -3;
  const int fam = std::max(famx, famc);
  const bool do_grad = fam >= 2, do_tau = fam >= 4;
  std::vector<size_t> off(nel + 1, 0);
  for (int e = 0; e < nel; e++) off[e + 1] = off[e] + (size_t)(last[e] - first[e] + 1) * nq;
  double exc = 0.0, nelsum = 0.0;
  // even elements first, then odd ones, as the reference does to avoid write races (:883-924)
  for (int parity = 0; parity < 2; parity++) {
#ifdef _OPENMP
#pragma omp parallel for reduction(+ : exc, nelsum) schedule(dynamic)
#endif
    for (int iel = parity; iel < nel; iel += 2) {
      const int f0 = first[iel], np = last[iel] - f0 + 1;
      const double* B = bf + off[iel];   // B(p,i) = B[p + i*nq]
      const double* dB = df + off[iel];
      const double* rr = r + (size_t)iel * nq;
      const double* wr = wrad + (size_t)iel * nq;
      vec wtot(nq);
      for (int p = 0; p < nq; p++) wtot[p] = 4.0 * M_PI * wr[p] * rr[p] * rr[p];
      // update_density
      vec rho(nspin * nq, 0.0), grho(nspin * nq, 0.0), sigma((nspin == 1 ? 1 : 3) * nq, 0.0), tau(nspin * nq, 0.0);
      for (int s = 0; s < nspin; s++) {
        const double* Pc = s ? Pb : Pa;
        vec Pt((size_t)np * np, 0.0), Pl((size_t)np * np, 0.0);
        for (int l = 0; l < nslices; l++)
          for (int b = 0; b < np; b++)
            for (int a = 0; a < np; a++) {
              const double v = Pc[l * n2 + (f0 + a) + (size_t)(f0 + b) * nrad];
              Pt[a + b * np] += v;
              Pl[a + b * np] += l * (l + 1) * v;
            }
        vec Pv((size_t)np * nq, 0.0), Pvp, Plv;
        if (do_tau) { Pvp.assign((size_t)np * nq, 0.0); Plv.assign((size_t)np * nq, 0.0); }
        for (int p = 0; p < nq; p++)
          for (int b = 0; b < np; b++) {
            const double bv = B[p + b * nq];
            const double dv = do_grad ? dB[p + b * nq] : 0.0;
            for (int a = 0; a < np; a++) {
              Pv[a + p * np] += Pt[a + b * np] * bv;
              if (do_tau) {
                Pvp[a + p * np] += Pt[a + b * np] * dv;
                Plv[a + p * np] += Pl[a + b * np] * bv;
              }
            }
          }
        for (int p = 0; p < nq; p++) {
          double d = 0.0, g = 0.0, t1 = 0.0, t2 = 0.0;
          for (int a = 0; a < np; a++) {
            d += Pv[a + p * np] * B[p + a * nq];
            if (do_grad) g += Pv[a + p * np] * dB[p + a * nq];
            if (do_tau) { t1 += Pvp[a + p * np] * dB[p + a * nq]; t2 += Plv[a + p * np] * B[p + a * nq]; }
          }
          rho[nspin * p + s] = d;
          grho[nspin * p + s] = 2.0 * g;
          if (do_tau) tau[nspin * p + s] = 0.5 * (t1 + std::max(t2 / (rr[p] * rr[p]), 0.0));
        }
      }
      if (do_grad)
        for (int p = 0; p < nq; p++) {
          if (nspin == 1) sigma[p] = grho[p] * grho[p];
          else {
            sigma[3 * p] = grho[2 * p] * grho[2 * p];
            sigma[3 * p + 1] = grho[2 * p] * grho[2 * p + 1];
            sigma[3 * p + 2] = grho[2 * p + 1] * grho[2 * p + 1];
          }
        }
      for (int p = 0; p < nq; p++) nelsum += wtot[p] * (nspin == 1 ? rho[p] : rho[2 * p] + rho[2 * p + 1]);
      // compute_xc for x and c, accumulating
      vec e(nq, 0.0), vxc(nspin * nq, 0.0), vsigma((nspin == 1 ? 1 : 3) * nq, 0.0), vtau(nspin * nq, 0.0);
      for (int which = 0; which < 2; which++) {
        const int id = which ? c_func : x_func;
        if (id <= 0) continue;
        if (nspin == 1) oracle_xc_unpol(id, nq, thr, rho.data(), sigma.data(), tau.data(), e.data(), vxc.data(), vsigma.data(), vtau.data());
        else oracle_xc_pol(id, nq, thr, rho.data(), sigma.data(), tau.data(), e.data(), vxc.data(), vsigma.data(), vtau.data());
      }
      for (int p = 0; p < nq; p++) exc += wtot[p] * e[p] * (nspin == 1 ? rho[p] : rho[2 * p] + rho[2 * p + 1]);
      // eval_Fxc
      for (int s = 0; s < nspin; s++) {
        if (s == 1 && !beta) continue;
        vec H((size_t)np * np, 0.0), Hl((size_t)np * np, 0.0);
        vec a(nq), g(nq, 0.0), c(nq, 0.0), dl(nq, 0.0);
        for (int p = 0; p < nq; p++) {
          a[p] = vxc[nspin * p + s] * wtot[p];
          if (do_grad) {
            if (nspin == 1) g[p] = 2.0 * (wtot[p] * vsigma[p]) * grho[p];
            else {
              const int o = 1 - s;
              g[p] = wtot[p] * (2.0 * vsigma[3 * p + 2 * s] * grho[2 * p + s] + vsigma[3 * p + 1] * grho[2 * p + o]);
            }
          }
          if (do_tau) {
            c[p] = 0.5 * vtau[nspin * p + s] * wtot[p];
            dl[p] = vtau[nspin * p + s] * (0.5 * wr[p] * 4.0 * M_PI);
          }
        }
        for (int j = 0; j < np; j++)
          for (int i = 0; i < np; i++) {
            double h = 0.0, hl = 0.0;
            for (int p = 0; p < nq; p++) {
              const double bi = B[p + i * nq], bj = B[p + j * nq];
              h += a[p] * bi * bj;
              if (do_grad) {
                const double di = dB[p + i * nq], dj = dB[p + j * nq];
                h += g[p] * (di * bj + bi * dj);
                if (do_tau) { h += c[p] * di * dj; hl += dl[p] * bi * bj; }
              }
            }
            H[i + j * np] = h;
            Hl[i + j * np] = hl;
          }
        double* Ho = s ? Hb : Ha;
        for (int l = 0; l < nslices; l++)
          for (int j = 0; j < np; j++)
            for (int i = 0; i < np; i++) Ho[l * n2 + (f0 + i) + (size_t)(f0 + j) * nrad] += H[i + j * np] + l * (l + 1) * Hl[i + j * np];
      }
    }
  }
  *Exc = exc;
  *Nel = nelsum;
  return 0;
}

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/// bench.py under torchrun inherits OMP_NUM_THREADS=1; the CPU arm asks for all host threads explicitly
void oracle_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

}  // extern "C"
