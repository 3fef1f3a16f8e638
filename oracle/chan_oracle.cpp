// ORACLE (test infrastructure, not product code): CPU restatement of the reference's Gaunt-coupled
// Coulomb / exchange operators of the `atomic` program, following the reference's own loop structure
// with dense Nrad x Nrad helper matrices.
//   oracle_atomic_coulomb    /root/reference/src/atomic/TwoDBasis.cpp:817-955
//   oracle_atomic_exchange   /root/reference/src/atomic/TwoDBasis.cpp:957-1140
// Stored integrals are INPUTS (arrays of pointers, entry [L*nel + iel]); Gaunt coefficients come from
// the oracle's own Racah-formula implementation (hotpath_oracle.cpp), not from the product.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

extern "C" double oracle_gaunt(int L, int M, int l, int m, int lp, int mp);

namespace {
typedef std::vector<double> vec;
inline double gcoeff(int L, int M, int l, int m, int lp, int mp) {
  if (std::abs(M) > L || std::abs(m) > l || std::abs(mp) > lp) return 0.0;  // Gaunt::coeff guards, gaunt.cpp:113-115
  return oracle_gaunt(L, M, l, m, lp, mp);
}
}  // namespace

extern "C" {

int oracle_atomic_coulomb(int nang, const int* lval, const int* mval, int nel, const int* first, const int* last,
                          const double* const* dL, const double* const* dm1L, const double* const* tei, const double* P,
                          double* J) {
  const int nrad = last[nel - 1] + 1, ntot = nang * nrad;
  const size_t n2 = (size_t)nrad * nrad;
  int lmax = 0, mmin = 0, mmx = 0;
  for (int a = 0; a < nang; a++) { lmax = std::max(lmax, lval[a]); mmin = std::min(mmin, mval[a]); mmx = std::max(mmx, mval[a]); }
  const int Mmax = mmx - mmin, nL = 2 * lmax + 1, nM = 2 * Mmax + 1;
  std::vector<vec> Paux((size_t)nL * nM), Jaux((size_t)nL * nM);
  for (int L = 0; L < nL; L++)
    for (int M = -std::min(L, Mmax); M <= std::min(L, Mmax); M++) {
      Paux[L * nM + M + Mmax].assign(n2, 0.0);
      Jaux[L * nM + M + Mmax].assign(n2, 0.0);
    }
  // contract ket (:845-864)
  for (int k = 0; k < nang; k++)
    for (int l = 0; l < nang; l++) {
      const int lk = lval[k], mk = mval[k], ll = lval[l], ml = mval[l], M = mk - ml;
      for (int L = std::max(std::abs(lk - ll), std::abs(M)); L <= lk + ll; L++) {
        const double cpl = gcoeff(lk, mk, L, M, ll, ml);
        if (cpl == 0.0) continue;
        vec& Pa = Paux[L * nM + M + Mmax];
        for (int b = 0; b < nrad; b++)
          for (int a = 0; a < nrad; a++) Pa[a + (size_t)b * nrad] += cpl * P[(size_t)(k * nrad + a) + (size_t)(l * nrad + b) * ntot];
      }
    }
  // radial contraction (:872-927)
  for (int L = 0; L < nL; L++) {
    const double Lfac = 4.0 * M_PI / (2 * L + 1);
    for (int M = -std::min(L, Mmax); M <= std::min(L, Mmax); M++) {
      const vec& Pa = Paux[L * nM + M + Mmax];
      vec& Ja = Jaux[L * nM + M + Mmax];
      for (int jel = 0; jel < nel; jel++) {
        const int jf = first[jel], Nj = last[jel] - jf + 1;
        vec Psub((size_t)Nj * Nj);
        for (int b = 0; b < Nj; b++)
          for (int a = 0; a < Nj; a++) Psub[a + b * Nj] = Pa[(jf + a) + (size_t)(jf + b) * nrad];
        double ts = 0.0, tb = 0.0;
        for (int a = 0; a < Nj; a++)
          for (int b = 0; b < Nj; b++) {
            ts += dL[L * nel + jel][a + b * Nj] * Psub[b + a * Nj];
            tb += dm1L[L * nel + jel][a + b * Nj] * Psub[b + a * Nj];
          }
        const double jsmall = Lfac * ts, jbig = Lfac * tb;
        for (int iel = 0; iel < jel; iel++) {
          const int f = first[iel], Ni = last[iel] - f + 1;
          for (int b = 0; b < Ni; b++)
            for (int a = 0; a < Ni; a++) Ja[(f + a) + (size_t)(f + b) * nrad] += jbig * dL[L * nel + iel][a + b * Ni];
        }
        for (int iel = jel + 1; iel < nel; iel++) {
          const int f = first[iel], Ni = last[iel] - f + 1;
          for (int b = 0; b < Ni; b++)
            for (int a = 0; a < Ni; a++) Ja[(f + a) + (size_t)(f + b) * nrad] += jsmall * dm1L[L * nel + iel][a + b * Ni];
        }
        const size_t m2 = (size_t)Nj * Nj;
        vec Jsub(m2, 0.0);
        const double* T = tei[L * nel + jel];
        for (size_t c = 0; c < m2; c++)
          for (size_t r = 0; r < m2; r++) Jsub[r] += T[r + c * m2] * Psub[c];
        for (int b = 0; b < Nj; b++)
          for (int a = 0; a < Nj; a++) Ja[(jf + a) + (size_t)(jf + b) * nrad] += Lfac * Jsub[a + b * Nj];
      }
    }
  }
  // bra (:932-952)
  std::fill(J, J + (size_t)ntot * ntot, 0.0);
  for (int i = 0; i < nang; i++)
    for (int j = 0; j < nang; j++) {
      const int li = lval[i], mi = mval[i], lj = lval[j], mj = mval[j], M = mj - mi;
      for (int L = std::max(std::abs(lj - li), std::abs(M)); L <= lj + li; L++) {
        const double cpl = gcoeff(lj, mj, L, M, li, mi);
        if (cpl == 0.0) continue;
        const vec& Ja = Jaux[L * nM + M + Mmax];
        for (int b = 0; b < nrad; b++)
          for (int a = 0; a < nrad; a++) J[(size_t)(i * nrad + a) + (size_t)(j * nrad + b) * ntot] += cpl * Ja[a + (size_t)b * nrad];
      }
    }
  return 0;
}

int oracle_atomic_exchange(int nang, const int* lval, const int* mval, int nel, const int* first, const int* last,
                           const double* const* dL, const double* const* dm1L, const double* const* ktei, const double* P,
                           double* K) {
  const int nrad = last[nel - 1] + 1, ntot = nang * nrad;
  const size_t n2 = (size_t)nrad * nrad;
  int lmax = 0;
  for (int a = 0; a < nang; a++) lmax = std::max(lmax, lval[a]);
  const int nL = 2 * lmax + 1;
  std::fill(K, K + (size_t)ntot * ntot, 0.0);
#ifdef _OPENMP
#pragma omp parallel for collapse(2) schedule(dynamic)
#endif
  for (int jang = 0; jang < nang; jang++)
    for (int kang = 0; kang < nang; kang++) {
      const int lj = lval[jang], mj = mval[jang], lk = lval[kang], mk = mval[kang];
      std::vector<vec> Rmat(nL, vec(n2, 0.0));
      std::vector<bool> couple(nL, false);
      for (int iang = 0; iang < nang; iang++)
        for (int lang = 0; lang < nang; lang++) {
          const int li = lval[iang], mi = mval[iang], ll = lval[lang], ml = mval[lang];
          const int M = mj - mi, Mp = mk - ml;
          if (M != Mp) continue;
          double nrm = 0.0;
          for (int b = 0; b < nrad; b++)
            for (int a = 0; a < nrad; a++) {
              const double v = P[(size_t)(iang * nrad + a) + (size_t)(lang * nrad + b) * ntot];
              nrm += v * v;
            }
          if (std::sqrt(nrm) < 10 * 2.220446049250313e-16) continue;
          const int Lmin = std::max(std::max(std::abs(li - lj), std::abs(lk - ll)), std::abs(M));
          const int Lmax = std::min(li + lj, lk + ll);
          for (int L = Lmin; L <= Lmax; L++) {
            const double cpl = gcoeff(lj, mj, L, M, li, mi) * gcoeff(lk, mk, L, M, ll, ml);
            if (cpl == 0.0) continue;
            const double Lfac = 4.0 * M_PI / (2 * L + 1);
            for (int b = 0; b < nrad; b++)
              for (int a = 0; a < nrad; a++)
                Rmat[L][a + (size_t)b * nrad] += (Lfac * cpl) * P[(size_t)(iang * nrad + a) + (size_t)(lang * nrad + b) * ntot];
            couple[L] = true;
          }
        }
      for (int iel = 0; iel < nel; iel++)
        for (int jel = 0; jel < nel; jel++) {
          const int fi = first[iel], fj = first[jel], Ni = last[iel] - fi + 1, Nj = last[jel] - fj + 1;
          vec Ksub((size_t)Ni * Nj, 0.0);
          for (int L = 0; L < nL; L++) {
            if (!couple[L]) continue;
            const vec& R = Rmat[L];
            if (iel == jel) {
              const size_t m2 = (size_t)Ni * Nj;
              const double* T = ktei[L * nel + iel];
              for (int b = 0; b < Nj; b++)
                for (int a = 0; a < Ni; a++) {
                  const double pv = R[(fi + a) + (size_t)(fj + b) * nrad];
                  const size_t c = a + (size_t)b * Ni;
                  for (size_t r = 0; r < m2; r++) Ksub[r] += T[r + c * m2] * pv;
                }
            } else {
              const double* iint = (iel > jel) ? dm1L[L * nel + iel] : dL[L * nel + iel];
              const double* jint = (iel > jel) ? dL[L * nel + jel] : dm1L[L * nel + jel];
              vec T((size_t)Ni * Nj, 0.0);
              for (int k = 0; k < Nj; k++)
                for (int b = 0; b < Nj; b++) {
                  const double jv = jint[k + b * Nj];
                  for (int a = 0; a < Ni; a++) T[a + k * Ni] += R[(fi + a) + (size_t)(fj + b) * nrad] * jv;
                }
              for (int k = 0; k < Nj; k++)
                for (int c = 0; c < Ni; c++) {
                  const double tv = T[c + k * Ni];
                  for (int a = 0; a < Ni; a++) Ksub[a + k * Ni] += iint[a + c * Ni] * tv;
                }
            }
          }
          for (int b = 0; b < Nj; b++)
            for (int a = 0; a < Ni; a++)
              K[(size_t)(jang * nrad + fi + a) + (size_t)(kang * nrad + fj + b) * ntot] -= Ksub[a + b * Ni];
        }
    }
  return 0;
}

// ---------------------------------------------------------------------------------------------
// diatomic: /root/reference/src/diatomic/basis.cpp:1359-1530 (coulomb), :1532-1733 (exchange).
// Matrices are in the DUMMY space (nang*nrad); expand/remove_boundaries is done by the caller.
// lm families sorted (L,|M|); disjoint arrays [ilm*nel+iel]; tei/ktei [4*(ilm*nel+iel) + 2a+b].
// ---------------------------------------------------------------------------------------------
static double mod_coeff(int lj, int mj, int L, int M, int li, int mi) {  // general/gaunt.cpp:167-180
  static const double const0(2.0 / 3.0 * sqrt(M_PI));
  static const double const2(4.0 / 15.0 * sqrt(5.0 * M_PI));
  const double cpl0 = gcoeff(L, M, 0, 0, L, M) * gcoeff(lj, mj, li, mi, L, M);
  double cpl2 = 0.0;
  for (int Lp = std::max(std::max(L - 2, 0), std::abs(M)); Lp <= L + 2; Lp++) cpl2 += gcoeff(Lp, M, 2, 0, L, M) * gcoeff(lj, mj, li, mi, Lp, M);
  return const0 * cpl0 + const2 * cpl2;
}
static int lmind(int nlm, const int* Lmap, const int* Mmap, int L, int M) {
  M = std::abs(M);
  for (int r = 0; r < nlm; r++)
    if (Lmap[r] == L && Mmap[r] == M) return r;
  return -1;
}
static double LMfac(double Rhalf, int L, int M) {  // basis.cpp:1421
  M = std::abs(M);
  double ratio = 1.0;
  for (int p = L + M; p > L - M; p--) ratio *= p;
  return 4.0 * M_PI * std::pow(Rhalf, 5) * std::pow(-1.0, M) / ratio;
}

int oracle_diatomic_coulomb(int nang, const int* lval, const int* mval, int nel, const int* first, const int* last, int nlm,
                            const int* Lmap, const int* Mmap, double Rhalf, const double* const* P0, const double* const* P2,
                            const double* const* Q0, const double* const* Q2, const double* const* tei, const double* P, double* J) {
  const int nrad = last[nel - 1] + 1, ntot = nang * nrad;
  const size_t n2 = (size_t)nrad * nrad;
  // LM channels keyed (L, M)
  std::vector<std::pair<int, int>> LM;
  auto LMind = [&](int L, int M) {
    for (size_t c = 0; c < LM.size(); c++)
      if (LM[c].first == L && LM[c].second == M) return (int)c;
    LM.push_back({L, M});
    return (int)LM.size() - 1;
  };
  std::vector<vec> Paux0, Paux2;
  for (int k = 0; k < nang; k++)
    for (int l = 0; l < nang; l++) {
      const int lk = lval[k], mk = mval[k], ll = lval[l], ml = mval[l], M = mk - ml;
      for (int L = std::max(std::abs(lk - ll) - 2, std::abs(M)); L <= lk + ll + 2; L++) {
        const int c = LMind(L, M);
        if ((int)Paux0.size() <= c) { Paux0.resize(c + 1, vec(n2, 0.0)); Paux2.resize(c + 1, vec(n2, 0.0)); }
        const double c0 = mod_coeff(lk, mk, L, M, ll, ml), c2 = gcoeff(lk, mk, L, M, ll, ml);
        for (int b = 0; b < nrad; b++)
          for (int a = 0; a < nrad; a++) {
            const double p = P[(size_t)(k * nrad + a) + (size_t)(l * nrad + b) * ntot];
            if (c0 != 0.0) Paux0[c][a + (size_t)b * nrad] += c0 * p;
            if (c2 != 0.0) Paux2[c][a + (size_t)b * nrad] += c2 * p;
          }
      }
    }
  std::vector<vec> Jaux0(LM.size(), vec(n2, 0.0)), Jaux2(LM.size(), vec(n2, 0.0));
  for (size_t c = 0; c < LM.size(); c++) {
    const int L = LM[c].first, M = LM[c].second, ilm = lmind(nlm, Lmap, Mmap, L, M);
    if (ilm < 0) return -1;
    const double fac = LMfac(Rhalf, L, M);
    for (int jel = 0; jel < nel; jel++) {
      const int jf = first[jel], Nj = last[jel] - jf + 1;
      vec Ps0((size_t)Nj * Nj), Ps2((size_t)Nj * Nj);
      for (int b = 0; b < Nj; b++)
        for (int a = 0; a < Nj; a++) {
          Ps0[a + b * Nj] = Paux0[c][(jf + a) + (size_t)(jf + b) * nrad];
          Ps2[a + b * Nj] = Paux2[c][(jf + a) + (size_t)(jf + b) * nrad];
        }
      auto trace = [&](const double* D, const vec& Ps) {
        double t = 0.0;
        for (int a = 0; a < Nj; a++)
          for (int b = 0; b < Nj; b++) t += D[a + b * Nj] * Ps[b + a * Nj];
        return t;
      };
      const double jsmall0 = fac * trace(P0[ilm * nel + jel], Ps0), jbig0 = fac * trace(Q0[ilm * nel + jel], Ps0);
      const double jsmall2 = fac * trace(P2[ilm * nel + jel], Ps2), jbig2 = fac * trace(Q2[ilm * nel + jel], Ps2);
      for (int iel = 0; iel < nel; iel++) {
        if (iel == jel) continue;
        const int f = first[iel], Ni = last[iel] - f + 1;
        const double ifac0 = iel < jel ? jbig0 - jbig2 : jsmall0 - jsmall2;
        const double* i0 = iel < jel ? P0[ilm * nel + iel] : Q0[ilm * nel + iel];
        const double* i2 = iel < jel ? P2[ilm * nel + iel] : Q2[ilm * nel + iel];
        for (int b = 0; b < Ni; b++)
          for (int a = 0; a < Ni; a++) {
            Jaux0[c][(f + a) + (size_t)(f + b) * nrad] += i0[a + b * Ni] * ifac0;
            Jaux2[c][(f + a) + (size_t)(f + b) * nrad] -= i2[a + b * Ni] * ifac0;
          }
      }
      const size_t m2 = (size_t)Nj * Nj;
      const double* const* T = tei + 4 * (ilm * nel + jel);
      for (size_t cc = 0; cc < m2; cc++)
        for (size_t r = 0; r < m2; r++) {
          const double j0 = fac * (T[0][r + cc * m2] * Ps0[cc] - T[1][r + cc * m2] * Ps2[cc]);
          const double j2 = fac * (-T[2][r + cc * m2] * Ps0[cc] + T[3][r + cc * m2] * Ps2[cc]);
          const int a = (int)(r % Nj), b = (int)(r / Nj);
          Jaux0[c][(jf + a) + (size_t)(jf + b) * nrad] += j0;
          Jaux2[c][(jf + a) + (size_t)(jf + b) * nrad] += j2;
        }
    }
  }
  std::fill(J, J + (size_t)ntot * ntot, 0.0);
  for (int i = 0; i < nang; i++)
    for (int j = 0; j < nang; j++) {
      const int li = lval[i], mi = mval[i], lj = lval[j], mj = mval[j], M = mj - mi;
      for (int L = std::max(std::abs(lj - li) - 2, std::abs(M)); L <= lj + li + 2; L++) {
        const int c = LMind(L, M);
        const double c0 = mod_coeff(lj, mj, L, M, li, mi), c2 = gcoeff(lj, mj, L, M, li, mi);
        for (int b = 0; b < nrad; b++)
          for (int a = 0; a < nrad; a++)
            J[(size_t)(i * nrad + a) + (size_t)(j * nrad + b) * ntot] += c0 * Jaux0[c][a + (size_t)b * nrad] + c2 * Jaux2[c][a + (size_t)b * nrad];
      }
    }
  return 0;
}

int oracle_diatomic_exchange(int nang, const int* lval, const int* mval, int nel, const int* first, const int* last, int nlm,
                             const int* Lmap, const int* Mmap, double Rhalf, const double* const* P0, const double* const* P2,
                             const double* const* Q0, const double* const* Q2, const double* const* ktei, const double* P, double* K) {
  const int nrad = last[nel - 1] + 1, ntot = nang * nrad;
  const size_t n2 = (size_t)nrad * nrad;
  std::fill(K, K + (size_t)ntot * ntot, 0.0);
#ifdef _OPENMP
#pragma omp parallel for collapse(2) schedule(dynamic)
#endif
  for (int jang = 0; jang < nang; jang++)
    for (int kang = 0; kang < nang; kang++) {
      const int lj = lval[jang], mj = mval[jang], lk = lval[kang], mk = mval[kang];
      std::vector<vec> R[4];
      for (int ab = 0; ab < 4; ab++) R[ab].assign(nlm, vec());
      std::vector<bool> couple(nlm, false);
      for (int iang = 0; iang < nang; iang++)
        for (int lang = 0; lang < nang; lang++) {
          const int li = lval[iang], mi = mval[iang], ll = lval[lang], ml = mval[lang];
          const int M = mj - mi;
          if (M != mk - ml) continue;
          double nrm = 0.0;
          for (int b = 0; b < nrad; b++)
            for (int a = 0; a < nrad; a++) {
              const double v = P[(size_t)(iang * nrad + a) + (size_t)(lang * nrad + b) * ntot];
              nrm += v * v;
            }
          if (std::sqrt(nrm) < 10 * 2.220446049250313e-16) continue;
          const int Lmin = std::max(std::max(std::abs(li - lj), std::abs(lk - ll)) - 2, std::abs(M));
          const int Lmax = std::min(li + lj, lk + ll) + 2;
          for (int L = Lmin; L <= Lmax; L++) {
            const double mji = mod_coeff(lj, mj, L, M, li, mi), gji = gcoeff(lj, mj, L, M, li, mi);
            const double mkl = mod_coeff(lk, mk, L, M, ll, ml), gkl = gcoeff(lk, mk, L, M, ll, ml);
            const double cpl[4] = {mji * mkl, -mji * gkl, -gji * mkl, gji * gkl};
            if (cpl[0] == 0.0 && cpl[1] == 0.0 && cpl[2] == 0.0 && cpl[3] == 0.0) continue;
            const int ilm = lmind(nlm, Lmap, Mmap, L, M);
            const double fac = LMfac(Rhalf, L, M);
            for (int ab = 0; ab < 4; ab++) {
              if (R[ab][ilm].empty()) R[ab][ilm].assign(n2, 0.0);
              for (int b = 0; b < nrad; b++)
                for (int a = 0; a < nrad; a++)
                  R[ab][ilm][a + (size_t)b * nrad] += (fac * cpl[ab]) * P[(size_t)(iang * nrad + a) + (size_t)(lang * nrad + b) * ntot];
            }
            couple[ilm] = true;
          }
        }
      for (int iel = 0; iel < nel; iel++)
        for (int jel = 0; jel < nel; jel++) {
          const int fi = first[iel], fj = first[jel], Ni = last[iel] - fi + 1, Nj = last[jel] - fj + 1;
          vec Ksub((size_t)Ni * Nj, 0.0);
          for (int ilm = 0; ilm < nlm; ilm++) {
            if (!couple[ilm]) continue;
            for (int ab = 0; ab < 4; ab++) {
              const vec& Rm = R[ab][ilm];
              const int a_ = ab >> 1, b_ = ab & 1;
              if (iel == jel) {
                const size_t m2 = (size_t)Ni * Nj;
                const double* T = ktei[4 * (ilm * nel + iel) + ab];
                for (int b = 0; b < Nj; b++)
                  for (int a = 0; a < Ni; a++) {
                    const double pv = Rm[(fi + a) + (size_t)(fj + b) * nrad];
                    const size_t c = a + (size_t)b * Ni;
                    for (size_t r = 0; r < m2; r++) Ksub[r] += T[r + c * m2] * pv;
                  }
              } else {
                // iel > jel: iel gets Q, jel gets P (:1703-1707)
                const double* iint = (iel > jel) ? (a_ ? Q2 : Q0)[ilm * nel + iel] : (a_ ? P2 : P0)[ilm * nel + iel];
                const double* jint = (iel > jel) ? (b_ ? P2 : P0)[ilm * nel + jel] : (b_ ? Q2 : Q0)[ilm * nel + jel];
                vec T((size_t)Ni * Nj, 0.0);
                for (int k = 0; k < Nj; k++)
                  for (int b = 0; b < Nj; b++) {
                    const double jv = jint[k + b * Nj];
                    for (int a = 0; a < Ni; a++) T[a + k * Ni] += Rm[(fi + a) + (size_t)(fj + b) * nrad] * jv;
                  }
                for (int k = 0; k < Nj; k++)
                  for (int c = 0; c < Ni; c++) {
                    const double tv = T[c + k * Ni];
                    for (int a = 0; a < Ni; a++) Ksub[a + k * Ni] += iint[a + c * Ni] * tv;
                  }
              }
            }
          }
          for (int b = 0; b < Nj; b++)
            for (int a = 0; a < Ni; a++)
              K[(size_t)(jang * nrad + fi + a) + (size_t)(kang * nrad + fj + b) * ntot] -= Ksub[a + b * Ni];
        }
    }
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Block-restricted forms of the two diatomic operators: the SAME loops as above, evaluated for a
// list of output angular blocks only, so that a basis far too large for the full oracle (BASELINE
// configs[3]: 283 angular functions, 19 810 dummy functions, ~1e13 flop of AXPYs per full exchange)
// can still be checked block by block against the reference's algorithm.
//   exchange: the body of the (jang,kang) loop, /root/reference/src/diatomic/basis.cpp:1576-1726
//   coulomb:  Paux / Jaux of the channels (L, M = m_j - m_i) the block couples to, :1380-1527
// P is the caller's matrix with leading dimension ldp; map (optional) sends a dummy-space index to
// the caller's row/column, -1 for the removed boundary functions (expand_boundaries, :1755-1773).
// Output: nblk blocks of nrad x nrad (column-major) in the dummy space.
// ---------------------------------------------------------------------------------------------
namespace {
struct PView {
  const double* P;
  size_t ld;
  const int* map;
  int nrad;
  void block(int iang, int lang, double* out) const {  // out[a + b*nrad] = P[(iang,a),(lang,b)]
    for (int b = 0; b < nrad; b++) {
      int c = lang * nrad + b;
      if (map) c = map[c];
      for (int a = 0; a < nrad; a++) {
        int r = iang * nrad + a;
        if (map) r = map[r];
        out[a + (size_t)b * nrad] = (r < 0 || c < 0) ? 0.0 : P[(size_t)r + (size_t)c * ld];
      }
    }
  }
};
}  // namespace

int oracle_diatomic_exchange_blocks(int nang, const int* lval, const int* mval, int nel, const int* first, const int* last, int nlm,
                                    const int* Lmap, const int* Mmap, double Rhalf, const double* const* P0, const double* const* P2,
                                    const double* const* Q0, const double* const* Q2, const double* const* ktei, const double* P,
                                    long ldp, const int* map, int nblk, const int* jlist, const int* klist, double* Kout) {
  const int nrad = last[nel - 1] + 1;
  const size_t n2 = (size_t)nrad * nrad;
  const PView Pv{P, (size_t)ldp, map, nrad};
  // Frobenius-norm screen of the input blocks (:1617), evaluated once for all output blocks
  std::vector<char> nz((size_t)nang * nang, 0);
#ifdef _OPENMP
#pragma omp parallel
#endif
  {
    vec blk(n2);
#ifdef _OPENMP
#pragma omp for collapse(2) schedule(dynamic, 8)
#endif
    for (int iang = 0; iang < nang; iang++)
      for (int lang = 0; lang < nang; lang++) {
        Pv.block(iang, lang, blk.data());
        double nrm = 0.0;
        for (size_t q = 0; q < n2; q++) nrm += blk[q] * blk[q];
        nz[(size_t)iang * nang + lang] = std::sqrt(nrm) < 10 * 2.220446049250313e-16 ? 0 : 1;
      }
  }
  int bad = 0;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic)
#endif
  for (int q = 0; q < nblk; q++) {
    const int jang = jlist[q], kang = klist[q];
    const int lj = lval[jang], mj = mval[jang], lk = lval[kang], mk = mval[kang];
    std::vector<vec> R[4];
    for (int ab = 0; ab < 4; ab++) R[ab].assign(nlm, vec());
    std::vector<bool> couple(nlm, false);
    vec Pblk(n2);
    for (int iang = 0; iang < nang; iang++)
      for (int lang = 0; lang < nang; lang++) {
        const int li = lval[iang], mi = mval[iang], ll = lval[lang], ml = mval[lang];
        const int M = mj - mi;
        if (M != mk - ml) continue;
        if (!nz[(size_t)iang * nang + lang]) continue;
        Pv.block(iang, lang, Pblk.data());
        const int Lmin = std::max(std::max(std::abs(li - lj), std::abs(lk - ll)) - 2, std::abs(M));
        const int Lmax = std::min(li + lj, lk + ll) + 2;
        for (int L = Lmin; L <= Lmax; L++) {
          const double mji = mod_coeff(lj, mj, L, M, li, mi), gji = gcoeff(lj, mj, L, M, li, mi);
          const double mkl = mod_coeff(lk, mk, L, M, ll, ml), gkl = gcoeff(lk, mk, L, M, ll, ml);
          const double cpl[4] = {mji * mkl, -mji * gkl, -gji * mkl, gji * gkl};
          if (cpl[0] == 0.0 && cpl[1] == 0.0 && cpl[2] == 0.0 && cpl[3] == 0.0) continue;
          const int ilm = lmind(nlm, Lmap, Mmap, L, M);
          if (ilm < 0) { bad = 1; continue; }
          const double fac = LMfac(Rhalf, L, M);
          for (int ab = 0; ab < 4; ab++) {
            if (R[ab][ilm].empty()) R[ab][ilm].assign(n2, 0.0);
            const double f = fac * cpl[ab];
            double* Rm = R[ab][ilm].data();
            for (size_t e = 0; e < n2; e++) Rm[e] += f * Pblk[e];
          }
          couple[ilm] = true;
        }
      }
    double* Kb = Kout + (size_t)q * n2;
    std::fill(Kb, Kb + n2, 0.0);
    for (int iel = 0; iel < nel; iel++)
      for (int jel = 0; jel < nel; jel++) {
        const int fi = first[iel], fj = first[jel], Ni = last[iel] - fi + 1, Nj = last[jel] - fj + 1;
        vec Ksub((size_t)Ni * Nj, 0.0);
        for (int ilm = 0; ilm < nlm; ilm++) {
          if (!couple[ilm]) continue;
          for (int ab = 0; ab < 4; ab++) {
            const vec& Rm = R[ab][ilm];
            const int a_ = ab >> 1, b_ = ab & 1;
            if (iel == jel) {
              const size_t m2 = (size_t)Ni * Nj;
              const double* T = ktei[4 * (ilm * nel + iel) + ab];
              for (int b = 0; b < Nj; b++)
                for (int a = 0; a < Ni; a++) {
                  const double pv = Rm[(fi + a) + (size_t)(fj + b) * nrad];
                  const size_t c = a + (size_t)b * Ni;
                  for (size_t r = 0; r < m2; r++) Ksub[r] += T[r + c * m2] * pv;
                }
            } else {
              const double* iint = (iel > jel) ? (a_ ? Q2 : Q0)[ilm * nel + iel] : (a_ ? P2 : P0)[ilm * nel + iel];
              const double* jint = (iel > jel) ? (b_ ? P2 : P0)[ilm * nel + jel] : (b_ ? Q2 : Q0)[ilm * nel + jel];
              vec T((size_t)Ni * Nj, 0.0);
              for (int k = 0; k < Nj; k++)
                for (int b = 0; b < Nj; b++) {
                  const double jv = jint[k + b * Nj];
                  for (int a = 0; a < Ni; a++) T[a + k * Ni] += Rm[(fi + a) + (size_t)(fj + b) * nrad] * jv;
                }
              for (int k = 0; k < Nj; k++)
                for (int c = 0; c < Ni; c++) {
                  const double tv = T[c + k * Ni];
                  for (int a = 0; a < Ni; a++) Ksub[a + k * Ni] += iint[a + c * Ni] * tv;
                }
            }
          }
        }
        for (int b = 0; b < Nj; b++)
          for (int a = 0; a < Ni; a++) Kb[(fi + a) + (size_t)(fj + b) * nrad] -= Ksub[a + b * Ni];
      }
  }
  return bad ? -1 : 0;
}

int oracle_diatomic_coulomb_blocks(int nang, const int* lval, const int* mval, int nel, const int* first, const int* last, int nlm,
                                   const int* Lmap, const int* Mmap, double Rhalf, const double* const* P0, const double* const* P2,
                                   const double* const* Q0, const double* const* Q2, const double* const* tei, const double* P,
                                   long ldp, const int* map, int nblk, const int* ilist, const int* jlist, double* Jout) {
  const int nrad = last[nel - 1] + 1;
  const size_t n2 = (size_t)nrad * nrad;
  const PView Pv{P, (size_t)ldp, map, nrad};
  // channels (L, M) the requested blocks couple to (:1499-1527)
  std::vector<std::pair<int, int>> LM;
  auto LMfind = [&](int L, int M) {
    for (size_t c = 0; c < LM.size(); c++)
      if (LM[c].first == L && LM[c].second == M) return (int)c;
    return -1;
  };
  for (int q = 0; q < nblk; q++) {
    const int i = ilist[q], j = jlist[q], li = lval[i], mi = mval[i], lj = lval[j], mj = mval[j], M = mj - mi;
    for (int L = std::max(std::abs(lj - li) - 2, std::abs(M)); L <= lj + li + 2; L++)
      if (LMfind(L, M) < 0) LM.push_back({L, M});
  }
  const int nch = (int)LM.size();
  std::vector<vec> Jaux0(nch, vec(n2, 0.0)), Jaux2(nch, vec(n2, 0.0));
  int bad = 0;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic)
#endif
  for (int c = 0; c < nch; c++) {
    const int L = LM[c].first, M = LM[c].second, ilm = lmind(nlm, Lmap, Mmap, L, M);
    if (ilm < 0) { bad = 1; continue; }
    // ket contraction for this channel (:1380-1403)
    vec Paux0(n2, 0.0), Paux2(n2, 0.0), Pblk(n2);
    for (int k = 0; k < nang; k++)
      for (int l = 0; l < nang; l++) {
        const int lk = lval[k], mk = mval[k], ll = lval[l], ml = mval[l];
        if (mk - ml != M) continue;
        if (L < std::max(std::abs(lk - ll) - 2, std::abs(M)) || L > lk + ll + 2) continue;
        const double c0 = mod_coeff(lk, mk, L, M, ll, ml), c2 = gcoeff(lk, mk, L, M, ll, ml);
        if (c0 == 0.0 && c2 == 0.0) continue;
        Pv.block(k, l, Pblk.data());
        for (size_t e = 0; e < n2; e++) {
          if (c0 != 0.0) Paux0[e] += c0 * Pblk[e];
          if (c2 != 0.0) Paux2[e] += c2 * Pblk[e];
        }
      }
    // radial part (:1411-1489)
    const double fac = LMfac(Rhalf, L, M);
    for (int jel = 0; jel < nel; jel++) {
      const int jf = first[jel], Nj = last[jel] - jf + 1;
      vec Ps0((size_t)Nj * Nj), Ps2((size_t)Nj * Nj);
      for (int b = 0; b < Nj; b++)
        for (int a = 0; a < Nj; a++) {
          Ps0[a + b * Nj] = Paux0[(jf + a) + (size_t)(jf + b) * nrad];
          Ps2[a + b * Nj] = Paux2[(jf + a) + (size_t)(jf + b) * nrad];
        }
      auto trace = [&](const double* D, const vec& Ps) {
        double t = 0.0;
        for (int a = 0; a < Nj; a++)
          for (int b = 0; b < Nj; b++) t += D[a + b * Nj] * Ps[b + a * Nj];
        return t;
      };
      const double jsmall0 = fac * trace(P0[ilm * nel + jel], Ps0), jbig0 = fac * trace(Q0[ilm * nel + jel], Ps0);
      const double jsmall2 = fac * trace(P2[ilm * nel + jel], Ps2), jbig2 = fac * trace(Q2[ilm * nel + jel], Ps2);
      for (int iel = 0; iel < nel; iel++) {
        if (iel == jel) continue;
        const int f = first[iel], Ni = last[iel] - f + 1;
        const double ifac0 = iel < jel ? jbig0 - jbig2 : jsmall0 - jsmall2;
        const double* i0 = iel < jel ? P0[ilm * nel + iel] : Q0[ilm * nel + iel];
        const double* i2 = iel < jel ? P2[ilm * nel + iel] : Q2[ilm * nel + iel];
        for (int b = 0; b < Ni; b++)
          for (int a = 0; a < Ni; a++) {
            Jaux0[c][(f + a) + (size_t)(f + b) * nrad] += i0[a + b * Ni] * ifac0;
            Jaux2[c][(f + a) + (size_t)(f + b) * nrad] -= i2[a + b * Ni] * ifac0;
          }
      }
      const size_t m2 = (size_t)Nj * Nj;
      const double* const* T = tei + 4 * (ilm * nel + jel);
      for (size_t cc = 0; cc < m2; cc++)
        for (size_t r = 0; r < m2; r++) {
          const double j0 = fac * (T[0][r + cc * m2] * Ps0[cc] - T[1][r + cc * m2] * Ps2[cc]);
          const double j2 = fac * (-T[2][r + cc * m2] * Ps0[cc] + T[3][r + cc * m2] * Ps2[cc]);
          const int a = (int)(r % Nj), b = (int)(r / Nj);
          Jaux0[c][(jf + a) + (size_t)(jf + b) * nrad] += j0;
          Jaux2[c][(jf + a) + (size_t)(jf + b) * nrad] += j2;
        }
    }
  }
  if (bad) return -1;
  for (int q = 0; q < nblk; q++) {
    const int i = ilist[q], j = jlist[q], li = lval[i], mi = mval[i], lj = lval[j], mj = mval[j], M = mj - mi;
    double* Jb = Jout + (size_t)q * n2;
    std::fill(Jb, Jb + n2, 0.0);
    for (int L = std::max(std::abs(lj - li) - 2, std::abs(M)); L <= lj + li + 2; L++) {
      const int c = LMfind(L, M);
      const double c0 = mod_coeff(lj, mj, L, M, li, mi), c2 = gcoeff(lj, mj, L, M, li, mi);
      for (size_t e = 0; e < n2; e++) Jb[e] += c0 * Jaux0[c][e] + c2 * Jaux2[c][e];
    }
  }
  return 0;
}

}  // extern "C"
