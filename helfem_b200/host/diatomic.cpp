// Host-side mirror of the reference's `diatomic` operator class (see diatomic.h).
#include "diatomic.h"
#include "famshard.h"

#include <algorithm>
#include <cmath>
#include <set>
#include <string>
#include "../../include/helfem_b200.h"
#include "legendre.h"

namespace hfb {
namespace diatomic {

static void throw_status(int rc) {
  if (rc == HFB200_OK) return;
  const std::string msg = std::string(hfb200_last_error()) + "\n";
  if (rc == HFB200_EINVAL || rc == HFB200_ENOTEI) throw std::logic_error(msg);
  throw std::runtime_error(msg);
}

// ---------------------------------------------------------------------------------------------
// RadialBasis
// ---------------------------------------------------------------------------------------------
RadialBasis::RadialBasis(const fem::FEBasis& fem_, int n_quad) : fem(fem_) { fem::chebyshev(n_quad, xq, wq); }

void RadialBasis::compute_legendre(int Lmax, int Mmax) {
  tabL = Lmax;
  tabM = Mmax;
  const size_t nq = xq.size(), nel = Nel(), sz = (size_t)(Mmax + 1) * (Lmax + 1);
  Pout.assign(nel * nq * sz, 0.0);
  Qout.assign(nel * nq * sz, 0.0);
  Pin.assign(nel * nq * nq * sz, 0.0);
#ifdef _OPENMP
#pragma omp parallel for collapse(2) schedule(dynamic)
#endif
  for (size_t iel = 0; iel < nel; iel++)
    for (size_t ip = 0; ip < nq; ip++) {
      const double mumin0 = fem.element_begin(iel), mumax0 = fem.element_end(iel);
      const double mid0 = 0.5 * (mumax0 + mumin0), len0 = 0.5 * (mumax0 - mumin0);
      const double mu = mid0 + len0 * xq[ip];
      legendre::plm_array(Lmax, Mmax, mu, &Pout[(iel * nq + ip) * sz]);
      legendre::qlm_array(Lmax, Mmax, mu, &Qout[(iel * nq + ip) * sz]);
      // sub-interval [mu_{ip-1}, mu_ip] of the inner integral (diatomic/quadrature.cpp:72-76)
      const double lo = ip == 0 ? mumin0 : mid0 + len0 * xq[ip - 1];
      const double smid = 0.5 * (mu + lo), slen = 0.5 * (mu - lo);
      for (size_t q = 0; q < nq; q++) legendre::plm_array(Lmax, Mmax, smid + slen * xq[q], &Pin[((iel * nq + ip) * nq + q) * sz]);
    }
}

// fem.matrix_element(iel, false, false, xq, wq, f) with f = sinh cosh^k {P,Q}_L^M   (basis.cpp:193-211)
Mat RadialBasis::legendre_integral(bool Q, int k, size_t iel, int L, int M) const {
  if (L > tabL || M > tabM) throw std::logic_error("Legendre table too small\n");
  const size_t nq = xq.size(), sz = (size_t)(tabM + 1) * (tabL + 1), np = Nprim(iel);
  const Vec mu = fem.eval_coord(xq, iel);
  const Mat bf = fem.eval_dnf(xq, 0, iel);
  const double sf = fem.scaling_factor(iel);
  Mat out(np, np);
  const std::vector<double>& tab = Q ? Qout : Pout;
  for (size_t p = 0; p < nq; p++) {
    double w = wq[p] * sf * sinh(mu[p]) * tab[(iel * nq + p) * sz + (size_t)M * (tabL + 1) + L];
    if (k) w *= pow(cosh(mu[p]), k);
    for (size_t j = 0; j < np; j++) {
      const double wj = w * bf(p, j);
      for (size_t i = 0; i < np; i++) out(i, j) += bf(p, i) * wj;
    }
  }
  return out;
}
Mat RadialBasis::Plm_integral(int k, size_t iel, int L, int M) const { return legendre_integral(false, k, iel, L, M); }
Mat RadialBasis::Qlm_integral(int k, size_t iel, int L, int M) const { return legendre_integral(true, k, iel, L, M); }

// diatomic/quadrature.cpp:22-125: wrk(k,l) = sum_p w_p sinh cosh^k Q_L^M (B_i B_j)(mu_p) * inner_l(kl; mu_p),
// inner_l = int_{mu_min}^{mu_p} sinh cosh^l P_L^M B_k B_l, accumulated over sub-intervals each with a fresh
// rule; tei(alpha,beta) = wrk(alpha,beta) + wrk(beta,alpha)^T.
void RadialBasis::twoe_integrals(size_t iel, int L, int M, Mat out[4]) const {
  if (L > tabL || M > tabM) throw std::logic_error("Legendre table too small\n");
  const size_t nq = xq.size(), sz = (size_t)(tabM + 1) * (tabL + 1), lmoff = (size_t)M * (tabL + 1) + L;
  const fem::PolyBasis p = fem.get_basis(iel);
  const size_t np = p.get_nbf(), np2 = np * np;
  const double mumin0 = fem.element_begin(iel), mumax0 = fem.element_end(iel);
  const double mid0 = 0.5 * (mumax0 + mumin0), len0 = 0.5 * (mumax0 - mumin0);
  // inner integrals for l = 0, 2: inner[l][p][a + b*np]
  std::vector<double> inner[2];
  inner[0].assign(nq * np2, 0.0);
  inner[1].assign(nq * np2, 0.0);
  std::vector<double> acc0(np2, 0.0), acc2(np2, 0.0);
  for (size_t ip = 0; ip < nq; ip++) {
    const double hi = mid0 + len0 * xq[ip];
    const double lo = ip == 0 ? mumin0 : mid0 + len0 * xq[ip - 1];
    const double smid = 0.5 * (hi + lo), slen = 0.5 * (hi - lo);
    Vec mu(nq), xpoly(nq);
    for (size_t q = 0; q < nq; q++) {
      mu[q] = smid + slen * xq[q];
      xpoly[q] = (mu[q] - mid0) / len0;
    }
    const Mat bf = p.eval_dnf(xpoly, 0, len0);
    for (size_t q = 0; q < nq; q++) {
      const double ch = cosh(mu[q]);
      const double w0 = wq[q] * slen * sinh(mu[q]) * Pin[((iel * nq + ip) * nq + q) * sz + lmoff];
      const double w2 = w0 * ch * ch;
      for (size_t b = 0; b < np; b++) {
        const double f0 = w0 * bf(q, b), f2 = w2 * bf(q, b);
        for (size_t a = 0; a < np; a++) {
          acc0[a + b * np] += bf(q, a) * f0;
          acc2[a + b * np] += bf(q, a) * f2;
        }
      }
    }
    std::copy(acc0.begin(), acc0.end(), inner[0].begin() + ip * np2);
    std::copy(acc2.begin(), acc2.end(), inner[1].begin() + ip * np2);
  }
  // outer products, weights for k = 0, 2
  const Mat bfo = p.eval_dnf(xq, 0, len0);
  Mat W[2][2];  // [k][l]
  for (int k = 0; k < 2; k++)
    for (int l = 0; l < 2; l++) W[k][l] = Mat(np2, np2);
  std::vector<double> prod(np2);
  for (size_t ip = 0; ip < nq; ip++) {
    const double mu = mid0 + len0 * xq[ip], ch = cosh(mu);
    const double wk0 = wq[ip] * len0 * sinh(mu) * Qout[(iel * nq + ip) * sz + lmoff];
    const double wk[2] = {wk0, wk0 * ch * ch};
    for (size_t fi = 0; fi < np; fi++)
      for (size_t fj = 0; fj < np; fj++) prod[fi * np + fj] = bfo(ip, fi) * bfo(ip, fj);
    for (int l = 0; l < 2; l++) {
      const double* in = &inner[l][ip * np2];
      for (int k = 0; k < 2; k++) {
        Mat& Wm = W[k][l];
        for (size_t c = 0; c < np2; c++) {
          const double v = wk[k] * in[c];
          if (v == 0.0) continue;
          double* col = &Wm(0, c);
          for (size_t r = 0; r < np2; r++) col[r] += prod[r] * v;
        }
      }
    }
  }
  for (int a = 0; a < 2; a++)
    for (int b = 0; b < 2; b++) {
      Mat t(np2, np2);
      for (size_t c = 0; c < np2; c++)
        for (size_t r = 0; r < np2; r++) t(r, c) = W[a][b](r, c) + W[b][a](c, r);
      out[2 * a + b] = t;
    }
}

// ---------------------------------------------------------------------------------------------
// TwoDBasis
// ---------------------------------------------------------------------------------------------
TwoDBasis::TwoDBasis(int Z1_, int Z2_, double Rhalf_, const fem::PolyBasis& poly, int n_quad, const Vec& bval,
                     const std::vector<int>& lval_, const std::vector<int>& mval_, int device_)
    : Z1(Z1_), Z2(Z2_), device(device_), Rhalf(Rhalf_), lval(lval_), mval(mval_) {
  // diatomic/basis.cpp:313-319: sigma orbitals may reach the axis; cleaned up for m != 0 by pure_indices()
  fem::FEBasis fe(poly, bval, false, false, true, true);
  radial = RadialBasis(fe, n_quad);
  // (L,|M|) families of all couplings (basis.cpp:336-376)
  std::set<lmidx_t> fams;
  int Lmax = 0, Mmax = 0;
  for (size_t i = 0; i < lval.size(); i++)
    for (size_t j = 0; j < lval.size(); j++) {
      const int M = mval[j] - mval[i];
      for (int L = std::max(std::abs(lval[j] - lval[i]) - 2, std::abs(M)); L <= lval[j] + lval[i] + 2; L++) {
        fams.insert(lmidx_t(L, std::abs(M)));
        Lmax = std::max(Lmax, L);
        Mmax = std::max(Mmax, std::abs(M));
      }
    }
  lm_map.assign(fams.begin(), fams.end());
  radial.compute_legendre(Lmax, Mmax);
}

TwoDBasis::~TwoDBasis() {
  if (ctx) hfb200_destroy(ctx);
}

size_t TwoDBasis::Nbf() const {
  size_t n = 0;
  for (size_t i = 0; i < mval.size(); i++) n += radial.Nbf() - (mval[i] != 0 ? 1 : 0);
  return n;
}

// diatomic/basis.cpp:482-499
std::vector<size_t> TwoDBasis::pure_indices() const {
  std::vector<size_t> idx;
  const size_t Nr = radial.Nbf();
  for (size_t i = 0; i < mval.size(); i++)
    for (size_t a = (mval[i] != 0 ? 1 : 0); a < Nr; a++) idx.push_back(i * Nr + a);
  return idx;
}

Mat TwoDBasis::expand_boundaries(const Mat& Pp) const {
  if (Pp.nr != Nbf() || Pp.nc != Nbf()) throw std::logic_error("Matrix does not have expected size!\n");
  const std::vector<size_t> idx = pure_indices();
  Mat P(Ndummy(), Ndummy());
  for (size_t j = 0; j < idx.size(); j++)
    for (size_t i = 0; i < idx.size(); i++) P(idx[i], idx[j]) = Pp(i, j);
  return P;
}

Mat TwoDBasis::remove_boundaries(const Mat& F) const {
  if (F.nr != Ndummy() || F.nc != Ndummy()) throw std::logic_error("Matrix does not have expected size!\n");
  const std::vector<size_t> idx = pure_indices();
  Mat Fp(idx.size(), idx.size());
  for (size_t j = 0; j < idx.size(); j++)
    for (size_t i = 0; i < idx.size(); i++) Fp(i, j) = F(idx[i], idx[j]);
  return Fp;
}

// diatomic/basis.cpp:1166-1302
void TwoDBasis::compute_tei(bool exchange) {
  const size_t Nel = radial.Nel(), nlm = lm_map.size();
  for (int w = 0; w < 4; w++) disjoint[w].assign(Nel * nlm, Mat());
  prim_tei.assign(4 * Nel * nlm, Mat());
  prim_ktei.assign(exchange ? 4 * Nel * nlm : 0, Mat());
  const std::vector<int> own = family_owner((int)nlm, owned_only ? shard_n : 1);  // the map the C ABI applies (host/famshard.h)
#ifdef _OPENMP
#pragma omp parallel for collapse(2) schedule(dynamic)
#endif
  for (size_t ilm = 0; ilm < nlm; ilm++)
    for (size_t iel = 0; iel < Nel; iel++) {
      if (owned_only && own[ilm] != shard_rank) continue;  // family held by another rank
      const int L = lm_map[ilm].first, M = lm_map[ilm].second;
      const size_t idx = ilm * Nel + iel;
      disjoint[0][idx] = radial.Plm_integral(0, iel, L, M);
      disjoint[1][idx] = radial.Plm_integral(2, iel, L, M);
      disjoint[2][idx] = radial.Qlm_integral(0, iel, L, M);
      disjoint[3][idx] = radial.Qlm_integral(2, iel, L, M);
      Mat t[4];
      radial.twoe_integrals(iel, L, M, t);
      const size_t Ni = radial.Nprim(iel);
      for (int ab = 0; ab < 4; ab++) {
        if (exchange) prim_ktei[4 * idx + ab] = fem::exchange_tei(t[ab], Ni, Ni, Ni, Ni);
        prim_tei[4 * idx + ab] = t[ab];
      }
    }
  tei_uploaded = false;
}

void TwoDBasis::set_shard(int rank, int nranks) {
  if (ctx) throw std::logic_error("set_shard must precede the first use of the device context\n");
  shard_rank = rank;
  shard_n = nranks;
}

hfb200_ctx* TwoDBasis::context() const {
  if (!ctx) {
    const size_t Nel = radial.Nel();
    std::vector<int> first(Nel), last(Nel);
    for (size_t iel = 0; iel < Nel; iel++) {
      size_t f, l;
      radial.get_idx(iel, f, l);
      first[iel] = (int)f;
      last[iel] = (int)l;
    }
    throw_status(hfb200_diatomic_create(&ctx, device, (int)lval.size(), lval.data(), mval.data(), (int)Nel, first.data(),
                                        last.data(), Rhalf));
    throw_status(hfb200_set_shard(ctx, shard_rank, shard_n));
  }
  if (!tei_uploaded && !prim_tei.empty()) {
    std::vector<int> Lm, Mm;
    for (const auto& p : lm_map) { Lm.push_back(p.first); Mm.push_back(p.second); }
    std::vector<const double*> d[4], t, k;
    for (int w = 0; w < 4; w++)
      for (const auto& m : disjoint[w]) d[w].push_back(m.ptr());
    for (const auto& m : prim_tei) t.push_back(m.ptr());
    for (const auto& m : prim_ktei) k.push_back(m.ptr());
    throw_status(hfb200_diatomic_set_tei(ctx, (int)lm_map.size(), Lm.data(), Mm.data(), d[0].data(), d[1].data(), d[2].data(),
                                         d[3].data(), t.data(), k.empty() ? nullptr : k.data()));
    tei_uploaded = true;
  }
  return ctx;
}

Mat TwoDBasis::coulomb(const Mat& P) const {
  if (prim_tei.empty()) throw std::logic_error("Primitive teis have not been computed!\n");
  if (P.nr != Nbf() || P.nc != Nbf()) throw std::logic_error("Matrix does not have expected size!\n");
  Mat J(P.nr, P.nc);
  throw_status(hfb200_diatomic_coulomb(context(), P.ptr(), (int)P.nr, J.ptr(), (int)J.nr));
  return J;
}

Mat TwoDBasis::exchange(const Mat& P) const {
  if (prim_ktei.empty()) throw std::logic_error("Primitive teis have not been computed!\n");
  if (P.nr != Nbf() || P.nc != Nbf()) throw std::logic_error("Matrix does not have expected size!\n");
  Mat K(P.nr, P.nc);
  throw_status(hfb200_diatomic_exchange(context(), P.ptr(), (int)P.nr, K.ptr(), (int)K.nr));
  return K;
}


// Raw-pointer forms of the two operators (n x n column-major, n = Nbf()): what the flat C entry points of
// hostapi.cpp call, so that a caller's buffers go straight to the C ABI without an intermediate Mat copy.
void TwoDBasis::coulomb(const double* P, double* J) const {
  if (prim_tei.empty()) throw std::logic_error("Primitive teis have not been computed!\n");
  throw_status(hfb200_diatomic_coulomb(context(), P, (int)Nbf(), J, (int)Nbf()));
}
void TwoDBasis::exchange(const double* P, double* K) const {
  if (prim_ktei.empty()) throw std::logic_error("Primitive teis have not been computed!\n");
  throw_status(hfb200_diatomic_exchange(context(), P, (int)Nbf(), K, (int)Nbf()));
}

}  // namespace diatomic
}  // namespace hfb
