// Host-side mirror of the reference's `diatomic` operator class (prolate spheroidal coordinates).
//
//   helfem::diatomic::basis::RadialBasis   /root/reference/src/diatomic/basis.cpp:37-285
//   helfem::diatomic::basis::TwoDBasis     /root/reference/src/diatomic/basis.cpp:307-412,1166-1773
//   in-element two-electron quadrature     /root/reference/src/diatomic/quadrature.cpp:22-125
//
// Same method names, argument meaning, index conventions (global index iang*Nrad + irad in the
// "dummy" space; the first radial function of every m != 0 shell is removed in the "pure" space,
// basis.cpp:482-499) and sign conventions (exchange returns -K).  coulomb / exchange run on the GPU
// through the C ABI; there is no host implementation of them here.
#pragma once
#include <stdexcept>
#include <utility>
#include "fem.h"

struct hfb200_ctx;

namespace hfb {
namespace diatomic {

typedef std::pair<int, int> lmidx_t;

class RadialBasis {
 public:
  RadialBasis() {}
  RadialBasis(const fem::FEBasis& fem, int n_quad);
  size_t Nel() const { return fem.get_nelem(); }
  size_t Nbf() const { return fem.get_nbf(); }
  size_t Nprim(size_t iel) const { return fem.get_nprim(iel); }
  int get_nquad() const { return (int)xq.size(); }
  void get_idx(size_t iel, size_t& f, size_t& l) const { fem.get_idx(iel, f, l); }
  /// B(mu), dB/dmu at the quadrature points (nq x Nprim(iel)); mu values and weights (basis.cpp:266-285)
  Mat get_bf(size_t iel) const { return fem.eval_dnf(xq, 0, iel); }
  Mat get_df(size_t iel) const { return fem.eval_dnf(xq, 1, iel); }
  Vec get_r(size_t iel) const { return fem.eval_coord(xq, iel); }
  Vec get_wrad(size_t iel) const {
    Vec w(wq);
    for (double& v : w) v *= fem.scaling_factor(iel);
    return w;
  }
  /// Legendre tables on every quadrature mu (outer points and the sub-interval points of the inner
  /// integrals), what LegendreTable holds in the reference (basis.cpp:229-264,394-398)
  void compute_legendre(int Lmax, int Mmax);
  /// <B_i| sinh(mu) cosh^k(mu) P_L^M(cosh mu) |B_j> in element iel (basis.cpp:193-201); Q likewise (:203-211)
  Mat Plm_integral(int k, size_t iel, int L, int M) const;
  Mat Qlm_integral(int k, size_t iel, int L, int M) const;
  /// all four in-element blocks (alpha,beta) in {0,2}^2 at once: out[2*a+b] (basis.cpp:218-227)
  void twoe_integrals(size_t iel, int L, int M, Mat out[4]) const;

 private:
  Mat legendre_integral(bool Q, int k, size_t iel, int L, int M) const;
  fem::FEBasis fem;
  Vec xq, wq;
  int tabL = -1, tabM = -1;
  // [iel][q][M*(Lmax+1)+L] at the outer points; [iel][isub][q][...] at the inner points
  std::vector<double> Pout, Qout, Pin;
};

class TwoDBasis {
 public:
  TwoDBasis() {}
  /// Reference ctor: diatomic/basis.cpp:307-412
  TwoDBasis(int Z1, int Z2, double Rhalf, const fem::PolyBasis& poly, int n_quad, const Vec& bval,
            const std::vector<int>& lval, const std::vector<int>& mval, int device = 0);
  ~TwoDBasis();
  TwoDBasis(const TwoDBasis&) = delete;
  TwoDBasis& operator=(const TwoDBasis&) = delete;

  size_t Ndummy() const { return lval.size() * radial.Nbf(); }
  size_t Nbf() const;  // pure functions
  size_t Nrad() const { return radial.Nbf(); }
  size_t Nang() const { return lval.size(); }
  double get_Rhalf() const { return Rhalf; }
  const std::vector<int>& get_lval() const { return lval; }
  const std::vector<int>& get_mval() const { return mval; }
  const RadialBasis& get_radial() const { return radial; }
  const std::vector<lmidx_t>& get_lm_map() const { return lm_map; }
  std::vector<size_t> pure_indices() const;
  Mat expand_boundaries(const Mat& Ppure) const;
  Mat remove_boundaries(const Mat& Fnob) const;

  /// diatomic/basis.cpp:1166-1302
  void compute_tei(bool exchange);
  /// J = coulomb(P), diatomic/basis.cpp:1359-1530 (pure-space matrices in and out)
  Mat coulomb(const Mat& P) const;
  /// -K = exchange(P), diatomic/basis.cpp:1532-1733
  Mat exchange(const Mat& P) const;
  void coulomb(const double* P, double* J) const;   // raw n x n column-major buffers, n = Nbf()
  void exchange(const double* P, double* K) const;

  // stored integrals, index [ilm*Nel + iel]; two-electron blocks [4*(ilm*Nel+iel) + 2a+b]
  const std::vector<Mat>& get_disjoint(int which) const { return disjoint[which]; }  // 0 P0, 1 P2, 2 Q0, 3 Q2
  const std::vector<Mat>& get_prim_tei() const { return prim_tei; }
  const std::vector<Mat>& get_prim_ktei() const { return prim_ktei; }
  /// multi-GPU: this object contracts only the integral families r % nranks == rank (see hfb200_set_shard)
  void set_shard(int rank, int nranks);  // throws std::logic_error once the device context exists (it is sharded at creation)
  /// with a shard set: compute_tei only builds the stored integrals of the owned families (1/nranks of the setup
  /// time and host memory); the others stay empty and are never read
  void set_owned_only(bool on) { owned_only = on; }
  hfb200_ctx* context() const;

 private:
  int Z1 = 0, Z2 = 0, device = 0;
  double Rhalf = 0.0;
  std::vector<int> lval, mval;
  RadialBasis radial;
  std::vector<lmidx_t> lm_map;  // sorted (L,|M|)
  std::vector<Mat> disjoint[4];
  std::vector<Mat> prim_tei, prim_ktei;
  int shard_rank = 0, shard_n = 1;
  bool owned_only = false;
  mutable hfb200_ctx* ctx = nullptr;
  mutable bool tei_uploaded = false;
};

}  // namespace diatomic
}  // namespace hfb
