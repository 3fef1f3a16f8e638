// Host-side finite-element setup for the Fock-build hot path.
//
// Everything in this header runs ONCE per basis (at TwoDBasis construction /
// compute_tei time) and produces the constant tables that are uploaded to the
// device context: radial basis values on the quadrature points, quadrature
// weights, and the stored one-/two-electron radial integral blocks.
//
// Behavioural specification (reference files, read-only):
//   quadrature rule          /root/reference/libhelfem/src/chebyshev.cpp:22-53
//   Lobatto nodes            /root/reference/libhelfem/src/lobatto.cpp:588-743
//   LIP / HIP shape funcs    /root/reference/libhelfem/src/LIPBasis.cpp, HIPBasis.cpp,
//                            LIPBasis_eval.cpp:8-..., HIPBasis_eval.cpp:8-...
//   element bookkeeping      /root/reference/libhelfem/src/FiniteElementBasis.cpp:39-51,117-120,203-211,269-277
//   radial functions B(r)/r  /root/reference/libhelfem/src/RadialBasis.cpp:32-133,575-730
//   in-element 2e integrals  /root/reference/libhelfem/src/quadrature.cpp:23-126
//   radial element grids     /root/reference/libhelfem/src/grid.cpp:18-87
#pragma once
#include <memory>
#include "mat.h"

namespace hfb {
namespace fem {

/// Modified Gauss-Chebyshev rule of the second kind on [-1,1], ascending nodes.
void chebyshev(int n, Vec& x, Vec& w);
/// Gauss-Lobatto nodes on [-1,1], ascending (weights are not used by the basis).
Vec lobatto_nodes(int n);
/// Radial element boundaries, igrid as in the reference (1 linear, 2 quadratic,
/// 3 polynomial, 4 exponential, 5 geometric).
Vec get_grid(double rmax, int num_el, int igrid, double zexp);

/// Primitive polynomial basis on [-1,1]: primbas 4 = LIP, 5 = HIP.
class PolyBasis {
 public:
  PolyBasis() {}
  PolyBasis(int primbas, int nnodes);
  int get_nprim() const { return nprim; }
  int get_nbf() const { return (int)enabled.size(); }
  int get_noverlap() const { return noverlap; }
  int get_id() const { return id; }
  int get_nnodes() const { return (int)x0.size(); }
  const Vec& get_nodes() const { return x0; }
  const std::vector<int>& get_enabled() const { return enabled; }
  void drop_first(bool zero_func, bool zero_deriv);
  void drop_last(bool zero_func, bool zero_deriv);
  /// n-th derivative of the enabled functions at x, (x.size() x nbf), divided by len^n.
  Mat eval_dnf(const Vec& x, int n, double element_length) const;

 private:
  Mat eval_lip(const Vec& x, int n) const;
  Mat eval_prim_dnf(const Vec& x, int n, double element_length) const;
  int id = 0, nprim = 0, noverlap = 0;
  Vec x0;       // nodes
  Vec lipxi;    // HIP: l_i'(x_i)
  std::vector<int> enabled;
};

class FEBasis {
 public:
  FEBasis() {}
  FEBasis(const PolyBasis& poly, const Vec& bval, bool zero_func_left, bool zero_deriv_left,
          bool zero_func_right, bool zero_deriv_right);
  size_t get_nelem() const { return bval.size() - 1; }
  size_t get_nbf() const { return last.back() + 1; }
  size_t get_max_nprim() const { return poly.get_nprim(); }
  size_t get_nprim(size_t iel) const { return get_basis(iel).get_nbf(); }
  void get_idx(size_t iel, size_t& ifirst, size_t& ilast) const { ifirst = first[iel]; ilast = last[iel]; }
  PolyBasis get_basis(size_t iel) const;
  double element_begin(size_t iel) const { return bval[iel]; }
  double element_end(size_t iel) const { return bval[iel + 1]; }
  double element_length(size_t iel) const { return bval[iel + 1] - bval[iel]; }
  double element_midpoint(size_t iel) const { return 0.5 * (bval[iel] + bval[iel + 1]); }
  double scaling_factor(size_t iel) const { return element_length(iel) / 2; }
  Vec eval_coord(const Vec& x, size_t iel) const;
  Vec eval_prim(const Vec& r, size_t iel) const;
  Mat eval_dnf(const Vec& x, int n, size_t iel) const;
  const Vec& get_bval() const { return bval; }
  const PolyBasis& get_poly() const { return poly; }
  /// Max C0/C1 discontinuity across element boundaries (the reference throws above sqrt(eps)).
  double check_bf_continuity() const;

 private:
  PolyBasis poly;
  Vec bval;
  bool zfl = false, zdl = false, zfr = false, zdr = false;
  std::vector<size_t> first, last;
};

class RadialBasis {
 public:
  RadialBasis() {}
  RadialBasis(const FEBasis& fem, int n_quad, int taylor_order);
  size_t Nbf() const { return fem.get_nbf(); }
  size_t Nel() const { return fem.get_nelem(); }
  size_t Nprim(size_t iel) const { return fem.get_nprim(iel); }
  size_t max_Nprim() const { return fem.get_max_nprim(); }
  int get_nquad() const { return (int)xq.size(); }
  void get_idx(size_t iel, size_t& f, size_t& l) const { fem.get_idx(iel, f, l); }
  const Vec& get_xq() const { return xq; }
  const Vec& get_wq() const { return wq; }
  const FEBasis& get_fem() const { return fem; }
  double get_small_r_taylor_cutoff() const { return small_r_taylor_cutoff; }
  double get_taylor_diff() const { return taylor_diff; }
  int get_taylor_order() const { return taylor_order; }

  /// B(r)/r and its first and second r-derivatives at primitive points x: (nx x Nprim(iel))
  Mat get_bf(const Vec& x, size_t iel) const;
  Mat get_df(const Vec& x, size_t iel) const;
  Mat get_lf(const Vec& x, size_t iel) const;
  Mat get_bf(size_t iel) const { return get_bf(xq, iel); }
  Mat get_df(size_t iel) const { return get_df(xq, iel); }
  Mat get_lf(size_t iel) const { return get_lf(xq, iel); }
  Vec get_r(size_t iel) const { return fem.eval_coord(xq, iel); }
  Vec get_wrad(size_t iel) const;

  /// <B_i/r | r^(Rexp+2) | B_j/r> in element iel
  Mat radial_integral(int Rexp, size_t iel) const;
  Mat overlap(size_t iel) const;
  Mat kinetic(size_t iel) const;
  Mat kinetic_l(size_t iel) const;
  /// In-element two-electron block (Np^2 x Np^2), rows (i,j) i-fast, cols (k,l) k-fast.
  Mat twoe_integral(int L, size_t iel) const;

 private:
  void get_taylor(const Vec& r, size_t ntaylor, Mat& val, int ider) const;
  void set_small_r_taylor_cutoff();
  FEBasis fem;
  Vec xq, wq;
  int taylor_order = 0;
  std::vector<Vec> taylor_df;  // [order-1][ifun]
  double small_r_taylor_cutoff = 0.0;
  double taylor_diff = 0.0;
};

/// (ij|kl) -> rows (j,k) cols (i,l): ktei(k*Nj+j, l*Ni+i) = tei(j*Ni+i, l*Nk+k)
/// (/root/reference/libhelfem/src/utils.cpp:130-156)
Mat exchange_tei(const Mat& tei, size_t Ni, size_t Nj, size_t Nk, size_t Nl);

/// In-element two-electron quadrature on bare polynomials
/// (/root/reference/libhelfem/src/quadrature.cpp:92-126); exposed for the known-answer test
/// of /root/reference/src/atomic/inttest.cpp:58-98.
Mat twoe_integral(double rmin, double rmax, const Vec& x, const Vec& wx, const PolyBasis& poly, int L);

}  // namespace fem
}  // namespace hfb
