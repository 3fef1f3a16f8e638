// Host-side mirror of the reference's spherically-averaged-atom operator classes.
//
//   helfem::sadatom::basis::TwoDBasis    /root/reference/src/sadatom/basis.h:38-170
//   helfem::sadatom::dftgrid::DFTGrid    /root/reference/src/sadatom/dftgrid.h:150-176
//
// Same method names, argument meaning, sign/scale conventions and error behaviour
// (std::logic_error for missing integrals / size mismatch, std::runtime_error for an
// unsupported functional); hfb::Mat/hfb::Cube stand in for arma::mat/arma::cube (same
// column-major layout, see mat.h).  coulomb / exchange / eval_Fxc run on the GPU through the
// C ABI of include/helfem_b200.h; there is no host implementation of them here.
#pragma once
#include <memory>
#include "fem.h"

struct hfb200_ctx;

namespace hfb {
namespace sadatom {

class TwoDBasis {
 public:
  TwoDBasis() {}
  /// Reference ctor: sadatom/basis.cpp:40-53 (Z, poly, zeroder, n_quad, bval, taylor_order, lmax);
  /// the nuclear model is always the point nucleus here.
  TwoDBasis(int Z, const fem::PolyBasis& poly, bool zeroder, int n_quad, const Vec& bval, int taylor_order,
            int lmax, int device = 0);
  ~TwoDBasis();
  TwoDBasis(const TwoDBasis&) = delete;
  TwoDBasis& operator=(const TwoDBasis&) = delete;

  size_t Nbf() const { return radial.Nbf(); }
  int charge() const { return Z; }
  int get_lmax() const { return lmax; }
  size_t get_rad_Nel() const { return radial.Nel(); }
  const fem::RadialBasis& get_radial() const { return radial; }

  Mat radial_integral(int Rexp) const;
  Mat overlap() const { return radial_integral(0); }
  Mat kinetic() const;
  Mat kinetic_l() const;
  Mat nuclear() const;

  /// sadatom/basis.cpp:181-218.  nL<=0 means all L = 0..2*lmax as in the reference; a pure
  /// functional only ever reads L=0 (coulomb, basis.cpp:306), so callers may pass nL=1.
  void compute_tei(int nL = 0, bool exchange = true);
  /// J = coulomb(P), sadatom/basis.cpp:296-358
  Mat coulomb(const Mat& P) const;
  /// -K = exchange(P), sadatom/basis.cpp:360-526
  Cube exchange(const Cube& P) const;

  Mat eval_bf(size_t iel) const { return radial.get_bf(iel); }
  Mat eval_df(size_t iel) const { return radial.get_df(iel); }
  Mat eval_lf(size_t iel) const { return radial.get_lf(iel); }
  Vec get_r(size_t iel) const { return radial.get_r(iel); }
  Vec get_wrad(size_t iel) const { return radial.get_wrad(iel); }
  std::vector<size_t> bf_list(size_t iel) const;

  // stored integrals (for tests and for the device upload)
  const std::vector<Mat>& get_disjoint_L() const { return disjoint_L; }
  const std::vector<Mat>& get_disjoint_m1L() const { return disjoint_m1L; }
  const std::vector<Mat>& get_prim_tei() const { return prim_tei; }
  const std::vector<Mat>& get_prim_ktei() const { return prim_ktei; }
  int get_nL() const { return nL_stored; }

  /// Device context (created lazily on first use; throws std::runtime_error without a GPU).
  hfb200_ctx* context() const;

 private:
  int Z = 0, lmax = 0, device = 0;
  fem::RadialBasis radial;
  int nL_stored = 0;
  std::vector<Mat> disjoint_L, disjoint_m1L, prim_tei, prim_ktei;  // index [L*Nel+iel]
  mutable hfb200_ctx* ctx = nullptr;
  mutable bool tei_uploaded = false;
};

class DFTGrid {
 public:
  DFTGrid() {}
  explicit DFTGrid(const TwoDBasis* basp_) : basp(basp_) {}
  /// sadatom/dftgrid.cpp:874-939
  void eval_Fxc(int x_func, const Vec& x_pars, int c_func, const Vec& c_pars, const Cube& P, Cube& H,
                double& Exc, double& Nel, double thr);
  /// sadatom/dftgrid.cpp:941-993
  void eval_Fxc(int x_func, const Vec& x_pars, int c_func, const Vec& c_pars, const Cube& Pa, const Cube& Pb,
                Cube& Ha, Cube& Hb, double& Exc, double& Nel, bool beta, double thr);

 private:
  const TwoDBasis* basp = nullptr;
};

}  // namespace sadatom
}  // namespace hfb
