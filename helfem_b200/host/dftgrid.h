// Host-side mirrors of the reference's product-grid XC operator classes.
//
//   helfem::atomic::dftgrid::DFTGrid     /root/reference/src/atomic/dftgrid.h:160-187
//   helfem::diatomic::dftgrid::DFTGrid   /root/reference/src/diatomic/dftgrid.h:160-187
//
// Same constructor arguments (basis pointer, lang, mang), same eval_Fxc overloads, argument meaning and error
// behaviour.  The tables DFTGridWorker::compute_bf rebuilds for every element of every call in the reference
// (angular rule, Y_lm and d/dtheta Y_lm at the angular points, radial functions at the quadrature points) are
// constants of (basis, lang, mang): they are built once here with the reference's formulas and uploaded through
// hfb200_set_grid; eval_Fxc itself runs on the GPU through the C ABI -- there is no host implementation.
#pragma once
#include "atomic.h"
#include "diatomic.h"

namespace hfb {
namespace atomic {

class DFTGrid {
 public:
  DFTGrid() {}
  DFTGrid(const TwoDBasis* basp_, int lang_, int mang_) : basp(basp_), lang(lang_), mang(mang_) {}
  /// atomic/dftgrid.cpp:810-870
  void eval_Fxc(int x_func, const Vec& x_pars, int c_func, const Vec& c_pars, const Mat& P, Mat& H, double& Exc, double& Nel,
                double& Ekin, double thr);
  /// atomic/dftgrid.cpp:872-928
  void eval_Fxc(int x_func, const Vec& x_pars, int c_func, const Vec& c_pars, const Mat& Pa, const Mat& Pb, Mat& Ha, Mat& Hb,
                double& Exc, double& Nel, double& Ekin, bool beta, double thr);
  /// the same two operators on caller-owned column-major storage of order n (what arma::mat(ptr, n, n, false) wraps): no copies,
  /// H fully overwritten
  void eval_Fxc(int x_func, int c_func, int nspin, size_t n, const double* Pa, const double* Pb, double* Ha, double* Hb, double& Exc,
                double& Nel, double& Ekin, bool beta, double thr);
  int n_angular_points() const { return lang * mang; }

 private:
  void upload();
  const TwoDBasis* basp = nullptr;
  int lang = 0, mang = 0;
  bool uploaded = false;
};

}  // namespace atomic

namespace diatomic {

class DFTGrid {
 public:
  DFTGrid() {}
  DFTGrid(const TwoDBasis* basp_, int lang_, int mang_) : basp(basp_), lang(lang_), mang(mang_) {}
  /// diatomic/dftgrid.cpp:769-810 (pure-space matrices in and out)
  void eval_Fxc(int x_func, const Vec& x_pars, int c_func, const Vec& c_pars, const Mat& P, Mat& H, double& Exc, double& Nel,
                double& Ekin, double thr);
  /// diatomic/dftgrid.cpp:812-856
  void eval_Fxc(int x_func, const Vec& x_pars, int c_func, const Vec& c_pars, const Mat& Pa, const Mat& Pb, Mat& Ha, Mat& Hb,
                double& Exc, double& Nel, double& Ekin, bool beta, double thr);
  /// the same two operators on caller-owned column-major storage of order n (what arma::mat(ptr, n, n, false) wraps): no copies,
  /// H fully overwritten
  void eval_Fxc(int x_func, int c_func, int nspin, size_t n, const double* Pa, const double* Pb, double* Ha, double* Hb, double& Exc,
                double& Nel, double& Ekin, bool beta, double thr);
  int n_angular_points() const { return lang * mang; }

 private:
  void upload();
  const TwoDBasis* basp = nullptr;
  int lang = 0, mang = 0;
  bool uploaded = false;
};

}  // namespace diatomic
}  // namespace hfb
