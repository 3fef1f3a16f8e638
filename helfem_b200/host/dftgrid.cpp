// Host-side mirrors of atomic / diatomic DFTGrid (see dftgrid.h).
#include "dftgrid.h"

#include <cmath>
#include <complex>
#include <string>
#include "../../include/helfem_b200.h"
#include "angular.h"

namespace hfb {
namespace {

void throw_status(int rc) {
  if (rc == HFB200_OK) return;
  const std::string msg = std::string(hfb200_last_error()) + "\n";
  if (rc == HFB200_EINVAL || rc == HFB200_ENOTEI) throw std::logic_error(msg);
  throw std::runtime_error(msg);
}

// Constants of DFTGridWorker::compute_bf for all elements at once:
//   angular rule            general/angular.cpp:64-71 (Chebyshev in cos(theta) x uniform phi)
//   Y_lm at the points      atomic/TwoDBasis.cpp:1368-1383 (eval_bf)
//   d/dtheta Y_lm           atomic/TwoDBasis.cpp:1407-1419 (eval_df): m cot(th) Y_lm + sqrt((l-m)(l+m+1)) e^{-i phi} Y_{l,m+1}
template <class Radial>
void upload_grid(hfb200_ctx* ctx, const Radial& radial, const std::vector<int>& lval, const std::vector<int>& mval, int lang, int mang) {
  Vec cth, phi, wang;
  angular::angular_chebyshev(lang, mang, cth, phi, wang);
  const size_t npts = cth.size(), nang = lval.size();
  std::vector<double> Y(2 * nang * npts), dY(2 * nang * npts);
  for (size_t ia = 0; ia < npts; ia++) {
    const double cotth = cth[ia] / std::sqrt(1.0 - cth[ia] * cth[ia]);
    for (size_t a = 0; a < nang; a++) {
      const int l = lval[a], m = mval[a];
      const std::complex<double> sph = angular::spherical_harmonics(l, m, cth[ia], phi[ia]);
      std::complex<double> angfac = (double)m * cotth * sph;
      if (m < l)
        angfac += std::sqrt((double)((l - m) * (l + m + 1))) * std::exp(std::complex<double>(0.0, -phi[ia])) *
                  angular::spherical_harmonics(l, m + 1, cth[ia], phi[ia]);
      Y[2 * (a + nang * ia)] = sph.real();
      Y[2 * (a + nang * ia) + 1] = sph.imag();
      dY[2 * (a + nang * ia)] = angfac.real();
      dY[2 * (a + nang * ia) + 1] = angfac.imag();
    }
  }
  const size_t nel = radial.Nel(), nq = (size_t)radial.get_nquad();
  std::vector<double> r, wr, bf, df;
  for (size_t iel = 0; iel < nel; iel++) {
    const Vec re = radial.get_r(iel), we = radial.get_wrad(iel);
    r.insert(r.end(), re.begin(), re.end());
    wr.insert(wr.end(), we.begin(), we.end());
    const Mat b = radial.get_bf(iel), d = radial.get_df(iel);
    bf.insert(bf.end(), b.d.begin(), b.d.end());
    df.insert(df.end(), d.d.begin(), d.d.end());
  }
  throw_status(hfb200_set_grid(ctx, (int)nq, r.data(), wr.data(), bf.data(), df.data(), (int)npts, cth.data(), wang.data(), Y.data(),
                               dY.data()));
}

void check_pars(const Vec& x_pars, const Vec& c_pars) {
  if (!x_pars.empty() || !c_pars.empty())
    throw std::runtime_error("External functional parameters are not supported by the device functionals.\n");
}

}  // namespace

namespace atomic {

void DFTGrid::upload() {
  if (uploaded) return;
  upload_grid(basp->context(), basp->get_radial(), basp->get_lval(), basp->get_mval(), lang, mang);
  uploaded = true;
}

void DFTGrid::eval_Fxc(int x_func, const Vec& x_pars, int c_func, const Vec& c_pars, const Mat& P, Mat& H, double& Exc, double& Nel,
                       double& Ekin, double thr) {
  check_pars(x_pars, c_pars);
  if (!P.d.size()) throw std::runtime_error("Error - density matrix is empty!\n");
  if (P.nr != basp->Nbf() || P.nc != basp->Nbf()) throw std::logic_error("Density matrix does not match basis set!\n");
  upload();
  H.zeros(P.nr, P.nr);
  throw_status(hfb200_atomic_eval_fxc(basp->context(), x_func, c_func, 1, P.ptr(), nullptr, (int)P.nr, thr, H.ptr(), nullptr, (int)H.nr,
                                      0, &Exc, &Nel, &Ekin));
}

void DFTGrid::eval_Fxc(int x_func, const Vec& x_pars, int c_func, const Vec& c_pars, const Mat& Pa, const Mat& Pb, Mat& Ha, Mat& Hb,
                       double& Exc, double& Nel, double& Ekin, bool beta, double thr) {
  check_pars(x_pars, c_pars);
  if (!Pa.d.size() || !Pb.d.size()) throw std::runtime_error("Error - density matrix is empty!\n");
  if (Pa.nr != basp->Nbf() || Pa.nc != basp->Nbf() || Pb.nr != Pa.nr || Pb.nc != Pa.nc)
    throw std::logic_error("Density matrix does not match basis set!\n");
  upload();
  Ha.zeros(Pa.nr, Pa.nr);
  Hb.zeros(Pb.nr, Pb.nr);
  throw_status(hfb200_atomic_eval_fxc(basp->context(), x_func, c_func, 2, Pa.ptr(), Pb.ptr(), (int)Pa.nr, thr, Ha.ptr(), Hb.ptr(),
                                      (int)Ha.nr, beta ? 1 : 0, &Exc, &Nel, &Ekin));
}

void DFTGrid::eval_Fxc(int x_func, int c_func, int nspin, size_t n, const double* Pa, const double* Pb, double* Ha, double* Hb, double& Exc,
                       double& Nel, double& Ekin, bool beta, double thr) {
  if (!Pa || (nspin == 2 && !Pb)) throw std::runtime_error("Error - density matrix is empty!\n");
  if (n != basp->Nbf()) throw std::logic_error("Density matrix does not match basis set!\n");
  upload();
  throw_status(hfb200_atomic_eval_fxc(basp->context(), x_func, c_func, nspin, Pa, Pb, (int)n, thr, Ha, Hb, (int)n, beta ? 1 : 0, &Exc, &Nel, &Ekin));
}


}  // namespace atomic

namespace diatomic {

void DFTGrid::upload() {
  if (uploaded) return;
  upload_grid(basp->context(), basp->get_radial(), basp->get_lval(), basp->get_mval(), lang, mang);
  uploaded = true;
}

void DFTGrid::eval_Fxc(int x_func, const Vec& x_pars, int c_func, const Vec& c_pars, const Mat& P, Mat& H, double& Exc, double& Nel,
                       double& Ekin, double thr) {
  check_pars(x_pars, c_pars);
  if (!P.d.size()) throw std::runtime_error("Error - density matrix is empty!\n");
  if (P.nr != basp->Nbf() || P.nc != basp->Nbf()) throw std::logic_error("Matrix does not have expected size!\n");
  upload();
  H.zeros(P.nr, P.nr);
  throw_status(hfb200_diatomic_eval_fxc(basp->context(), x_func, c_func, 1, P.ptr(), nullptr, (int)P.nr, thr, H.ptr(), nullptr,
                                        (int)H.nr, 0, &Exc, &Nel, &Ekin));
}

void DFTGrid::eval_Fxc(int x_func, const Vec& x_pars, int c_func, const Vec& c_pars, const Mat& Pa, const Mat& Pb, Mat& Ha, Mat& Hb,
                       double& Exc, double& Nel, double& Ekin, bool beta, double thr) {
  check_pars(x_pars, c_pars);
  if (!Pa.d.size() || !Pb.d.size()) throw std::runtime_error("Error - density matrix is empty!\n");
  if (Pa.nr != basp->Nbf() || Pa.nc != basp->Nbf() || Pb.nr != Pa.nr || Pb.nc != Pa.nc)
    throw std::logic_error("Matrix does not have expected size!\n");
  upload();
  Ha.zeros(Pa.nr, Pa.nr);
  Hb.zeros(Pb.nr, Pb.nr);
  throw_status(hfb200_diatomic_eval_fxc(basp->context(), x_func, c_func, 2, Pa.ptr(), Pb.ptr(), (int)Pa.nr, thr, Ha.ptr(), Hb.ptr(),
                                        (int)Ha.nr, beta ? 1 : 0, &Exc, &Nel, &Ekin));
}

void DFTGrid::eval_Fxc(int x_func, int c_func, int nspin, size_t n, const double* Pa, const double* Pb, double* Ha, double* Hb, double& Exc,
                       double& Nel, double& Ekin, bool beta, double thr) {
  if (!Pa || (nspin == 2 && !Pb)) throw std::runtime_error("Error - density matrix is empty!\n");
  if (n != basp->Nbf()) throw std::logic_error("Matrix does not have expected size!\n");
  upload();
  throw_status(hfb200_diatomic_eval_fxc(basp->context(), x_func, c_func, nspin, Pa, Pb, (int)n, thr, Ha, Hb, (int)n, beta ? 1 : 0, &Exc, &Nel, &Ekin));
}


}  // namespace diatomic
}  // namespace hfb
