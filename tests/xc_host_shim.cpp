// TEST-ONLY: compiles the product's device XC source (helfem_b200/csrc/xc_point.cuh) for the host so
// the CPU test-suite can compare the very same formulas against the oracle without a GPU.
// Never linked into libhelfem_b200.so.
#include "../helfem_b200/csrc/xc_point.cuh"

using namespace hfb;

template <int FAM>
static void run(int x, int c, int nspin, int n, double dens_thr, const double* rho, const double* sigma, const double* tau,
                double* exc, double* vrho, double* vsigma, double* vtau) {
  const xc::XcThresholds thr = xc::make_thresholds(dens_thr);
  for (int i = 0; i < n; i++) {
    if (nspin == 1) {
      double e = 0, vr = 0, vs = 0, vt = 0;
      const double sg = FAM >= xc::FAM_GGA ? sigma[i] : 0.0, ta = FAM >= xc::FAM_MGGA ? tau[i] : 0.0;
      xc::xc_point_unpol<FAM>(x, thr, rho[i], sg, ta, e, vr, vs, vt);
      xc::xc_point_unpol<FAM>(c, thr, rho[i], sg, ta, e, vr, vs, vt);
      exc[i] = e; vrho[i] = vr;
      if (FAM >= xc::FAM_GGA) vsigma[i] = vs;
      if (FAM >= xc::FAM_MGGA) vtau[i] = vt;
    } else {
      double e = 0, vr[2] = {0, 0}, vs[3] = {0, 0, 0}, vt[2] = {0, 0};
      double r2[2] = {rho[2 * i], rho[2 * i + 1]}, s3[3] = {0, 0, 0}, t2[2] = {0, 0};
      if (FAM >= xc::FAM_GGA) { s3[0] = sigma[3 * i]; s3[1] = sigma[3 * i + 1]; s3[2] = sigma[3 * i + 2]; }
      if (FAM >= xc::FAM_MGGA) { t2[0] = tau[2 * i]; t2[1] = tau[2 * i + 1]; }
      xc::xc_point_pol<FAM>(x, thr, r2, s3, t2, e, vr, vs, vt);
      xc::xc_point_pol<FAM>(c, thr, r2, s3, t2, e, vr, vs, vt);
      exc[i] = e; vrho[2 * i] = vr[0]; vrho[2 * i + 1] = vr[1];
      if (FAM >= xc::FAM_GGA) { vsigma[3 * i] = vs[0]; vsigma[3 * i + 1] = vs[1]; vsigma[3 * i + 2] = vs[2]; }
      if (FAM >= xc::FAM_MGGA) { vtau[2 * i] = vt[0]; vtau[2 * i + 1] = vt[1]; }
    }
  }
}

extern "C" int xc_host_eval(int x, int c, int nspin, int n, double dens_thr, const double* rho, const double* sigma,
                            const double* tau, double* exc, double* vrho, double* vsigma, double* vtau) {
  const int fx = xc::xc_family(x), fc = xc::xc_family(c), fam = fx > fc ? fx : fc;
  if (fam == xc::FAM_LDA) run<xc::FAM_LDA>(x, c, nspin, n, dens_thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
  else if (fam == xc::FAM_GGA) run<xc::FAM_GGA>(x, c, nspin, n, dens_thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
  else if (fam == xc::FAM_MGGA) run<xc::FAM_MGGA>(x, c, nspin, n, dens_thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
  else return -3;
  return 0;
}
