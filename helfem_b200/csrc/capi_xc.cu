// C ABI: pointwise XC on the device (validation entry point, include/helfem_b200.h).
#include "../../include/helfem_b200.h"
#include "context.h"
#include "xc_point.cuh"

using namespace hfb;

namespace {

template <int FAM, int NSPIN>
__global__ void xc_eval_kernel(int x_func, int c_func, int n, double dens_thr, const double* rho, const double* sigma,
                               const double* tau, double* exc, double* vrho, double* vsigma, double* vtau) {
  const xc::XcThresholds thr = xc::make_thresholds(dens_thr);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (NSPIN == 1) {
      double e = 0.0, vr = 0.0, vs = 0.0, vt = 0.0;
      const double sg = (FAM >= xc::FAM_GGA) ? sigma[i] : 0.0;
      const double ta = (FAM >= xc::FAM_MGGA) ? tau[i] : 0.0;
      xc::xc_point_unpol<FAM>(x_func, thr, rho[i], sg, ta, e, vr, vs, vt);
      xc::xc_point_unpol<FAM>(c_func, thr, rho[i], sg, ta, e, vr, vs, vt);
      exc[i] = e;
      vrho[i] = vr;
      if (FAM >= xc::FAM_GGA) vsigma[i] = vs;
      if (FAM >= xc::FAM_MGGA) vtau[i] = vt;
    } else {
      double e = 0.0, vr[2] = {0.0, 0.0}, vs[3] = {0.0, 0.0, 0.0}, vt[2] = {0.0, 0.0};
      double r2[2] = {rho[2 * i], rho[2 * i + 1]};
      double s3[3] = {0.0, 0.0, 0.0}, t2[2] = {0.0, 0.0};
      if (FAM >= xc::FAM_GGA) { s3[0] = sigma[3 * i]; s3[1] = sigma[3 * i + 1]; s3[2] = sigma[3 * i + 2]; }
      if (FAM >= xc::FAM_MGGA) { t2[0] = tau[2 * i]; t2[1] = tau[2 * i + 1]; }
      xc::xc_point_pol<FAM>(x_func, thr, r2, s3, t2, e, vr, vs, vt);
      xc::xc_point_pol<FAM>(c_func, thr, r2, s3, t2, e, vr, vs, vt);
      exc[i] = e;
      vrho[2 * i] = vr[0]; vrho[2 * i + 1] = vr[1];
      if (FAM >= xc::FAM_GGA) { vsigma[3 * i] = vs[0]; vsigma[3 * i + 1] = vs[1]; vsigma[3 * i + 2] = vs[2]; }
      if (FAM >= xc::FAM_MGGA) { vtau[2 * i] = vt[0]; vtau[2 * i + 1] = vt[1]; }
    }
  }
}

}  // namespace

extern "C" int hfb200_xc_eval(int device, int x_func, int c_func, int nspin, int n, double dens_thr, const double* rho,
                              const double* sigma, const double* tau, double* exc, double* vrho, double* vsigma,
                              double* vtau) {
  const int fx = xc::xc_family(x_func), fc = xc::xc_family(c_func);
  if ((x_func && (!fx || !xc::xc_is_exchange(x_func))) || (c_func && (!fc || xc::xc_is_exchange(c_func))) || !(fx | fc)) {
    set_error("functional pair (%d,%d) has no device implementation (no CPU fallback)", x_func, c_func);
    return HFB200_EFUNC;
  }
  const int fam = fx > fc ? fx : fc;
  if (n < 1 || (nspin != 1 && nspin != 2) || !rho || !exc || !vrho || (fam >= 2 && (!sigma || !vsigma)) ||
      (fam >= 4 && (!tau || !vtau))) {
    set_error("hfb200_xc_eval: invalid argument");
    return HFB200_EINVAL;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    set_error("no CUDA device available: helfem_b200 has no CPU fallback");
    return HFB200_ENODEVICE;
  }
  HFB_CUDA_OK(cudaSetDevice(device));
  const size_t nr = (size_t)n * nspin, ns = (size_t)n * (nspin == 1 ? 1 : 3);
  double *d = nullptr;
  const size_t total = 2 * nr + 2 * ns + 2 * nr + n;  // rho,vrho | sigma,vsigma | tau,vtau | exc
  HFB_CUDA_OK(cudaMalloc(&d, total * sizeof(double)));
  double *dRho = d, *dVrho = d + nr, *dSig = d + 2 * nr, *dVsig = dSig + ns, *dTau = dVsig + ns, *dVtau = dTau + nr, *dExc = dVtau + nr;
  cudaMemset(d, 0, total * sizeof(double));
  cudaMemcpy(dRho, rho, nr * sizeof(double), cudaMemcpyHostToDevice);
  if (fam >= 2) cudaMemcpy(dSig, sigma, ns * sizeof(double), cudaMemcpyHostToDevice);
  if (fam >= 4) cudaMemcpy(dTau, tau, nr * sizeof(double), cudaMemcpyHostToDevice);
  const int blocks = (n + 127) / 128 < 148 * 8 ? (n + 127) / 128 : 148 * 8;
#define HFB_XC_LAUNCH(F, S) xc_eval_kernel<F, S><<<blocks, 128>>>(x_func, c_func, n, dens_thr, dRho, dSig, dTau, dExc, dVrho, dVsig, dVtau)
  if (nspin == 1) {
    if (fam == 1) HFB_XC_LAUNCH(xc::FAM_LDA, 1); else if (fam == 2) HFB_XC_LAUNCH(xc::FAM_GGA, 1); else HFB_XC_LAUNCH(xc::FAM_MGGA, 1);
  } else {
    if (fam == 1) HFB_XC_LAUNCH(xc::FAM_LDA, 2); else if (fam == 2) HFB_XC_LAUNCH(xc::FAM_GGA, 2); else HFB_XC_LAUNCH(xc::FAM_MGGA, 2);
  }
#undef HFB_XC_LAUNCH
  cudaError_t e = cudaDeviceSynchronize();
  if (e == cudaSuccess) {
    cudaMemcpy(exc, dExc, n * sizeof(double), cudaMemcpyDeviceToHost);
    cudaMemcpy(vrho, dVrho, nr * sizeof(double), cudaMemcpyDeviceToHost);
    if (fam >= 2) cudaMemcpy(vsigma, dVsig, ns * sizeof(double), cudaMemcpyDeviceToHost);
    if (fam >= 4) cudaMemcpy(vtau, dVtau, nr * sizeof(double), cudaMemcpyDeviceToHost);
    e = cudaGetLastError();
  }
  cudaFree(d);
  if (e != cudaSuccess) { set_error("hfb200_xc_eval: CUDA error %s", cudaGetErrorString(e)); return HFB200_ECUDA; }
  return HFB200_OK;
}
