// Per-grid-point XC evaluation with libxc's calling conventions.
//
// Mirrors what one xc_{lda,gga,mgga}_exc_vxc call returns for one point
// (/root/reference/src/sadatom/dftgrid.cpp:349-398; results of the exchange and the
// correlation functional are ADDED, :479-489):
//   unpolarised: rho, sigma=|grad rho|^2, tau (totals)      -> exc, vrho, vsigma, vtau
//   polarised:   (rho_a,rho_b), (s_aa,s_ab,s_bb), (tau_a,tau_b) -> exc, vrho[2], vsigma[3], vtau[2]
// exc is the energy per particle; the v* are partial derivatives of the energy per volume.
// Threshold handling follows libxc >= 5.1: points with total density below dens_threshold
// return zeros; inputs are floored (rho_s >= dens_thr, sigma_ss >= sigma_thr, tau_s >= tau_thr),
// sigma_ss <= 8 rho_s tau_s is enforced for meta-GGAs, |s_ab| <= (s_aa+s_bb)/2; a spin channel
// at the density floor contributes no exchange.
#pragma once
#include "xc_functionals.cuh"

namespace hfb {
namespace xc {

// Number of independent variables carried per family.
template <int FAM> struct NVar { enum { x = (FAM == FAM_LDA ? 1 : (FAM == FAM_GGA ? 2 : 3)) }; };

/// Unpolarised point.  FAM is the highest family among the two functionals.
/// Outputs are accumulated (+=) so the caller zeroes them first.
template <int FAM>
HFB_HD void xc_point_unpol(int func_id, const XcThresholds& thr, double rho, double sigma, double tau,
                           double& exc, double& vrho, double& vsigma, double& vtau) {
  if (func_id == XC_NONE) return;
  if (rho < thr.dens) return;
  constexpr int NV = NVar<FAM>::x;
  typedef Dual<NV> D;
  const int fam = xc_family(func_id);  // clamps depend on the functional's own family
  rho = fmax(rho, thr.dens);
  if (fam >= FAM_GGA) sigma = fmax(sigma, thr.sigma);
  if (fam >= FAM_MGGA) {
    tau = fmax(tau, thr.tau);
    sigma = fmin(sigma, 8.0 * rho * tau);
  }
  D n = D::var(rho, 0);
  D sg = (NV >= 2) ? D::var(sigma, NV >= 2 ? 1 : 0) : D(0.0);
  D tu = (NV >= 3) ? D::var(tau, NV >= 3 ? 2 : 0) : D(0.0);
  D e;
  if (xc_is_exchange(func_id)) {
    if (0.5 * rho <= thr.dens) return;
    e = 2.0 * exchange_spin<D>(func_id, 0.5 * n, 0.25 * sg, 0.5 * tu);
  } else {
    e = correlation<D>(func_id, n, D(0.0), sg, tu, false, thr.zeta);
  }
  exc += e.v / rho;
  vrho += e.d[0];
  if (NV >= 2) vsigma += e.d[NV >= 2 ? 1 : 0];
  if (NV >= 3) vtau += e.d[NV >= 3 ? 2 : 0];
}

/// Polarised point.
template <int FAM>
HFB_HD void xc_point_pol(int func_id, const XcThresholds& thr, const double* rho_in, const double* sigma_in,
                         const double* tau_in, double& exc, double* vrho, double* vsigma, double* vtau) {
  if (func_id == XC_NONE) return;
  if (rho_in[0] + rho_in[1] < thr.dens) return;
  double rho[2] = {fmax(rho_in[0], thr.dens), fmax(rho_in[1], thr.dens)};
  double sigma[3] = {0.0, 0.0, 0.0}, tau[2] = {0.0, 0.0};
  const int fam = xc_family(func_id);  // clamps depend on the functional's own family
  if (fam >= FAM_GGA) {
    sigma[0] = fmax(sigma_in[0], thr.sigma);
    sigma[2] = fmax(sigma_in[2], thr.sigma);
  }
  if (fam >= FAM_MGGA) {
    tau[0] = fmax(tau_in[0], thr.tau);
    tau[1] = fmax(tau_in[1], thr.tau);
    sigma[0] = fmin(sigma[0], 8.0 * rho[0] * tau[0]);
    sigma[2] = fmin(sigma[2], 8.0 * rho[1] * tau[1]);
  }
  if (fam >= FAM_GGA) {
    const double save = 0.5 * (sigma[0] + sigma[2]);
    sigma[1] = fmin(fmax(sigma_in[1], -save), save);
  }
  const double dens = rho[0] + rho[1];

  if (xc_is_exchange(func_id)) {
    constexpr int NV = NVar<FAM>::x;
    typedef Dual<NV> D;
#pragma unroll
    for (int s = 0; s < 2; s++) {
      if (rho[s] <= thr.dens) continue;  // screened channel
      D r = D::var(rho[s], 0);
      D sg = (NV >= 2) ? D::var(sigma[2 * s], NV >= 2 ? 1 : 0) : D(0.0);
      D tu = (NV >= 3) ? D::var(tau[s], NV >= 3 ? 2 : 0) : D(0.0);
      D e = exchange_spin<D>(func_id, r, sg, tu);
      exc += e.v / dens;
      vrho[s] += e.d[0];
      if (NV >= 2) vsigma[2 * s] += e.d[NV >= 2 ? 1 : 0];
      if (NV >= 3) vtau[s] += e.d[NV >= 3 ? 2 : 0];
    }
  } else {
    constexpr int NV = NVar<FAM>::x + 1;  // (n, zeta, sigma, tau)
    typedef Dual<NV> D;
    const double z = (rho[0] - rho[1]) / dens;
    D n = D::var(dens, 0);
    D zeta = D::var(z, 1);
    D sg = (NV >= 3) ? D::var(sigma[0] + 2.0 * sigma[1] + sigma[2], NV >= 3 ? 2 : 0) : D(0.0);
    D tu = (NV >= 4) ? D::var(tau[0] + tau[1], NV >= 4 ? 3 : 0) : D(0.0);
    D e = correlation<D>(func_id, n, zeta, sg, tu, true, thr.zeta);
    exc += e.v / dens;
    vrho[0] += e.d[0] + e.d[1] * (1.0 - z) / dens;
    vrho[1] += e.d[0] - e.d[1] * (1.0 + z) / dens;
    if (NV >= 3) {
      const double vs = e.d[NV >= 3 ? 2 : 0];
      vsigma[0] += vs;
      vsigma[1] += 2.0 * vs;
      vsigma[2] += vs;
    }
    if (NV >= 4) {
      const double vt = e.d[NV >= 4 ? 3 : 0];
      vtau[0] += vt;
      vtau[1] += vt;
    }
  }
}

}  // namespace xc
}  // namespace hfb
