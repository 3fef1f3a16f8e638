// C ABI, `atomic` / `diatomic` programs: DFTGrid::eval_Fxc on the angular x radial product grid.
//   helfem::atomic::dftgrid::DFTGrid::eval_Fxc     /root/reference/src/atomic/dftgrid.cpp:810-928
//   helfem::diatomic::dftgrid::DFTGrid::eval_Fxc   /root/reference/src/diatomic/dftgrid.cpp:769-856
// Kernels and the factorisation: grid2d.cu.
#include <cstring>
#include "../../include/helfem_b200.h"
#include "context.h"
#include "grid2d.h"
#include "xc_functionals.cuh"

using namespace hfb;

namespace {

bool valid_pair(int x_func, int c_func, int* fam) {
  const int fx = xc::xc_family(x_func), fc = xc::xc_family(c_func);
  if (x_func != 0 && (fx == 0 || !xc::xc_is_exchange(x_func))) return false;
  if (c_func != 0 && (fc == 0 || xc::xc_is_exchange(c_func))) return false;
  *fam = fx > fc ? fx : fc;
  return *fam != 0;
}

int eval_common(hfb200_ctx* c, const char* who, int x_func, int c_func, int nspin, const double* Pa, const double* Pb, int ldp,
                double dens_thr, double* Ha, double* Hb, int ldh, int beta, double* Exc, double* Nel, double* Ekin, const int* skip) {
  if (!Pa || !Ha || (nspin != 1 && nspin != 2) || (nspin == 2 && (!Pb || (beta && !Hb)))) {
    set_error("%s: invalid argument / empty density matrix", who);
    return HFB200_EINVAL;
  }
  if (!c->g2d || !c->g2d->ready) { set_error("%s: hfb200_set_grid has not been called", who); return HFB200_EINVAL; }
  int fam = 0;
  if (!valid_pair(x_func, c_func, &fam)) {
    set_error("functional pair (%d,%d) has no device implementation (no CPU fallback)", x_func, c_func);
    return HFB200_EFUNC;
  }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  double sc[3] = {0, 0, 0};
  const int rc = c->g2d->eval_fxc(fam, x_func, c_func, nspin, Pa, Pb, ldp, dens_thr, Ha, Hb, ldh, beta, sc, c->stream, skip);
  if (rc) return rc;
  if (Exc) *Exc = sc[0];
  if (Nel) *Nel = sc[1];
  if (Ekin) *Ekin = sc[2];
  return HFB200_OK;
}

}  // namespace

extern "C" {

int hfb200_set_grid(hfb200_ctx* c, int nq, const double* r, const double* wrad, const double* bf, const double* df,
                    int nangpts, const double* cth, const double* wang, const double* ylm, const double* dth_ylm) {
  if (!c || (c->kind != CTX_ATOMIC && c->kind != CTX_DIATOMIC) || nq < 1 || !r || !wrad || !bf || !df || nangpts < 1 || !cth ||
      !wang || !ylm || !dth_ylm) {
    set_error("hfb200_set_grid: invalid argument");
    return HFB200_EINVAL;
  }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  if (!c->g2d) c->g2d = new Grid2D();
  return c->g2d->setup(c->kind == CTX_ATOMIC ? 0 : 1, c->Rhalf, c->lval, c->mval, c->nel, c->first, c->last, nq, r, wrad, bf, df,
                       nangpts, cth, wang, ylm, dth_ylm);
}

int hfb200_atomic_eval_fxc(hfb200_ctx* c, int x_func, int c_func, int nspin, const double* Pa, const double* Pb, int ldp,
                           double dens_thr, double* Ha, double* Hb, int ldh, int beta, double* Exc, double* Nel, double* Ekin) {
  if (!c || c->kind != CTX_ATOMIC) { set_error("hfb200_atomic_eval_fxc: invalid argument"); return HFB200_EINVAL; }
  const int ntot = (int)c->lval.size() * c->nrad;
  if (ldp < ntot || ldh < ntot) { set_error("Density matrix does not match basis set!"); return HFB200_EINVAL; }
  return eval_common(c, "hfb200_atomic_eval_fxc", x_func, c_func, nspin, Pa, Pb, ldp, dens_thr, Ha, Hb, ldh, beta, Exc, Nel, Ekin, nullptr);
}

int hfb200_diatomic_eval_fxc(hfb200_ctx* c, int x_func, int c_func, int nspin, const double* Pa, const double* Pb, int ldp,
                             double dens_thr, double* Ha, double* Hb, int ldh, int beta, double* Exc, double* Nel, double* Ekin) {
  if (!c || c->kind != CTX_DIATOMIC) { set_error("hfb200_diatomic_eval_fxc: invalid argument"); return HFB200_EINVAL; }
  // pure <-> dummy index map (diatomic/basis.cpp:482-499): the first radial function of every m != 0 angular function is absent
  // from the caller's matrices.  expand_boundaries (diatomic/basis.cpp:1754-1773) and remove_boundaries (dftgrid.cpp:809,854-855)
  // happen while the angular blocks are packed / unpacked (grid2d.cu).
  std::vector<int> skip(c->mval.size());
  int npure = 0;
  for (size_t i = 0; i < c->mval.size(); i++) { skip[i] = c->mval[i] != 0 ? 1 : 0; npure += c->nrad - skip[i]; }
  if (ldp < npure || ldh < npure) { set_error("Matrix does not have expected size!"); return HFB200_EINVAL; }
  if (!Pa || !Ha || (nspin == 2 && (!Pb || (beta && !Hb)))) { set_error("hfb200_diatomic_eval_fxc: empty density matrix"); return HFB200_EINVAL; }
  return eval_common(c, "hfb200_diatomic_eval_fxc", x_func, c_func, nspin, Pa, Pb, ldp, dens_thr, Ha, Hb, ldh, beta, Exc, Nel, Ekin, skip.data());
}

int hfb200_grid_launches(hfb200_ctx* c) { return (c && c->g2d) ? c->g2d->launches : 0; }
int hfb200_grid_last_call(hfb200_ctx* c, double* out3) {
  if (!c || !c->g2d || !out3) { set_error("hfb200_grid_last_call: invalid argument"); return HFB200_EINVAL; }
  out3[0] = (double)c->g2d->device_ms; out3[1] = c->g2d->h2d_bytes; out3[2] = c->g2d->d2h_bytes;
  return HFB200_OK;
}

}  // extern "C"
