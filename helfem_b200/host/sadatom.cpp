// Host-side mirror of the reference's sadatom operator classes (see sadatom.h).
#include "sadatom.h"

#include <cmath>
#include <string>
#include "../../include/helfem_b200.h"

namespace hfb {
namespace sadatom {

static void throw_status(int rc) {
  if (rc == HFB200_OK) return;
  const std::string msg = std::string(hfb200_last_error()) + "\n";
  // the reference throws std::logic_error for missing integrals / size mismatches
  // (sadatom/basis.cpp:297-298,361-376) and std::runtime_error from the grid code
  if (rc == HFB200_EINVAL || rc == HFB200_ENOTEI) throw std::logic_error(msg);
  throw std::runtime_error(msg);
}

TwoDBasis::TwoDBasis(int Z_, const fem::PolyBasis& poly, bool zeroder, int n_quad, const Vec& bval, int taylor_order,
                     int lmax_, int device_)
    : Z(Z_), lmax(lmax_), device(device_) {
  // sadatom/basis.cpp:46-50
  fem::FEBasis fe(poly, bval, true, false, true, zeroder);
  radial = fem::RadialBasis(fe, n_quad, taylor_order);
}

TwoDBasis::~TwoDBasis() {
  if (ctx) hfb200_destroy(ctx);
}

// sadatom/basis.cpp:73-90
Mat TwoDBasis::radial_integral(int Rexp) const {
  const size_t Nrad = radial.Nbf();
  Mat O(Nrad, Nrad);
  for (size_t iel = 0; iel < radial.Nel(); iel++) {
    size_t f, l;
    radial.get_idx(iel, f, l);
    Mat sub = radial.radial_integral(Rexp, iel);
    for (size_t j = 0; j < sub.nc; j++)
      for (size_t i = 0; i < sub.nr; i++) O(f + i, f + j) += sub(i, j);
  }
  return O;
}

static Mat assemble(const fem::RadialBasis& radial, Mat (fem::RadialBasis::*fn)(size_t) const) {
  const size_t Nrad = radial.Nbf();
  Mat O(Nrad, Nrad);
  for (size_t iel = 0; iel < radial.Nel(); iel++) {
    size_t f, l;
    radial.get_idx(iel, f, l);
    Mat sub = (radial.*fn)(iel);
    for (size_t j = 0; j < sub.nc; j++)
      for (size_t i = 0; i < sub.nr; i++) O(f + i, f + j) += sub(i, j);
  }
  return O;
}

Mat TwoDBasis::kinetic() const { return assemble(radial, &fem::RadialBasis::kinetic); }
Mat TwoDBasis::kinetic_l() const { return assemble(radial, &fem::RadialBasis::kinetic_l); }

// sadatom/basis.cpp:131-150 (point nucleus)
Mat TwoDBasis::nuclear() const {
  Mat V = radial_integral(-1);
  for (auto& v : V.d) v *= -Z;
  return V;
}

std::vector<size_t> TwoDBasis::bf_list(size_t iel) const {
  size_t f, l;
  radial.get_idx(iel, f, l);
  std::vector<size_t> idx(l - f + 1);
  for (size_t j = 0; j < idx.size(); j++) idx[j] = f + j;
  return idx;
}

// sadatom/basis.cpp:181-218
void TwoDBasis::compute_tei(int nL, bool exchange) {
  const size_t N_L = (nL > 0) ? (size_t)nL : (size_t)(2 * lmax + 1);
  const size_t Nel = radial.Nel();
  disjoint_L.assign(Nel * N_L, Mat());
  disjoint_m1L.assign(Nel * N_L, Mat());
  prim_tei.assign(Nel * N_L, Mat());
  prim_ktei.assign(exchange ? Nel * N_L : 0, Mat());
#ifdef _OPENMP
#pragma omp parallel for collapse(2) schedule(dynamic)
#endif
  for (size_t L = 0; L < N_L; L++)
    for (size_t iel = 0; iel < Nel; iel++) {
      disjoint_L[L * Nel + iel] = radial.radial_integral((int)L, iel);
      disjoint_m1L[L * Nel + iel] = radial.radial_integral(-(int)L - 1, iel);
      prim_tei[L * Nel + iel] = radial.twoe_integral((int)L, iel);
      if (exchange) {
        const size_t Ni = radial.Nprim(iel);
        prim_ktei[L * Nel + iel] = fem::exchange_tei(prim_tei[L * Nel + iel], Ni, Ni, Ni, Ni);
      }
    }
  nL_stored = (int)N_L;
  tei_uploaded = false;
}

hfb200_ctx* TwoDBasis::context() const {
  if (!ctx) {
    const size_t Nel = radial.Nel();
    const int nq = radial.get_nquad();
    std::vector<int> first(Nel), last(Nel);
    Vec r, w, bf, df, lf;
    for (size_t iel = 0; iel < Nel; iel++) {
      size_t f, l;
      radial.get_idx(iel, f, l);
      first[iel] = (int)f;
      last[iel] = (int)l;
      Vec ri = radial.get_r(iel), wi = radial.get_wrad(iel);
      r.insert(r.end(), ri.begin(), ri.end());
      w.insert(w.end(), wi.begin(), wi.end());
      Mat b = radial.get_bf(iel), d = radial.get_df(iel), l2 = radial.get_lf(iel);
      bf.insert(bf.end(), b.d.begin(), b.d.end());
      df.insert(df.end(), d.d.begin(), d.d.end());
      lf.insert(lf.end(), l2.d.begin(), l2.d.end());
    }
    throw_status(hfb200_sadatom_create(&ctx, device, lmax, (int)Nel, nq, first.data(), last.data(), r.data(), w.data(),
                                       bf.data(), df.data(), lf.data()));
  }
  if (!tei_uploaded && !prim_tei.empty()) {
    std::vector<const double*> a, b, t, k;
    for (size_t i = 0; i < prim_tei.size(); i++) {
      a.push_back(disjoint_L[i].ptr());
      b.push_back(disjoint_m1L[i].ptr());
      t.push_back(prim_tei[i].ptr());
      if (!prim_ktei.empty()) k.push_back(prim_ktei[i].ptr());
    }
    throw_status(hfb200_sadatom_set_tei(ctx, nL_stored, a.data(), b.data(), t.data(), k.empty() ? nullptr : k.data()));
    tei_uploaded = true;
  }
  return ctx;
}

Mat TwoDBasis::coulomb(const Mat& P) const {
  if (prim_tei.empty()) throw std::logic_error("Primitive teis have not been computed!\n");
  if (P.nr != Nbf() || P.nc != Nbf()) throw std::logic_error("Density matrix does not match basis set!\n");
  Mat J(P.nr, P.nc);
  throw_status(hfb200_sadatom_coulomb(context(), P.ptr(), (int)P.nr, J.ptr(), (int)J.nr));
  return J;
}

Cube TwoDBasis::exchange(const Cube& P) const {
  if (prim_ktei.empty()) throw std::logic_error("Primitive teis have not been computed!\n");
  if (P.ns != (size_t)lmax + 1) throw std::logic_error("Density matrix am does not match basis set!\n");
  if (P.nr != Nbf() || P.nc != Nbf()) throw std::logic_error("Density matrix does not match basis set!\n");
  Cube K(P.nr, P.nc, P.ns);
  throw_status(hfb200_sadatom_exchange(context(), P.ptr(), (int)P.nr, (int)P.ns, K.ptr(), (int)K.nr));
  return K;
}

void DFTGrid::eval_Fxc(int x_func, const Vec& x_pars, int c_func, const Vec& c_pars, const Cube& P, Cube& H, double& Exc,
                       double& Nel, double thr) {
  if (!x_pars.empty() || !c_pars.empty())
    throw std::runtime_error("External functional parameters are not supported by the device functionals.\n");
  if (P.nr != basp->Nbf() || P.nc != basp->Nbf()) throw std::logic_error("Density matrix does not match basis set!\n");
  H.zeros(P.nr, P.nr, P.ns);
  throw_status(hfb200_sadatom_eval_fxc(basp->context(), x_func, c_func, 1, P.ptr(), nullptr, (int)P.nr, (int)P.ns, thr,
                                       H.ptr(), nullptr, (int)H.nr, 0, &Exc, &Nel));
}

void DFTGrid::eval_Fxc(int x_func, const Vec& x_pars, int c_func, const Vec& c_pars, const Cube& Pa, const Cube& Pb, Cube& Ha,
                       Cube& Hb, double& Exc, double& Nel, bool beta, double thr) {
  if (!x_pars.empty() || !c_pars.empty())
    throw std::runtime_error("External functional parameters are not supported by the device functionals.\n");
  if (!Pa.d.size() || !Pb.d.size()) throw std::runtime_error("Error - density matrix is empty!\n");
  if (Pa.nr != basp->Nbf() || Pb.nr != basp->Nbf() || Pa.ns != Pb.ns)
    throw std::logic_error("Density matrix does not match basis set!\n");
  Ha.zeros(Pa.nr, Pa.nr, Pa.ns);
  Hb.zeros(Pb.nr, Pb.nr, Pb.ns);
  throw_status(hfb200_sadatom_eval_fxc(basp->context(), x_func, c_func, 2, Pa.ptr(), Pb.ptr(), (int)Pa.nr, (int)Pa.ns, thr,
                                       Ha.ptr(), Hb.ptr(), (int)Ha.nr, beta ? 1 : 0, &Exc, &Nel));
}

}  // namespace sadatom
}  // namespace hfb
