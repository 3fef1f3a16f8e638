// Host-side mirror of the reference's `atomic` operator class (see atomic.h).
#include "atomic.h"

#include <algorithm>
#include <string>
#include "../../include/helfem_b200.h"

namespace hfb {
namespace atomic {

static void throw_status(int rc) {
  if (rc == HFB200_OK) return;
  const std::string msg = std::string(hfb200_last_error()) + "\n";
  if (rc == HFB200_EINVAL || rc == HFB200_ENOTEI) throw std::logic_error(msg);
  throw std::runtime_error(msg);
}

TwoDBasis::TwoDBasis(int Z_, const fem::PolyBasis& poly, bool zeroder, int n_quad, const Vec& bval, int taylor_order,
                     const std::vector<int>& lval_, const std::vector<int>& mval_, int device_)
    : Z(Z_), device(device_), lval(lval_), mval(mval_) {
  // atomic/TwoDBasis.cpp:48-53
  fem::FEBasis fe(poly, bval, true, false, true, zeroder);
  radial = fem::RadialBasis(fe, n_quad, taylor_order);
}

TwoDBasis::~TwoDBasis() {
  if (ctx) hfb200_destroy(ctx);
}

// which: 0 overlap, 1 kinetic, 2 kinetic_l, 3 r^-1
Mat TwoDBasis::radial_assemble(int which) const {
  const size_t Nr = radial.Nbf();
  Mat O(Nr, Nr);
  for (size_t iel = 0; iel < radial.Nel(); iel++) {
    size_t f, l;
    radial.get_idx(iel, f, l);
    Mat sub = which == 0 ? radial.radial_integral(0, iel)
            : which == 1 ? radial.kinetic(iel)
            : which == 2 ? radial.kinetic_l(iel)
                         : radial.radial_integral(-1, iel);
    for (size_t j = 0; j < sub.nc; j++)
      for (size_t i = 0; i < sub.nr; i++) O(f + i, f + j) += sub(i, j);
  }
  return O;
}

Mat TwoDBasis::overlap() const {
  const size_t Nr = radial.Nbf();
  Mat S = radial_assemble(0), O(Nbf(), Nbf());
  for (size_t a = 0; a < lval.size(); a++)
    for (size_t j = 0; j < Nr; j++)
      for (size_t i = 0; i < Nr; i++) O(a * Nr + i, a * Nr + j) = S(i, j);
  return O;
}

Mat TwoDBasis::kinetic() const {
  const size_t Nr = radial.Nbf();
  Mat T = radial_assemble(1), Tl = radial_assemble(2), O(Nbf(), Nbf());
  for (size_t a = 0; a < lval.size(); a++) {
    const double ll = lval[a] * (lval[a] + 1);
    for (size_t j = 0; j < Nr; j++)
      for (size_t i = 0; i < Nr; i++) O(a * Nr + i, a * Nr + j) = T(i, j) + ll * Tl(i, j);
  }
  return O;
}

Mat TwoDBasis::nuclear() const {
  const size_t Nr = radial.Nbf();
  Mat V = radial_assemble(3), O(Nbf(), Nbf());
  for (size_t a = 0; a < lval.size(); a++)
    for (size_t j = 0; j < Nr; j++)
      for (size_t i = 0; i < Nr; i++) O(a * Nr + i, a * Nr + j) = -Z * V(i, j);
  return O;
}

// atomic/TwoDBasis.cpp:666-739
void TwoDBasis::compute_tei(bool exchange) {
  const size_t N_L = 2 * (size_t)*std::max_element(lval.begin(), lval.end()) + 1;
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

void TwoDBasis::set_shard(int rank, int nranks) {
  if (ctx) throw std::logic_error("set_shard must precede the first use of the device context\n");
  shard_rank = rank;
  shard_n = nranks;
}

hfb200_ctx* TwoDBasis::context() const {
  if (!ctx) {
    const size_t Nel = radial.Nel();
    std::vector<int> first(Nel), last(Nel);
    for (size_t iel = 0; iel < Nel; iel++) {
      size_t f, l;
      radial.get_idx(iel, f, l);
      first[iel] = (int)f;
      last[iel] = (int)l;
    }
    throw_status(hfb200_atomic_create(&ctx, device, (int)lval.size(), lval.data(), mval.data(), (int)Nel, first.data(), last.data()));
    throw_status(hfb200_set_shard(ctx, shard_rank, shard_n));
  }
  if (!tei_uploaded && !prim_tei.empty()) {
    std::vector<const double*> a, b, t, k;
    for (size_t i = 0; i < prim_tei.size(); i++) {
      a.push_back(disjoint_L[i].ptr());
      b.push_back(disjoint_m1L[i].ptr());
      t.push_back(prim_tei[i].ptr());
      if (!prim_ktei.empty()) k.push_back(prim_ktei[i].ptr());
    }
    throw_status(hfb200_atomic_set_tei(ctx, nL_stored, a.data(), b.data(), t.data(), k.empty() ? nullptr : k.data()));
    tei_uploaded = true;
  }
  return ctx;
}

Mat TwoDBasis::coulomb(const Mat& P) const {
  if (prim_tei.empty()) throw std::logic_error("Primitive teis have not been computed!\n");
  if (P.nr != Nbf() || P.nc != Nbf()) throw std::logic_error("Matrix does not have expected size!\n");
  Mat J(P.nr, P.nc);
  throw_status(hfb200_atomic_coulomb(context(), P.ptr(), (int)P.nr, J.ptr(), (int)J.nr));
  return J;
}

Mat TwoDBasis::exchange(const Mat& P) const {
  if (prim_ktei.empty()) throw std::logic_error("Primitive teis have not been computed!\n");
  if (P.nr != Nbf() || P.nc != Nbf()) throw std::logic_error("Matrix does not have expected size!\n");
  Mat K(P.nr, P.nc);
  throw_status(hfb200_atomic_exchange(context(), P.ptr(), (int)P.nr, K.ptr(), (int)K.nr));
  return K;
}


// Raw-pointer forms of the two operators (n x n column-major, n = Nbf()): what the flat C entry points of
// hostapi.cpp call, so that a caller's buffers go straight to the C ABI without an intermediate Mat copy.
void TwoDBasis::coulomb(const double* P, double* J) const {
  if (prim_tei.empty()) throw std::logic_error("Primitive teis have not been computed!\n");
  throw_status(hfb200_atomic_coulomb(context(), P, (int)Nbf(), J, (int)Nbf()));
}
void TwoDBasis::exchange(const double* P, double* K) const {
  if (prim_ktei.empty()) throw std::logic_error("Primitive teis have not been computed!\n");
  throw_status(hfb200_atomic_exchange(context(), P, (int)Nbf(), K, (int)Nbf()));
}

}  // namespace atomic
}  // namespace hfb
