// C ABI, `atomic` program: Gaunt-coupled coulomb / exchange on the (l,m) x radial product basis.
//   helfem::atomic::basis::TwoDBasis::coulomb   /root/reference/src/atomic/TwoDBasis.cpp:817-955
//   helfem::atomic::basis::TwoDBasis::exchange  /root/reference/src/atomic/TwoDBasis.cpp:957-1140
// The angular algebra (which (L,M) channels exist, which angular pairs couple to them with which
// Gaunt coefficient) is a constant of the basis and is flattened once into CSR lists; the radial
// work runs through the shared channel engine (chan_engine.h).
#include <cmath>
#include <cstring>
#include "../../include/helfem_b200.h"
#include "../host/angular.h"
#include "chan_engine.h"
#include "context.h"
#include "../host/famshard.h"

using namespace hfb;

extern "C" {

int hfb200_atomic_create(hfb200_ctx** out, int device, int nang, const int* lval, const int* mval, int nel,
                         const int* first, const int* last) {
  if (!out || nang < 1 || !lval || !mval || nel < 1 || !first || !last) {
    set_error("hfb200_atomic_create: invalid argument");
    return HFB200_EINVAL;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    set_error("no CUDA device available: helfem_b200 has no CPU fallback");
    return HFB200_ENODEVICE;
  }
  HFB_CUDA_OK(cudaSetDevice(device));
  hfb200_ctx* c = new hfb200_ctx();
  c->kind = CTX_ATOMIC;
  c->device = device;
  c->nel = nel;
  c->first.assign(first, first + nel);
  c->last.assign(last, last + nel);
  c->nrad = last[nel - 1] + 1;
  c->lval.assign(lval, lval + nang);
  c->mval.assign(mval, mval + nang);
  for (int e = 0; e < nel; e++) c->np_max = std::max(c->np_max, last[e] - first[e] + 1);
  {
    const cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {  // do not leak the half-built context
      set_error("cudaStreamCreateWithFlags failed: %s", cudaGetErrorString(e));
      hfb200_destroy(c);
      return e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? HFB200_ENODEVICE : HFB200_ECUDA;
    }
  }
  *out = c;
  return HFB200_OK;
}

/* Stored integrals of TwoDBasis::compute_tei (atomic/TwoDBasis.cpp:666-739), L = 0..nL-1 with
 * nL = 2 max(lval) + 1; arrays of nL*nel pointers, entry [L*nel + iel] (the reference's in-element
 * index Nel*Nel*L + iel*Nel + iel).  prim_ktei may be NULL if exchange() is never called. */
int hfb200_atomic_set_tei(hfb200_ctx* c, int nL, const double* const* disjoint_L, const double* const* disjoint_m1L,
                          const double* const* prim_tei, const double* const* prim_ktei) {
  if (!c || c->kind != CTX_ATOMIC || !disjoint_L || !disjoint_m1L || !prim_tei) {
    set_error("hfb200_atomic_set_tei: invalid argument");
    return HFB200_EINVAL;
  }
  const int nang = (int)c->lval.size();
  int lmax = 0, mmin = 0, mmx = 0;
  for (int a = 0; a < nang; a++) {
    lmax = std::max(lmax, c->lval[a]);
    mmin = std::min(mmin, c->mval[a]);
    mmx = std::max(mmx, c->mval[a]);
  }
  if (nL < 2 * lmax + 1) { set_error("hfb200_atomic_set_tei: need L = 0..%d", 2 * lmax); return HFB200_EINVAL; }
  nL = 2 * lmax + 1;
  const int Mmax = mmx - mmin;
  HFB_CUDA_OK(cudaSetDevice(c->device));
  ChanSpec S;
  S.nel = c->nel; S.nang = nang; S.stack = 1; S.nr = nL;
  S.first = c->first; S.last = c->last;
  // tiling hint of the exchange contractions: functions with equal m AND equal parity of l couple to the same partners
  // (a Gaunt or cos^2-modified Gaunt coefficient <l_j m_j| L M |l_i m_i> vanishes unless l_j + L + l_i is even), so a
  // tile of such rows / items has a dense coefficient table instead of a checkerboard
  S.ang_group.resize(nang);
  for (int a = 0; a < nang; a++) S.ang_group[a] = 2 * c->mval[a] + (c->lval[a] & 1);
  S.ang_m = c->mval;
  S.fam_owned.resize(nL);
  c->fam_owner = family_owner(nL, c->shard_n);
  const std::vector<int> own = c->fam_owner;
  for (int L = 0; L < nL; L++) S.fam_owned[L] = (own[L] == c->shard_rank);  // couplings of ALL families are listed below
  S.fac.resize(nL);
  for (int L = 0; L < nL; L++) S.fac[L] = 4.0 * M_PI / (2 * L + 1);
  // channels (L, M), sorted by L (atomic/TwoDBasis.cpp:836-842)
  std::vector<int> chan_of((size_t)nL * (2 * Mmax + 1), -1);
  for (int L = 0; L < nL; L++)
    for (int M = -std::min(L, Mmax); M <= std::min(L, Mmax); M++) {
      chan_of[(size_t)L * (2 * Mmax + 1) + M + Mmax] = (int)S.chan_r.size();
      S.chan_r.push_back(L);
    }
  // ket couplings (:845-864); the bra side (:932-952) uses the list of the transposed pair
  S.pair_coupl.assign((size_t)nang * nang, {});
  angular::gaunt_coefficient(2 * lmax, 0, 2 * lmax, 0, 0, 0);  // size the shared Gaunt table before the parallel loop
#pragma omp parallel for schedule(dynamic)
  for (int k = 0; k < nang; k++)
    for (int l = 0; l < nang; l++) {
      const int lk = c->lval[k], mk = c->mval[k], ll = c->lval[l], ml = c->mval[l], M = mk - ml;
      for (int L = std::max(std::abs(lk - ll), std::abs(M)); L <= lk + ll; L++) {
        const double g = angular::gaunt_coefficient(lk, mk, L, M, ll, ml);
        if (g == 0.0) continue;
        S.pair_coupl[(size_t)k * nang + l].push_back({chan_of[(size_t)L * (2 * Mmax + 1) + M + Mmax], {g, 0.0}});
      }
    }
  const int nel = c->nel;
  const int sr = c->shard_rank;
  typedef const double* cptr;
  S.tei = [=](int r, int e, int, int, double* sg) { *sg = 1.0; return own[r] == sr ? prim_tei[r * nel + e] : (cptr) nullptr; };
  if (prim_ktei) S.ktei = [=](int r, int e, int, int, double* sg) { *sg = 1.0; return own[r] == sr ? prim_ktei[r * nel + e] : (cptr) nullptr; };
  S.small_f = [=](int r, int e, int, double* sg) { *sg = 1.0; return own[r] == sr ? disjoint_L[r * nel + e] : (cptr) nullptr; };
  S.big_f = [=](int r, int e, int, double* sg) { *sg = 1.0; return own[r] == sr ? disjoint_m1L[r * nel + e] : (cptr) nullptr; };
  if (c->chan) { c->chan->release(); delete c->chan; }
  c->chan = new ChanEngine();
  c->chan->comm = c->comm;
  c->chan->comm_rank = c->shard_rank;
  c->chan->comm_size = c->comm ? c->shard_n : 1;
  const int rc = c->chan->build(S, prim_ktei != nullptr);
  if (rc) return rc;
  c->nL = nL;
  c->have_ktei = prim_ktei != nullptr;
  return HFB200_OK;
}

int hfb200_atomic_coulomb(hfb200_ctx* c, const double* P, int ldp, double* J, int ldj) {
  if (!c || c->kind != CTX_ATOMIC || !P || !J) { set_error("hfb200_atomic_coulomb: invalid argument"); return HFB200_EINVAL; }
  if (!c->chan) { set_error("Primitive teis have not been computed!"); return HFB200_ENOTEI; }
  const int ntot = (int)c->lval.size() * c->nrad;
  if (ldp < ntot || ldj < ntot) { set_error("Density matrix does not match basis set!"); return HFB200_EINVAL; }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  return c->chan->coulomb_dense(P, ldp, J, ldj, c->stream);
}

int hfb200_atomic_exchange(hfb200_ctx* c, const double* P, int ldp, double* K, int ldk) {
  if (!c || c->kind != CTX_ATOMIC || !P || !K) { set_error("hfb200_atomic_exchange: invalid argument"); return HFB200_EINVAL; }
  if (!c->chan || !c->have_ktei) { set_error("Primitive teis have not been computed!"); return HFB200_ENOTEI; }
  const int ntot = (int)c->lval.size() * c->nrad;
  if (ldp < ntot || ldk < ntot) { set_error("Density matrix does not match basis set!"); return HFB200_EINVAL; }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  return c->chan->exchange_dense(P, ldp, K, ldk, c->stream);
}

}  // extern "C"
