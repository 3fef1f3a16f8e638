// C ABI, `diatomic` program (prolate spheroidal coordinates): Gaunt-coupled coulomb / exchange.
//   helfem::diatomic::basis::TwoDBasis::coulomb   /root/reference/src/diatomic/basis.cpp:1359-1530
//   helfem::diatomic::basis::TwoDBasis::exchange  /root/reference/src/diatomic/basis.cpp:1532-1733
// The Neumann expansion couples every angular pair through two channels per (L,M): the plain Gaunt
// coefficient ("2", from cosh^2 mu) and the cos^2-modified one ("0").  Both, with the four integral
// families 00/02/20/22 and their signs, are mapped onto the stacked form of the shared channel
// engine (chan_kernels.cuh header comment); matrices cross the ABI in the "pure" basis and the
// dummy first radial function of every m != 0 shell (basis.cpp:482-499) is handled here.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include "../../include/helfem_b200.h"
#include "../host/angular.h"
#include "chan_engine.h"
#include "context.h"
#include "../host/famshard.h"

using namespace hfb;

namespace {
// Gaunt::coeff guards (general/gaunt.cpp:113-115)
inline double gcoeff(int L, int M, int l, int m, int lp, int mp) {
  if (std::abs(M) > L || std::abs(m) > l || std::abs(mp) > lp) return 0.0;
  return angular::gaunt_coefficient(L, M, l, m, lp, mp);
}
// Gaunt::mod_coeff (general/gaunt.cpp:167-180)
double mod_coeff(int lj, int mj, int L, int M, int li, int mi) {
  static const double const0(2.0 / 3.0 * sqrt(M_PI));
  static const double const2(4.0 / 15.0 * sqrt(5.0 * M_PI));
  const double cpl0 = gcoeff(L, M, 0, 0, L, M) * gcoeff(lj, mj, li, mi, L, M);
  double cpl2 = 0.0;
  for (int Lp = std::max(std::max(L - 2, 0), std::abs(M)); Lp <= L + 2; Lp++)
    cpl2 += gcoeff(Lp, M, 2, 0, L, M) * gcoeff(lj, mj, li, mi, Lp, M);
  return const0 * cpl0 + const2 * cpl2;
}
// pure <-> dummy index map
std::vector<int> pure_indices(const hfb200_ctx* c) {
  std::vector<int> idx;
  for (size_t i = 0; i < c->mval.size(); i++)
    for (int a = (c->mval[i] != 0 ? 1 : 0); a < c->nrad; a++) idx.push_back((int)i * c->nrad + a);
  return idx;
}
}  // namespace

extern "C" {

int hfb200_diatomic_create(hfb200_ctx** out, int device, int nang, const int* lval, const int* mval, int nel,
                           const int* first, const int* last, double Rhalf) {
  if (!out || nang < 1 || !lval || !mval || nel < 1 || !first || !last || !(Rhalf > 0.0)) {
    set_error("hfb200_diatomic_create: invalid argument");
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
  c->kind = CTX_DIATOMIC;
  c->device = device;
  c->nel = nel;
  c->Rhalf = Rhalf;
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

int hfb200_diatomic_set_tei(hfb200_ctx* c, int nlm, const int* Lmap, const int* Mmap, const double* const* P0,
                            const double* const* P2, const double* const* Q0, const double* const* Q2,
                            const double* const* prim_tei, const double* const* prim_ktei) {
  if (!c || c->kind != CTX_DIATOMIC || nlm < 1 || !Lmap || !Mmap || !P0 || !P2 || !Q0 || !Q2 || !prim_tei) {
    set_error("hfb200_diatomic_set_tei: invalid argument");
    return HFB200_EINVAL;
  }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  const int nang = (int)c->lval.size(), nel = c->nel;
  std::map<std::pair<int, int>, int> fam;
  for (int r = 0; r < nlm; r++) fam[{Lmap[r], Mmap[r]}] = r;
  ChanSpec S;
  S.nel = nel; S.nang = nang; S.stack = 2; S.nr = nlm;
  S.first = c->first; S.last = c->last;
  // tiling hint of the exchange contractions: functions with equal m AND equal parity of l couple to the same partners
  // (a Gaunt or cos^2-modified Gaunt coefficient <l_j m_j| L M |l_i m_i> vanishes unless l_j + L + l_i is even), so a
  // tile of such rows / items has a dense coefficient table instead of a checkerboard
  S.ang_group.resize(nang);
  for (int a = 0; a < nang; a++) S.ang_group[a] = 2 * c->mval[a] + (c->lval[a] & 1);
  S.ang_m = c->mval;
  S.fam_owned.resize(nlm);
  c->fam_owner = family_owner(nlm, c->shard_n);
  const std::vector<int> own = c->fam_owner;
  for (int r = 0; r < nlm; r++) S.fam_owned[r] = (own[r] == c->shard_rank);  // couplings of ALL families are listed below
  S.fac.resize(nlm);
  for (int r = 0; r < nlm; r++) {
    // LMfac = 4 pi Rh^5 (-1)^M (L-|M|)!/(L+|M|)!   (basis.cpp:1421)
    const int L = Lmap[r], M = Mmap[r];
    double ratio = 1.0;
    for (int p = L + M; p > L - M; p--) ratio *= p;
    S.fac[r] = 4.0 * M_PI * std::pow(c->Rhalf, 5) * ((M % 2) ? -1.0 : 1.0) / ratio;
  }
  // channels (L, M) of all pairs, sorted by family then M
  std::map<std::pair<int, int>, int> chan_tmp;  // (family, M) -> channel (after sorting)
  for (int k = 0; k < nang; k++)
    for (int l = 0; l < nang; l++) {
      const int lk = c->lval[k], ll = c->lval[l], M = c->mval[k] - c->mval[l];
      for (int L = std::max(std::abs(lk - ll) - 2, std::abs(M)); L <= lk + ll + 2; L++) {
        auto it = fam.find({L, std::abs(M)});
        if (it == fam.end()) { set_error("family (L=%d,|M|=%d) missing from the integral tables", L, std::abs(M)); return HFB200_EINVAL; }
        chan_tmp[{it->second, M}] = 0;
      }
    }
  for (auto& kv : chan_tmp) {
    kv.second = (int)S.chan_r.size();
    S.chan_r.push_back(kv.first.first);
  }
  // couplings (basis.cpp:1380-1403): coef[0] = cos^2-modified, coef[1] = plain Gaunt
  S.pair_coupl.assign((size_t)nang * nang, {});
  {
    int lmx = 0, mmx = 0;
    for (int k = 0; k < nang; k++) { lmx = std::max(lmx, c->lval[k]); mmx = std::max(mmx, std::abs(c->mval[k])); }
    gcoeff(2 * lmx + 4, 2 * mmx, 2 * lmx + 4, 2 * mmx, 0, 0);  // size the shared Gaunt table once, before the parallel loop
  }
  const std::map<std::pair<int, int>, int>& cfam = fam;
  const std::map<std::pair<int, int>, int>& cchan = chan_tmp;
#pragma omp parallel for schedule(dynamic)
  for (int k = 0; k < nang; k++)
    for (int l = 0; l < nang; l++) {
      const int lk = c->lval[k], mk = c->mval[k], ll = c->lval[l], ml = c->mval[l], M = mk - ml;
      for (int L = std::max(std::abs(lk - ll) - 2, std::abs(M)); L <= lk + ll + 2; L++) {
        const int r = cfam.find({L, std::abs(M)})->second;
        const double c0 = mod_coeff(lk, mk, L, M, ll, ml), c2 = gcoeff(lk, mk, L, M, ll, ml);
        if (c0 == 0.0 && c2 == 0.0) continue;
        S.pair_coupl[(size_t)k * nang + l].push_back({cchan.find({r, M})->second, {c0, c2}});
      }
    }
  // stacked integrals: fac [[T00, -T02], [-T20, T22]],  s = fac [P0; -P2],  b = [Q0; -Q2]
  const int sr = c->shard_rank;
  typedef const double* cptr;
  S.tei = [=](int r, int e, int ho, int hi, double* sg) { *sg = (ho != hi) ? -1.0 : 1.0; return own[r] == sr ? prim_tei[4 * (r * nel + e) + 2 * ho + hi] : (cptr) nullptr; };
  if (prim_ktei) S.ktei = [=](int r, int e, int a, int b, double* sg) { *sg = 1.0; return own[r] == sr ? prim_ktei[4 * (r * nel + e) + 2 * a + b] : (cptr) nullptr; };
  S.small_f = [=](int r, int e, int h, double* sg) { *sg = h ? -1.0 : 1.0; return own[r] != sr ? (cptr) nullptr : h ? P2[r * nel + e] : P0[r * nel + e]; };
  S.big_f = [=](int r, int e, int h, double* sg) { *sg = h ? -1.0 : 1.0; return own[r] != sr ? (cptr) nullptr : h ? Q2[r * nel + e] : Q0[r * nel + e]; };
  if (c->chan) { c->chan->release(); delete c->chan; }
  c->chan = new ChanEngine();
  c->chan->comm = c->comm;
  c->chan->comm_rank = c->shard_rank;
  c->chan->comm_size = c->comm ? c->shard_n : 1;
  const int rc = c->chan->build(S, prim_ktei != nullptr);
  if (rc) return rc;
  c->nL = nlm;
  c->have_ktei = prim_ktei != nullptr;
  return HFB200_OK;
}

static int diatomic_apply(hfb200_ctx* c, bool exchange, const double* P, int ldp, double* F, int ldf) {
  const std::vector<int> idx = pure_indices(c);
  const int npure = (int)idx.size(), nd = (int)c->lval.size() * c->nrad;
  if (ldp < npure || ldf < npure) { set_error("Matrix does not have expected size!"); return HFB200_EINVAL; }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  // expand_boundaries / remove_boundaries (basis.cpp:1735-1773) are folded into the engine's pack / unpack passes
  std::vector<int> map(nd, -1);
  for (int i = 0; i < npure; i++) map[idx[i]] = i;
  return exchange ? c->chan->exchange_dense(P, ldp, F, ldf, c->stream, map.data(), npure)
                  : c->chan->coulomb_dense(P, ldp, F, ldf, c->stream, map.data(), npure);
}

int hfb200_diatomic_coulomb(hfb200_ctx* c, const double* P, int ldp, double* J, int ldj) {
  if (!c || c->kind != CTX_DIATOMIC || !P || !J) { set_error("hfb200_diatomic_coulomb: invalid argument"); return HFB200_EINVAL; }
  if (!c->chan) { set_error("Primitive teis have not been computed!"); return HFB200_ENOTEI; }
  return diatomic_apply(c, false, P, ldp, J, ldj);
}

int hfb200_diatomic_exchange(hfb200_ctx* c, const double* P, int ldp, double* K, int ldk) {
  if (!c || c->kind != CTX_DIATOMIC || !P || !K) { set_error("hfb200_diatomic_exchange: invalid argument"); return HFB200_EINVAL; }
  if (!c->chan || !c->have_ktei) { set_error("Primitive teis have not been computed!"); return HFB200_ENOTEI; }
  return diatomic_apply(c, true, P, ldp, K, ldk);
}

}  // extern "C"
