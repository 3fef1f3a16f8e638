// Host-side driver of the channel-coupled Coulomb / exchange kernels; see chan_engine.h.
#include "chan_engine.h"
#include "../host/blockscan.h"

#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <map>
#include "../../include/helfem_b200.h"
#include "chan_exchange.cuh"
#include "chan_exchange_fused.cuh"

namespace hfb {

int comm_allreduce_sum(void* comm, double* buf, size_t n, cudaStream_t st);  // comm.cu
int comm_broadcast(void* comm, double* buf, size_t n, int root, cudaStream_t st);

namespace {
template <class T>
int up(ChanEngine* E, const T** dst, const std::vector<T>& src) {
  T* p = nullptr;
  const size_t n = std::max<size_t>(src.size(), 1);
  HFB_CUDA_OK(cudaMalloc(&p, n * sizeof(T)));
  if (!src.empty()) HFB_CUDA_OK(cudaMemcpy(p, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
  E->owned.push_back(p);
  *dst = p;
  return 0;
}
}  // namespace

int ChanEngine::build(const ChanSpec& S, bool with_exchange) {
  const int nel = S.nel, nang = S.nang, st = S.stack, nr = S.nr;
  first = S.first;
  last = S.last;
  int np = 0;
  for (int e = 0; e < nel; e++) np = std::max(np, last[e] - first[e] + 1);
  nrad = last[nel - 1] + 1;
  nchan = (int)S.chan_r.size();
  npair = nang * nang;
  const int np2 = np * np, ntri = np * (np + 1) / 2;
  const int MP = (st * ntri + 7) & ~7, KP = (st * ntri + 3) & ~3, MT = MP / 8, KS = KP / 4;
  prm.nel = nel; prm.np = np; prm.np2 = np2; prm.ntri = ntri; prm.stack = st; prm.MT = MT; prm.KS = KS;
  prm.nr = nr; prm.nchan = nchan; prm.npair = npair;
  fam_owned = S.fam_owned.size() == (size_t)nr ? S.fam_owned : std::vector<char>(nr, 1);
  ang_m = S.ang_m.size() == (size_t)nang ? S.ang_m : std::vector<int>(nang, 0);

  std::vector<int2> tri(ntri);
  for (int j = 0; j < np; j++)
    for (int i = 0; i <= j; i++) tri[i + j * (j + 1) / 2] = make_int2(i, j);
  int rc;
  if ((rc = up(this, &prm.tri, tri))) return rc;

  // ---- packed integrals and disjoint vectors of every family ----
  {
    std::vector<double> F((size_t)nr * nel * MT * KS * 32, 0.0), sv((size_t)nr * nel * MP, 0.0), bv((size_t)nr * nel * MP, 0.0);
#pragma omp parallel for collapse(2) schedule(dynamic)
    for (int r = 0; r < nr; r++)
      for (int e = 0; e < nel; e++) {
        const int npe = last[e] - first[e] + 1;
        const size_t n2 = (size_t)npe * npe;
        double* Fe = F.data() + ((size_t)r * nel + e) * MT * KS * 32;
        for (int h = 0; h < st; h++) {
          double sgs = 1.0, sgb = 1.0;
          const double* a = S.small_f(r, e, h, &sgs);
          const double* b = S.big_f(r, e, h, &sgb);
          if (!a || !b) continue;  // family not owned by this shard
          for (int m = 0; m < ntri; m++) {
            const int i = tri[m].x, j = tri[m].y;
            if (j >= npe) continue;
            sv[((size_t)r * nel + e) * MP + h * ntri + m] = S.fac[r] * sgs * a[i + j * npe];
            bv[((size_t)r * nel + e) * MP + h * ntri + m] = sgb * b[i + j * npe];
          }
          for (int hi = 0; hi < st; hi++) {
            double sg = 1.0;
            const double* t = S.tei(r, e, h, hi, &sg);
            if (!t) continue;
            for (int m = 0; m < ntri; m++) {
              const int i = tri[m].x, j = tri[m].y;
              if (j >= npe) continue;
              const int mm = h * ntri + m;
              for (int kk = 0; kk < ntri; kk++) {
                const int k = tri[kk].x, l = tri[kk].y;
                if (l >= npe) continue;
                const int kc = hi * ntri + kk;
                const double v = S.fac[r] * sg * t[(size_t)(i + j * npe) + (size_t)(k + l * npe) * n2];
                Fe[((size_t)(mm >> 3) * KS + (kc >> 2)) * 32 + (mm & 7) * 4 + (kc & 3)] = v;
              }
            }
          }
        }
      }
    if ((rc = up(this, &prm.teiF, F)) || (rc = up(this, &prm.svec, sv)) || (rc = up(this, &prm.bvec, bv))) return rc;
  }
  // ---- channel tables and the two CSR coupling lists ----
  {
    std::vector<int> fam_c0(nr + 1, 0);
    for (int c = 0; c < nchan; c++) {
      if (c && S.chan_r[c] < S.chan_r[c - 1]) { set_error("channels must be sorted by family"); return HFB200_EINVAL; }
      fam_c0[S.chan_r[c] + 1]++;
    }
    for (int r = 0; r < nr; r++) fam_c0[r + 1] += fam_c0[r];
    std::vector<std::vector<std::pair<int, const ChanCoupling*>>> by_chan(nchan);
    for (int pr = 0; pr < npair; pr++)
      for (const auto& cp : S.pair_coupl[pr])
        if (fam_owned[S.chan_r[cp.chan]]) by_chan[cp.chan].push_back({pr, &cp});  // ket contraction: owned channels only
    std::vector<int> pre_ptr(nchan + 1, 0), pre_pair, post_ptr(npair + 1, 0), post_chan, pall_ptr(npair + 1, 0), pall_chan;
    std::vector<double> pre_coef, post_coef, pall_coef;
    for (int c = 0; c < nchan; c++) {
      for (auto& en : by_chan[c]) {
        pre_pair.push_back(en.first);
        pre_coef.push_back(en.second->coef[0]);
        pre_coef.push_back(en.second->coef[1]);
      }
      pre_ptr[c + 1] = (int)pre_pair.size();
    }
    // bra contraction of block (i,j) uses the couplings of the pair (j,i): over the owned channels (partial operator, summed
    // by the caller) and over all channels (complete operator after the all-reduce of the channel results)
    for (int i = 0; i < nang; i++)
      for (int j = 0; j < nang; j++) {
        for (const auto& cp : S.pair_coupl[j * nang + i]) {
          pall_chan.push_back(cp.chan);
          pall_coef.push_back(cp.coef[0]);
          pall_coef.push_back(cp.coef[1]);
          if (!fam_owned[S.chan_r[cp.chan]]) continue;
          post_chan.push_back(cp.chan);
          post_coef.push_back(cp.coef[0]);
          post_coef.push_back(cp.coef[1]);
        }
        post_ptr[i * nang + j + 1] = (int)post_chan.size();
        pall_ptr[i * nang + j + 1] = (int)pall_chan.size();
      }
    if ((rc = up(this, &prm.chan_r, S.chan_r)) || (rc = up(this, &prm.fam_c0, fam_c0)) || (rc = up(this, &prm.pre_ptr, pre_ptr)) ||
        (rc = up(this, &prm.pre_pair, pre_pair)) || (rc = up(this, &prm.pre_coef, pre_coef)) || (rc = up(this, &prm.post_ptr, post_ptr)) ||
        (rc = up(this, &prm.post_chan, post_chan)) || (rc = up(this, &prm.post_coef, post_coef)) || (rc = up(this, &dPostAllPtr, pall_ptr)) ||
        (rc = up(this, &dPostAllChan, pall_chan)) || (rc = up(this, &dPostAllCoef, pall_coef)))
      return rc;
    fam_c0_host = fam_c0;
  }
  {
    std::vector<int> npe(nel), elem_of(2 * (size_t)nrad, -1);
    for (int e = 0; e < nel; e++) {
      npe[e] = last[e] - first[e] + 1;
      for (int a = first[e]; a <= last[e]; a++) {
        if (elem_of[2 * a] < 0) elem_of[2 * a] = e;
        else if (elem_of[2 * a + 1] < 0) elem_of[2 * a + 1] = e;
      }
    }
    if ((rc = up(this, &dFirstJ, first)) || (rc = up(this, &dNpeJ, npe)) || (rc = up(this, &dElemOfJ, elem_of))) return rc;
  }
  for (unsigned char** f : {&dPnz, &dCnz, &dOnz}) {
    const size_t nbytes = (f == &dCnz) ? (size_t)nchan : (size_t)npair;
    HFB_CUDA_OK(cudaMalloc(f, std::max<size_t>(nbytes, 1)));
    owned.push_back(*f);
  }
  {  // per-family ownership for the radial product: a rank of a sharded build skips the other ranks' families
    unsigned char* f = nullptr;
    HFB_CUDA_OK(cudaMalloc(&f, std::max<size_t>((size_t)nr, 1)));
    owned.push_back(f);
    HFB_CUDA_OK(cudaMemcpy(f, fam_owned.data(), (size_t)nr, cudaMemcpyHostToDevice));
    prm.fam_own = f;
  }
  have_k = false;
  if (with_exchange && S.ktei) return build_exchange(S);
  return 0;
}

// Exchange tables (chan_exchange.cuh): K-ordered in-element integrals swizzled into DMMA A-fragment order,
// full-square disjoint factors, and the coupling rows  (channel c, j) -> (i, c^0_c(j,i), c^1_c(j,i))
// (atomic/TwoDBasis.cpp:1023-1059, diatomic/basis.cpp:1604-1650).
int ChanEngine::build_exchange(const ChanSpec& S) {
  const int nel = S.nel, nang = S.nang, st = S.stack, nr = S.nr, np = prm.np, np2 = prm.np2, st2 = st * st;
  xprm.nel = nel; xprm.np = np; xprm.np2 = np2; xprm.stack = st; xprm.nrad = nrad; xprm.nang = nang;
  if (np > 32) { set_error("exchange: more than 32 radial functions per element are not supported"); return HFB200_EINVAL; }
  int rc;
  std::vector<int> npe(nel), elem_of(2 * (size_t)nrad, -1);
  for (int e = 0; e < nel; e++) {
    npe[e] = last[e] - first[e] + 1;
    for (int a = first[e]; a <= last[e]; a++) {
      if (elem_of[2 * a] < 0) elem_of[2 * a] = e;
      else if (elem_of[2 * a + 1] < 0) elem_of[2 * a + 1] = e;
      else { set_error("radial function %d belongs to more than two elements", a); return HFB200_EINVAL; }
    }
  }
  if ((rc = up(this, &xprm.first, first)) || (rc = up(this, &xprm.npe, npe)) || (rc = up(this, &xprm.elem_of, elem_of))) return rc;
  xprm.PB = (np2 + 7) & ~7;  // W sub-block of an element pair: whole 8-row m-tiles (64-byte multiples: bulk-copy granularity)
  xprm.WB = nel * nel * xprm.PB;
  fam_fac = S.fac;
  // disjoint factors, zero padded squares
  {
    std::vector<double> sm((size_t)nr * nel * st * np2, 0.0), bg((size_t)nr * nel * st * np2, 0.0);
    for (int r = 0; r < nr; r++)
      for (int e = 0; e < nel; e++)
        for (int h = 0; h < st; h++) {
          double s1 = 1.0, s2 = 1.0;
          const double* a = S.small_f(r, e, h, &s1);
          const double* b = S.big_f(r, e, h, &s2);
          if (!a || !b) continue;  // family not owned by this shard
          // signs of the stacked Coulomb vectors do not apply here: exchange signs live in the couplings
          for (int j = 0; j < npe[e]; j++)
            for (int i = 0; i < npe[e]; i++) {
              sm[(((size_t)r * nel + e) * st + h) * np2 + i + j * np] = a[i + j * npe[e]];
              bg[(((size_t)r * nel + e) * st + h) * np2 + i + j * np] = b[i + j * npe[e]];
            }
        }
    if ((rc = up(this, &xprm.small_f, sm)) || (rc = up(this, &xprm.big_f, bg))) return rc;
  }
  // K-ordered integrals: rows (j,k) x cols (i,l), zero padded to np, in A-fragment order [mtile][kstep][lane]
  {
    const int MT = (np2 + 7) / 8, KS = (np2 + 3) / 4;
    int LDK = 4 * KS;
    while ((LDK & 15) != 4) LDK++;
    xprm.MT = MT; xprm.KS = KS; xprm.LDK = LDK;
    const size_t blk = (size_t)MT * KS * 32;
    double* d = nullptr;
    HFB_CUDA_OK(cudaMalloc(&d, (size_t)nr * nel * st2 * blk * sizeof(double)));
    owned.push_back(d);
    xprm.teiK = d;
    HFB_CUDA_OK(cudaMemset(d, 0, (size_t)nr * nel * st2 * blk * sizeof(double)));
    const int nstage = 16;
    std::vector<std::vector<double>> stage(nstage, std::vector<double>(blk));
    const long nblk = (long)nr * nel * st2;
    for (long b0 = 0; b0 < nblk; b0 += nstage) {
      const int nb = (int)std::min<long>(nstage, nblk - b0);
      std::vector<char> present(nb, 0);
#pragma omp parallel for schedule(dynamic)
      for (int q = 0; q < nb; q++) {
        const long id = b0 + q;
        const int ab = (int)(id % st2), e = (int)((id / st2) % nel), r = (int)(id / ((long)st2 * nel));
        double sg = 1.0;
        const double* t = S.ktei(r, e, ab / st, ab % st, &sg);
        if (!t) continue;
        present[q] = 1;
        std::vector<double>& F = stage[q];
        std::fill(F.begin(), F.end(), 0.0);
        const int n = npe[e];
        const size_t n2 = (size_t)n * n;
        for (int cs = 0; cs < n; cs++)
          for (int cf = 0; cf < n; cf++) {
            const int k = cf + cs * np;
            const double* col = t + (size_t)(cf + cs * n) * n2;
            for (int rs = 0; rs < n; rs++)
              for (int rf = 0; rf < n; rf++) {
                const int m = rf + rs * np;
                F[((size_t)(m >> 3) * KS + (k >> 2)) * 32 + (m & 7) * 4 + (k & 3)] = col[rf + rs * n];
              }
          }
      }
      for (int q = 0; q < nb; q++)
        if (present[q]) HFB_CUDA_OK(cudaMemcpy(d + (size_t)(b0 + q) * blk, stage[q].data(), blk * sizeof(double), cudaMemcpyHostToDevice));
    }
  }
  // coupling rows by channel
  {
    std::vector<int> cnt((size_t)nchan * nang + 1, 0);
    for (int j = 0; j < nang; j++)
      for (int i = 0; i < nang; i++)
        for (const auto& cp : S.pair_coupl[(size_t)j * nang + i])
          if (fam_owned[S.chan_r[cp.chan]] && (cp.coef[0] != 0.0 || (st == 2 && cp.coef[1] != 0.0))) cnt[(size_t)cp.chan * nang + j + 1]++;
    for (size_t q = 0; q < (size_t)nchan * nang; q++) cnt[q + 1] += cnt[q];
    xrowptr = cnt;
    xrow_i.assign(cnt.back(), 0);
    std::vector<double>& coef = xrow_coef;
    coef.assign(2 * (size_t)cnt.back(), 0.0);
    std::vector<int> fill(cnt.begin(), cnt.end() - 1);
    for (int j = 0; j < nang; j++)
      for (int i = 0; i < nang; i++)  // i ascending within a row
        for (const auto& cp : S.pair_coupl[(size_t)j * nang + i])
          if (fam_owned[S.chan_r[cp.chan]] && (cp.coef[0] != 0.0 || (st == 2 && cp.coef[1] != 0.0))) {
            const int pos = fill[(size_t)cp.chan * nang + j]++;
            xrow_i[pos] = i;
            coef[2 * (size_t)pos] = cp.coef[0];
            coef[2 * (size_t)pos + 1] = st == 2 ? cp.coef[1] : 0.0;
          }
    fam_rows.assign(nr, 0);  // non-empty coupling rows of a family (0: family not coupled / held by another rank)
    nch_max = 1;
    for (int r = 0; r < nr; r++) {
      nch_max = std::max(nch_max, fam_c0_host[r + 1] - fam_c0_host[r]);
      for (int c = fam_c0_host[r]; c < fam_c0_host[r + 1]; c++)
        for (int j = 0; j < nang; j++)
          if (xrowptr[(size_t)c * nang + j + 1] > xrowptr[(size_t)c * nang + j]) fam_rows[r]++;
    }
    // x1 tiles: rows of a channel grouped by angular group (same m => same coupling partners), CHX_T rows each
    ang_group = S.ang_group.size() == (size_t)nang ? S.ang_group : std::vector<int>(nang, 0);
    std::vector<int> tiles1;
    fam_t1.assign(nr + 1, 0);
    for (int r = 0; r < nr; r++) {
      for (int c = fam_c0_host[r]; c < fam_c0_host[r + 1]; c++) {
        std::vector<int> js;
        for (int j = 0; j < nang; j++)
          if (xrowptr[(size_t)c * nang + j + 1] > xrowptr[(size_t)c * nang + j]) js.push_back(j);
        std::stable_sort(js.begin(), js.end(), [&](int a, int b) { return ang_group[a] < ang_group[b]; });
        for (size_t q = 0; q < js.size();) {
          size_t e = q;
          while (e < js.size() && e - q < (size_t)CHX_T && ang_group[js[e]] == ang_group[js[q]]) e++;
          tiles1.push_back(c);
          tiles1.push_back((int)(e - q));
          for (int t = 0; t < CHX_T; t++) tiles1.push_back(q + t < e ? js[q + t] : js[q]);
          q = e;
        }
      }
      fam_t1[r + 1] = (int)(tiles1.size() / (2 + CHX_T));
    }
    const int* dt1 = nullptr;
    if ((rc = up(this, &dt1, tiles1))) return rc;
    dTiles1 = const_cast<int*>(dt1);
    if ((rc = up(this, &xprm.rowptr, xrowptr)) || (rc = up(this, &xprm.row_i, xrow_i)) || (rc = up(this, &xprm.row_coef, coef))) return rc;
  }
  {
    unsigned char* p = nullptr;
    HFB_CUDA_OK(cudaMalloc(&p, 2 * ((size_t)nang * nang + 8)));  // + the symmetry flag of P; second half: exact non-zero flags
    owned.push_back(p);
    dNzP = p;
    dNzX = p + (size_t)nang * nang + 8;
    if (hNzP) cudaFreeHost(hNzP);
    HFB_CUDA_OK(cudaMallocHost(&hNzP, (size_t)nang * nang + 8));
  }
  if ((rc = up(this, &dFamFac, fam_fac))) return rc;
  // items per CTA of the fused kernels: as many as the shared-memory tiles allow with two CTAs per SM
  ni_same = 32;
  while (ni_same > 8 && ((size_t)ni_same * xprm.LDK + chxf_inline_ring_doubles(xprm.PB)) * sizeof(double) > 110 * 1024) ni_same >>= 1;
  ni_pair = (np <= 16) ? 8 : 4;
  mask_cached.clear();
  have_k = true;
  return 0;
}

namespace {
template <class T>
int up_list(ChanEngine::DevList& L, const std::vector<T>& v) {
  const size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(T);
  if (bytes > L.cap) {
    if (L.p) cudaFree(L.p);
    L.p = nullptr;
    L.cap = 0;
    HFB_CUDA_OK(cudaMalloc(&L.p, bytes));
    L.cap = bytes;
  }
  if (!v.empty()) HFB_CUDA_OK(cudaMemcpy(L.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}
}  // namespace

// Everything that depends on the sparsity pattern of P (one block mask), built once per pattern:
//  * L(c,j) = { l : some partner i of the coupling row (c,j) has P[i,l] != 0 } -> compact layout of W
//  * items of family r: (j,k) such that some channel c of r has rows (c,j), (c,k) and L(c,j) meets row (c,k);
//    with an exactly symmetric P only j <= k (K[k,j] = K[j,k]^T is written by x4)
//  * tiles (<= CHX_T consecutive items with one j), their slot tables for the fused kernels, item groups,
//    the CSR of distinct output blocks for x4
//  * batches of consecutive families whose W and output workspaces fit the budget
int ChanEngine::build_items(const unsigned char* nz, bool sym) {
  const int nang = xprm.nang, nr = prm.nr, nw = (nang + 63) / 64, st = prm.stack, nel = prm.nel, np2 = prm.np2;
  std::vector<uint64_t> nzrow((size_t)nang * nw, 0);  // nzrow[i] = { l : P[i,l] != 0 }
  for (int i = 0; i < nang; i++)
    for (int l = 0; l < nang; l++)
      if (nz[(size_t)i * nang + l]) nzrow[(size_t)i * nw + (l >> 6)] |= (uint64_t)1 << (l & 63);
  std::vector<unsigned long long> wbits((size_t)nchan * nang * nw, 0);
  std::vector<int> wcount((size_t)nchan * nang, 0);
  std::vector<std::vector<int2>> fam_items(nr);
  // Symmetric P: of every pair of output blocks (j,k), (k,j) only ONE is contracted (x4 writes the other as its transpose).  Which
  // one is chosen CIRCULARLY inside an m class -- (j,k) is kept when k follows j by at most half the class size -- so that every row
  // j keeps about half of its partners (a triangle j <= k would leave rows with 1 .. n partners and half-empty tiles of 8).
  std::vector<int> mpos(nang, 0), mcnt(nang, 0);
  for (int a = 0; a < nang; a++)
    for (int b = 0; b < nang; b++)
      if (ang_m[b] == ang_m[a]) { mcnt[a]++; if (b < a) mpos[a]++; }
  auto keep = [&](int j, int k) {
    if (j == k) return true;
    if (ang_m[j] != ang_m[k]) return j < k;
    const int n = mcnt[j], d = (mpos[k] - mpos[j] + n) % n;
    return 2 * d < n || (2 * d == n && mpos[j] < mpos[k]);
  };
#pragma omp parallel
  {
    std::vector<uint64_t> rowset((size_t)nang * nw);
    std::vector<unsigned char> hit((size_t)nang * nang);
    std::vector<int> js;
#pragma omp for schedule(dynamic)
    for (int r = 0; r < nr; r++) {
      if (fam_rows[r] == 0) continue;
      std::fill(hit.begin(), hit.end(), 0);
      for (int c = fam_c0_host[r]; c < fam_c0_host[r + 1]; c++) {
        js.clear();
        for (int j = 0; j < nang; j++) {
          const int e0 = xrowptr[(size_t)c * nang + j], e1 = xrowptr[(size_t)c * nang + j + 1];
          if (e0 == e1) continue;
          js.push_back(j);
          unsigned long long* ls = &wbits[((size_t)c * nang + j) * nw];
          uint64_t* rs = &rowset[(size_t)j * nw];
          for (int w = 0; w < nw; w++) rs[w] = 0;
          for (int e = e0; e < e1; e++) {
            const int i = xrow_i[e];
            rs[i >> 6] |= (uint64_t)1 << (i & 63);
            for (int w = 0; w < nw; w++) ls[w] |= nzrow[(size_t)i * nw + w];
          }
          int n = 0;
          for (int w = 0; w < nw; w++) n += __builtin_popcountll(ls[w]);
          wcount[(size_t)c * nang + j] = n;
        }
        for (int j : js) {
          const unsigned long long* ls = &wbits[((size_t)c * nang + j) * nw];
          for (int k : js) {
            if (sym && !keep(j, k)) continue;
            if (hit[(size_t)j * nang + k]) continue;
            uint64_t any = 0;
            for (int w = 0; w < nw; w++) any |= ls[w] & rowset[(size_t)k * nw + w];
            if (any) hit[(size_t)j * nang + k] = 1;
          }
        }
      }
      for (int j = 0; j < nang; j++) {
        const size_t b0 = fam_items[r].size();
        for (int k = 0; k < nang; k++)
          if (hit[(size_t)j * nang + k]) fam_items[r].push_back(make_int2(j, k));
        // within one j: order by angular group so that the tiles (consecutive items) share their partners
        std::stable_sort(fam_items[r].begin() + b0, fam_items[r].end(),
                         [&](const int2& a, const int2& b) { return ang_group[a.y] < ang_group[b.y]; });
      }
      if (sym) {  // rows (c, j) that no kept item (j, k) uses produce nothing that is read: drop their W blocks
        for (int c = fam_c0_host[r]; c < fam_c0_host[r + 1]; c++)
          for (int j = 0; j < nang; j++) {
            if (!wcount[(size_t)c * nang + j]) continue;
            bool used = false;
            for (int k = 0; k < nang && !used; k++) used = hit[(size_t)j * nang + k] != 0;
            if (!used) {
              wcount[(size_t)c * nang + j] = 0;
              for (int w = 0; w < nw; w++) wbits[((size_t)c * nang + j) * nw + w] = 0;
            }
          }
      }
    }
  }
  // global lists + batches
  // workspace budgets of a batch (HFB200_X_WCAP_MB / HFB200_X_RCAP_MB override them: the tests force many small batches)
  size_t wcap_bytes = 16ull << 30, rcap_bytes = 12ull << 30;
  if (const char* e = getenv("HFB200_X_WCAP_MB")) wcap_bytes = std::max<size_t>(1, (size_t)atoll(e)) << 20;
  if (const char* e = getenv("HFB200_X_RCAP_MB")) rcap_bytes = std::min<size_t>(std::max<size_t>(1, (size_t)atoll(e)) << 20, 30ull << 30);
  const long long wcap = (long long)(wcap_bytes / ((size_t)st * xprm.WB * sizeof(double)));
  const long long icap = (long long)(rcap_bytes / ((size_t)nel * nel * np2 * sizeof(double)));  // output sub-blocks of the batch
  std::vector<int2> items, tiles2;
  std::vector<int4> groupsS, groupsP;
  std::vector<int> item_fam, out_ptr, out_item;
  std::vector<long long> wbase((size_t)nchan * nang, -1);
  batches.clear();
  max_wblocks = 0;
  max_batch_items = 0;
  XBatch cur{fam_t1[0], fam_t1[0], 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  auto close_batch = [&]() {
    cur.item1 = (int)items.size();
    cur.t2_1 = (int)tiles2.size();
    cur.gs1 = (int)groupsS.size();
    cur.gp1 = (int)groupsP.size();
    // distinct output blocks of the batch, items of one block in family order
    std::vector<int> idx(cur.item1 - cur.item0);
    for (size_t q = 0; q < idx.size(); q++) idx[q] = cur.item0 + (int)q;
    std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) {
      return (long long)items[a].x * nang + items[a].y < (long long)items[b].x * nang + items[b].y;
    });
    cur.out0 = (int)out_ptr.size();
    cur.nout = 0;
    for (size_t q = 0; q < idx.size(); q++) {
      if (q == 0 || items[idx[q]].x != items[idx[q - 1]].x || items[idx[q]].y != items[idx[q - 1]].y) {
        out_ptr.push_back((int)out_item.size());
        cur.nout++;
      }
      out_item.push_back(idx[q]);
    }
    out_ptr.push_back((int)out_item.size());
    if (cur.item1 > cur.item0 || cur.wblocks > 0) batches.push_back(cur);
    max_wblocks = std::max(max_wblocks, cur.wblocks);
    max_batch_items = std::max(max_batch_items, cur.item1 - cur.item0);
  };
  auto make_groups = [&](std::vector<int4>& groups, int t_begin, int ni) {  // dense packing of a family's tiles into groups of <= ni items
    for (int t = t_begin; t < (int)tiles2.size();) {
      int e = t, n = 0;
      while (e < (int)tiles2.size() && e - t < CHXF_TMAX && n + tiles2[e].y <= ni) n += tiles2[e++].y;
      groups.push_back(make_int4(tiles2[t].x, n, t, e - t));
      t = e;
    }
  };
  for (int r = 0; r < nr; r++) {
    long long wr = 0;
    for (int c = fam_c0_host[r]; c < fam_c0_host[r + 1]; c++)
      for (int j = 0; j < nang; j++) wr += wcount[(size_t)c * nang + j];
    const long long ir = (long long)fam_items[r].size();
    const bool nonempty = cur.wblocks > 0 || (int)items.size() > cur.item0;
    if (nonempty && (cur.wblocks + wr > wcap || (long long)items.size() - cur.item0 + ir > icap)) {
      cur.t1_1 = fam_t1[r];
      close_batch();
      cur = XBatch{fam_t1[r], fam_t1[r], (int)tiles2.size(), 0, (int)groupsS.size(), 0, (int)groupsP.size(), 0, (int)items.size(), 0, 0, 0, 0};
    }
    for (int c = fam_c0_host[r]; c < fam_c0_host[r + 1]; c++)
      for (int j = 0; j < nang; j++)
        if (wcount[(size_t)c * nang + j]) {
          wbase[(size_t)c * nang + j] = cur.wblocks;
          cur.wblocks += wcount[(size_t)c * nang + j];
        }
    const std::vector<int2>& it = fam_items[r];
    const int g0 = (int)items.size(), tb = (int)tiles2.size();
    for (size_t q = 0; q < it.size();) {  // tiles: <= CHX_T consecutive items with one j and one angular group of k
      size_t e = q;
      while (e < it.size() && e - q < (size_t)CHX_T && it[e].x == it[q].x && ang_group[it[e].y] == ang_group[it[q].y]) e++;
      tiles2.push_back(make_int2(g0 + (int)q, (int)(e - q)));
      q = e;
    }
    make_groups(groupsS, tb, ni_same);
    make_groups(groupsP, tb, ni_pair);
    items.insert(items.end(), it.begin(), it.end());
    item_fam.insert(item_fam.end(), it.size(), r);
    cur.t1_1 = fam_t1[r + 1];
  }
  cur.t1_1 = fam_t1[nr];
  close_batch();
  total_items = (long long)items.size();

  // ---- slot tables of the tiles (fused kernels): union of the W blocks (c, l) the tile's rows couple to, in
  //      ascending (channel, l) order, and the dense coefficient table [slot][b][item of the tile] ----
  const int ntile = (int)tiles2.size();
  std::vector<int2> tile_u(ntile);
  std::vector<std::vector<int>> tslots(ntile);
#pragma omp parallel
  {
    std::vector<int> dom((size_t)std::max(nch_max, 1) * nang);
#pragma omp for schedule(dynamic, 64)
    for (int t = 0; t < ntile; t++) {
      const int it0 = tiles2[t].x, nk = tiles2[t].y, j = items[it0].x, r = item_fam[it0];
      const int c0 = fam_c0_host[r], nch = fam_c0_host[r + 1] - c0;
      std::fill(dom.begin(), dom.begin() + (size_t)nch * nang, 0);
      for (int kk = 0; kk < nk; kk++) {
        const int k = items[it0 + kk].y;
        for (int cl = 0; cl < nch; cl++) {
          const size_t row = (size_t)(c0 + cl) * nang + k, jrow = (size_t)(c0 + cl) * nang + j;
          const unsigned long long* bits = &wbits[jrow * nw];
          for (int e = xrowptr[row]; e < xrowptr[row + 1]; e++) {
            const int l = xrow_i[e];
            if ((bits[l >> 6] >> (l & 63)) & 1ull) dom[(size_t)cl * nang + l] = 1;
          }
        }
      }
      for (int q = 0; q < nch * nang; q++)
        if (dom[q]) tslots[t].push_back(q);
    }
  }
  size_t nslots = 0;
  for (int t = 0; t < ntile; t++) {
    const int nu = ((int)tslots[t].size() + CHXF_SLOTS - 1) / CHXF_SLOTS * CHXF_SLOTS;
    tile_u[t] = make_int2((int)nslots, nu);
    nslots += nu;
  }
  if (nslots >= ((size_t)1 << 31)) { set_error("exchange: slot tables exceed 32-bit indexing"); return HFB200_EINVAL; }
  std::vector<long long> slot_wblk(std::max<size_t>(nslots, 1), 0);
  std::vector<double> slot_coef(std::max<size_t>(nslots, 1) * 16, 0.0);
#pragma omp parallel
  {
    std::vector<int> slot_of((size_t)std::max(nch_max, 1) * nang);
#pragma omp for schedule(dynamic, 64)
    for (int t = 0; t < ntile; t++) {
      const int it0 = tiles2[t].x, nk = tiles2[t].y, j = items[it0].x, r = item_fam[it0];
      const int c0 = fam_c0_host[r], nch = fam_c0_host[r + 1] - c0;
      const size_t u0 = (size_t)tile_u[t].x;
      const std::vector<int>& sl = tslots[t];
      for (size_t u = 0; u < sl.size(); u++) {
        const int cl = sl[u] / nang, l = sl[u] - cl * nang;
        const size_t jrow = (size_t)(c0 + cl) * nang + j;
        const unsigned long long* bits = &wbits[jrow * nw];
        long long idx = wbase[jrow];
        for (int w = 0; w < (l >> 6); w++) idx += __builtin_popcountll(bits[w]);
        idx += __builtin_popcountll(bits[l >> 6] & (((unsigned long long)1 << (l & 63)) - 1ull));
        slot_wblk[u0 + u] = idx;
        slot_of[sl[u]] = (int)u;
      }
      for (size_t u = sl.size(); u < (size_t)tile_u[t].y; u++) slot_wblk[u0 + u] = sl.empty() ? 0 : slot_wblk[u0];  // padding: zero coefficients
      for (int kk = 0; kk < nk; kk++) {
        const int k = items[it0 + kk].y;
        for (int cl = 0; cl < nch; cl++) {
          const size_t row = (size_t)(c0 + cl) * nang + k, jrow = (size_t)(c0 + cl) * nang + j;
          const unsigned long long* bits = &wbits[jrow * nw];
          for (int e = xrowptr[row]; e < xrowptr[row + 1]; e++) {
            const int l = xrow_i[e];
            if (!((bits[l >> 6] >> (l & 63)) & 1ull)) continue;
            const size_t u = u0 + slot_of[(size_t)cl * nang + l];
            slot_coef[u * 16 + kk] = xrow_coef[2 * (size_t)e];
            slot_coef[u * 16 + 8 + kk] = xrow_coef[2 * (size_t)e + 1];
          }
        }
      }
    }
  }
  // blocks (j,k) of K that can be non-zero for this mask: m_j - m_k = m_i - m_l for some non-zero P[i,l] (the same list on
  // every rank of a sharded run: it depends on the mask only)
  {
    int mlo = 0, mhi = 0;
    for (int a = 0; a < nang; a++) { mlo = std::min(mlo, ang_m[a]); mhi = std::max(mhi, ang_m[a]); }
    const int span = mhi - mlo;
    std::vector<char> dset(2 * span + 1, 0);
    for (int i = 0; i < nang; i++)
      for (int l = 0; l < nang; l++)
        if (nz[(size_t)i * nang + l]) dset[ang_m[i] - ang_m[l] + span] = 1;
    red_blocks.clear();
    for (int j = 0; j < nang; j++)
      for (int k = sym ? j : 0; k < nang; k++)
        if (dset[ang_m[j] - ang_m[k] + span]) red_blocks.push_back(make_int2(j, k));
  }
  {
    long long wtot = 0, nout = 0, real = 0, mac = 0;
    for (const XBatch& B : batches) { wtot += B.wblocks; nout += B.nout; }
    for (int t = 0; t < ntile; t++) { real += (long long)tslots[t].size(); mac += (long long)tslots[t].size() * tiles2[t].y; }
    counts[0] = total_items; counts[1] = ntile; counts[2] = real; counts[3] = (long long)nslots; counts[4] = wtot; counts[5] = xprm.WB;
    counts[6] = nout; counts[7] = (long long)batches.size(); counts[8] = sym ? 1 : 0; counts[9] = mac; counts[10] = (long long)red_blocks.size();
  }
  if (getenv("HFB200_X_VERBOSE")) {
    long long wtot = 0;
    for (const XBatch& B : batches) wtot += B.wblocks;
    size_t real = 0;
    for (int t = 0; t < ntile; t++) real += tslots[t].size();
    fprintf(stderr, "[hfb200] exchange lists: sym %d, items %lld, tiles %d, slots %zu (padded %zu), groups %zu / %zu, W blocks %lld (%.1f GB), "
            "batches %zu\n", (int)sym, total_items, ntile, real, nslots, groupsS.size(), groupsP.size(), wtot,
            (double)wtot * st * xprm.WB * 8e-9, batches.size());
  }
  int rc;
  if ((rc = up_list(dItems, items)) || (rc = up_list(dItemFam, item_fam)) || (rc = up_list(dTiles2, tiles2)) ||
      (rc = up_list(dGroupsS, groupsS)) || (rc = up_list(dGroupsP, groupsP)) || (rc = up_list(dTileU, tile_u)) ||
      (rc = up_list(dSlotWblk, slot_wblk)) || (rc = up_list(dSlotCoef, slot_coef)) || (rc = up_list(dWbase, wbase)) ||
      (rc = up_list(dWbits, wbits)) || (rc = up_list(dOutPtr, out_ptr)) || (rc = up_list(dOutItem, out_item)) ||
      (rc = up_list(dRedBlocks, red_blocks)))
    return rc;
  return 0;
}

namespace {
// the max-dynamic-shared-memory attribute is set per launch: it belongs to the current device's context (one process can
// hold contexts on several GPUs)
template <int NI>
int launch_xf_same(const ChanXParams& q, const int4* groups, int ngroups, const int2* tiles, int nsm, cudaStream_t st) {
  (void)nsm;
  const size_t smem = ((size_t)NI * q.LDK + chxf_inline_ring_doubles(q.PB)) * sizeof(double);
  HFB_CUDA_OK(cudaFuncSetAttribute(chan_xf_same_kernel<NI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  chan_xf_same_kernel<NI><<<dim3(ngroups, q.nel), 256, smem, st>>>(q, groups, tiles);
  HFB_CUDA_OK(cudaGetLastError());
  return 0;
}
template <int NI, int NTL>
int launch_xf_pair(const ChanXParams& q, const int4* groups, int ngroups, const int2* tiles, int nsm, cudaStream_t st) {
  if (q.nel < 2) return 0;
  const size_t smem = ((size_t)NI * q.stack * ChxfPair<NTL>::RS + (size_t)8 * ChxfPair<NTL>::YS + chxf_ring_doubles(q.PB) + chxf_pair_table_doubles(q.np2, q.stack)) * sizeof(double);
  HFB_CUDA_OK(cudaFuncSetAttribute(chan_xf_pair_kernel<NI, NTL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long units = (long long)ngroups * q.nel * (q.nel - 1);
  chan_xf_pair_kernel<NI, NTL><<<(unsigned)std::min<long long>(units, (NTL <= 2 ? 2LL : 1LL) * nsm), CHXF_THREADS, smem, st>>>(q, groups, ngroups, tiles);
  HFB_CUDA_OK(cudaGetLastError());
  return 0;
}
}  // namespace

// pre / DMMA / post kernels on the device bands dP -> dJ (buffers reserved by the caller)
int ChanEngine::coulomb_kernels(int nang, cudaStream_t st) {
  (void)nang;
  const int nel = prm.nel;
  ChanParams q = prm;
  q.nb = 1; q.P = dP.p; q.pc = dPc.p; q.traces = dTr.p; q.jc = dJc.p; q.J = dJ.p;
  q.pnz = dPnz; q.cnz = dCnz; q.onz = dOnz;  // dPnz was filled by the caller's pack pass
  if (4 * prm.KS <= 512) chan_pre_kernel<<<dim3(nel, nchan, 1), 128, 0, st>>>(q);
  else chan_pre_generic_kernel<<<dim3(nel, nchan, 1), 128, 0, st>>>(q);
  HFB_CUDA_OK(cudaGetLastError());
  if (comm) HFB_CUDA_OK(cudaMemsetAsync(dJc.p, 0, (size_t)nchan * nel * 8 * prm.MT * sizeof(double), st));  // skipped families sum as zeros
  int maxcol = 1;
  for (int r = 0; r < prm.nr; r++) maxcol = std::max(maxcol, fam_c0_host[r + 1] - fam_c0_host[r]);
  const int rts = (prm.MT + 3) / 4;
  if (maxcol <= 8) {
    chan_jmma_kernel<true, 1><<<dim3(nel * rts, 1, prm.nr), 256, 0, st>>>(q);
  } else if (maxcol <= 16) {
    chan_jmma_kernel<true, 2><<<dim3(nel * rts, 1, prm.nr), 256, 0, st>>>(q);
  } else {
    chan_jmma_kernel<false, 2><<<dim3(nel * rts, (maxcol + 127) / 128, prm.nr), 256, 0, st>>>(q);
  }
  HFB_CUDA_OK(cudaGetLastError());
  if (comm) {  // sum the per-channel radial results over the ranks (each holds its own families), then contract ALL channels
    int rc;
    if ((rc = comm_allreduce_sum(comm, dJc.p, (size_t)nchan * nel * 8 * prm.MT, st))) return rc;
    q.post_ptr = dPostAllPtr; q.post_chan = dPostAllChan; q.post_coef = dPostAllCoef;
  }
  chan_flag_kernel<<<nchan, 128, 0, st>>>(q);
  HFB_CUDA_OK(cudaGetLastError());
  HFB_CUDA_OK(cudaMemsetAsync(dOnz, 0, (size_t)npair, st));
  chan_post_kernel<<<dim3(npair, 1, 1), 256, 0, st>>>(q);
  HFB_CUDA_OK(cudaGetLastError());
  return 0;
}

int ChanEngine::coulomb_device(const double* dPm, int ldp, double* dJm, int ldj, cudaStream_t st, const unsigned char* nzx) {
  const int nel = prm.nel, np2 = prm.np2, nang = (int)std::lround(std::sqrt((double)npair));
  const size_t nband = (size_t)npair * nel * np2;
  const int KP = 4 * prm.KS, MP = 8 * prm.MT;
  int rc;
  if ((rc = dP.reserve(nband)) || (rc = dJ.reserve(nband)) || (rc = dPc.reserve((size_t)nchan * nel * KP)) ||
      (rc = dTr.reserve((size_t)nchan * nel * 2)) || (rc = dJc.reserve((size_t)nchan * nel * MP)))
    return rc;
  HFB_CUDA_OK(cudaMemsetAsync(dPnz, 0, (size_t)npair, st));
  chan_band_pack_kernel<<<dim3(nel * npair), 128, 0, st>>>(prm, dFirstJ, dNpeJ, nang, nrad, dPm, ldp, dP.p, dPnz, nzx);
  HFB_CUDA_OK(cudaGetLastError());
  if ((rc = coulomb_kernels(nang, st))) return rc;
  chan_band_unpack_kernel<<<(unsigned)npair, 256, 0, st>>>(prm, dFirstJ, dElemOfJ, nang, nrad, dJ.p, dJm, ldj, dOnz);
  HFB_CUDA_OK(cudaGetLastError());
  return 0;
}

// Host-matrix operators (the reference's interface: dense arma::mat in, dense zero-filled arma::mat out;
// atomic/TwoDBasis.cpp:817,957, diatomic/basis.cpp:1359,1532 with expand / remove_boundaries :1735-1773 folded in).
// map (optional): operator index -> row / column of the caller's nout x nout matrices, -1 for the rows that are identically
// zero (the diatomic boundary functions); an angular function's rows must be consecutive in the caller's matrix.
// Only what can be non-zero crosses PCIe: the host scans P for its non-zero angular blocks (the same Frobenius-free test
// "any non-zero entry"; blocks below the reference's 10 eps norm screen are dropped again on the device) and uploads those,
// packed; the result comes back as the packed blocks (j,k) with m_j - m_k = m_i - m_l for some non-zero P[i,l] -- every other
// block of J or K vanishes identically -- and is scattered into the caller's matrix, which is zero-filled while the GPU works.
int ChanEngine::dense_apply(bool exchange, const double* P, int ldp, double* F, int ldf, cudaStream_t st, const int* map, int nout) {
  if (exchange && !have_k) { set_error("Primitive teis have not been computed!"); return HFB200_ENOTEI; }
  const int nang = (int)std::lround(std::sqrt((double)npair)), ntot = nang * nrad;
  const size_t nrad2 = (size_t)nrad * nrad;
  if (!map) nout = ntot;
  int rc;
  // root-driven communicator (hfb200_comm_set_root): only the root passes host matrices; it broadcasts the list of non-zero
  // blocks and their packed contents over NVLink, the other ranks take part through serve() with P = F = nullptr
  const bool driven = comm && comm_root >= 0;
  const bool worker = driven && comm_rank != comm_root;
  if (worker != (P == nullptr)) {
    set_error(worker ? "this rank is a worker of a root-driven communicator: call hfb200_chan_serve" : "dense_apply: null matrix");
    return HFB200_EINVAL;
  }
  const size_t nhdr = 2 * (size_t)npair + 2;
  if (worker) {
    if ((rc = dHdr.reserve(nhdr)) || (rc = comm_broadcast(comm, dHdr.p, nhdr, comm_root, st))) return rc;
    std::vector<double> hdr(nhdr);
    HFB_CUDA_OK(cudaMemcpyAsync(hdr.data(), dHdr.p, nhdr * sizeof(double), cudaMemcpyDeviceToHost, st));
    HFB_CUDA_OK(cudaStreamSynchronize(st));
    const size_t nin = (size_t)hdr[0];
    std::vector<int2> inb(nin);
    for (size_t q = 0; q < nin; q++) inb[q] = make_int2((int)hdr[1 + 2 * q], (int)hdr[2 + 2 * q]);
    if ((rc = dPdense.reserve((size_t)ntot * ntot)) || (rc = dKdense.reserve((size_t)ntot * ntot)) ||
        (rc = dRed.reserve(std::max<size_t>(nin, 1) * nrad2)) || (rc = up_list(dIoBlocks, inb)))
      return rc;
    if ((rc = comm_broadcast(comm, dRed.p, nin * nrad2, comm_root, st))) return rc;
    HFB_CUDA_OK(cudaMemsetAsync(dPdense.p, 0, (size_t)ntot * ntot * sizeof(double), st));
    if (nin) {
      chan_blocks_scatter_kernel<<<(unsigned)nin, 256, 0, st>>>(dPdense.p, ntot, nrad, (const int2*)dIoBlocks.p, dRed.p, 0);
      HFB_CUDA_OK(cudaGetLastError());
    }
    if ((rc = exchange ? exchange_device(dPdense.p, ntot, dKdense.p, ntot, st) : coulomb_device(dPdense.p, ntot, dKdense.p, ntot, st))) return rc;
    HFB_CUDA_OK(cudaStreamSynchronize(st));
    return 0;
  }
  const bool verbose = getenv("HFB200_X_VERBOSE") != nullptr;
  auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  const double t_start = now();
  double t_scan = 0, t_pack = 0, t_dev = 0, t_zero = 0, t_wait = 0;
  // rows of angular function i in the caller's matrix: r0[i] .. r0[i] + cnt[i] - 1  <->  radial index skip[i] ..
  std::vector<int> r0(nang), skip(nang), cnt(nang);
  for (int i = 0; i < nang; i++) {
    if (!map) { r0[i] = i * nrad; skip[i] = 0; cnt[i] = nrad; continue; }
    int a0 = 0;
    while (a0 < nrad && map[i * nrad + a0] < 0) a0++;
    skip[i] = a0; cnt[i] = nrad - a0; r0[i] = a0 < nrad ? map[i * nrad + a0] : 0;
    for (int a = a0; a < nrad; a++)
      if (map[i * nrad + a] != r0[i] + (a - a0)) { set_error("index map: the rows of an angular function must be consecutive"); return HFB200_EINVAL; }
  }
  // ---- scan: non-zero angular blocks of P ----
  std::vector<unsigned char> nz((size_t)npair, 0);
  scan_nonzero_blocks(&P, 1, ldp, nang, r0.data(), cnt.data(), nz.data());
  t_scan = now() - t_start;
  std::vector<int2> inb;
  for (int i = 0; i < nang; i++)
    for (int l = 0; l < nang; l++)
      if (nz[(size_t)i * nang + l]) inb.push_back(make_int2(i, l));
  // output blocks that can be non-zero
  std::vector<int2> outb;
  {
    int mlo = 0, mhi = 0;
    for (int a = 0; a < nang; a++) { mlo = std::min(mlo, ang_m[a]); mhi = std::max(mhi, ang_m[a]); }
    const int span = mhi - mlo;
    std::vector<char> dset(2 * span + 1, 0);
    for (const int2& b : inb) dset[ang_m[b.x] - ang_m[b.y] + span] = 1;
    for (int j = 0; j < nang; j++)
      for (int k = 0; k < nang; k++)
        if (dset[ang_m[j] - ang_m[k] + span]) outb.push_back(make_int2(j, k));
  }
  if ((rc = dPdense.reserve((size_t)ntot * ntot)) || (rc = dKdense.reserve((size_t)ntot * ntot)) ||
      (rc = hX.reserve(std::max<size_t>(std::max(inb.size(), outb.size()), 1) * nrad2)) ||
      (rc = dRed.reserve(std::max<size_t>(std::max(inb.size(), outb.size()), 1) * nrad2)) || (rc = up_list(dIoBlocks, inb)))
    return rc;
  // ---- pack the non-zero blocks (dummy-space layout, zero rows / columns for the removed functions) and upload ----
#pragma omp parallel for schedule(dynamic, 4)
  for (long q = 0; q < (long)inb.size(); q++) {
    const int i = inb[q].x, l = inb[q].y;
    double* dst = hX.p + (size_t)q * nrad2;
    for (int b = 0; b < nrad; b++) {
      double* dc = dst + (size_t)b * nrad;
      if (b < skip[l]) { memset(dc, 0, sizeof(double) * nrad); continue; }
      for (int a = 0; a < skip[i]; a++) dc[a] = 0.0;
      memcpy(dc + skip[i], P + (size_t)r0[i] + (size_t)(r0[l] + b - skip[l]) * ldp, sizeof(double) * cnt[i]);
    }
  }
  HFB_CUDA_OK(cudaMemcpyAsync(dRed.p, hX.p, inb.size() * nrad2 * sizeof(double), cudaMemcpyHostToDevice, st));
  if (driven) {
    if ((rc = dHdr.reserve(nhdr))) return rc;
    hdr_host.assign(nhdr, 0.0);
    hdr_host[0] = (double)inb.size();
    for (size_t q = 0; q < inb.size(); q++) { hdr_host[1 + 2 * q] = inb[q].x; hdr_host[2 + 2 * q] = inb[q].y; }
    HFB_CUDA_OK(cudaMemcpyAsync(dHdr.p, hdr_host.data(), nhdr * sizeof(double), cudaMemcpyHostToDevice, st));
    if ((rc = comm_broadcast(comm, dHdr.p, nhdr, comm_root, st)) || (rc = comm_broadcast(comm, dRed.p, inb.size() * nrad2, comm_root, st))) return rc;
  }
  HFB_CUDA_OK(cudaMemsetAsync(dPdense.p, 0, (size_t)ntot * ntot * sizeof(double), st));
  if (!inb.empty()) {
    chan_blocks_scatter_kernel<<<(unsigned)inb.size(), 256, 0, st>>>(dPdense.p, ntot, nrad, (const int2*)dIoBlocks.p, dRed.p, 0);
    HFB_CUDA_OK(cudaGetLastError());
  }
  // hX is reused for the download: the upload must have left it
  HFB_CUDA_OK(cudaStreamSynchronize(st));
  t_pack = now() - t_start - t_scan;
  if ((rc = exchange ? exchange_device(dPdense.p, ntot, dKdense.p, ntot, st) : coulomb_device(dPdense.p, ntot, dKdense.p, ntot, st))) return rc;
  if ((rc = up_list(dIoBlocks, outb))) return rc;  // (synchronous copy: ordered after the kernels that read the input list)
  if (!outb.empty()) {
    chan_blocks_gather_kernel<<<(unsigned)outb.size(), 256, 0, st>>>(dKdense.p, ntot, nrad, (const int2*)dIoBlocks.p, dRed.p);
    HFB_CUDA_OK(cudaGetLastError());
    HFB_CUDA_OK(cudaMemcpyAsync(hX.p, dRed.p, outb.size() * nrad2 * sizeof(double), cudaMemcpyDeviceToHost, st));
  }
  t_dev = now() - t_start - t_scan - t_pack;  // host side of the operator call (list upkeep, launches)
  // zero-fill the caller's matrix while the device works (the reference returns a zero-filled dense matrix)
  const double t_z0 = now();
  zero_fill(F, ldf, nout);
  t_zero = now() - t_z0;
  HFB_CUDA_OK(cudaStreamSynchronize(st));
  t_wait = now() - t_z0 - t_zero;
#pragma omp parallel for schedule(dynamic, 4)
  for (long q = 0; q < (long)outb.size(); q++) {
    const int j = outb[q].x, k = outb[q].y;
    const double* src = hX.p + (size_t)q * nrad2;
    for (int b = skip[k]; b < nrad; b++)
      memcpy(F + (size_t)r0[j] + (size_t)(r0[k] + b - skip[k]) * ldf, src + (size_t)b * nrad + skip[j], sizeof(double) * cnt[j]);
  }
  if (verbose)
    fprintf(stderr, "[hfb200] %s host call: scan %.1f ms, pack+upload %.1f ms (%zu blocks), launches %.1f ms, zero-fill %.1f ms, wait %.1f ms, total %.1f ms (%zu blocks back)\n",
            exchange ? "exchange" : "coulomb", 1e3 * t_scan, 1e3 * t_pack, inb.size(), 1e3 * t_dev, 1e3 * t_zero, 1e3 * t_wait, 1e3 * (now() - t_start), outb.size());
  return 0;
}

// J and K of one Fock build: the Coulomb kernels (short, bandwidth-bound, 2 % of the build at configs[3] but a fixed cost that does
// not shrink with the number of ranks) run on a second stream of the engine next to the exchange pipeline (tensor-bound); the call
// is ordered on `st` on both sides.  The two operators share no workspace (chan_engine.h); with a communicator their all-reduces
// are issued in the same order on every rank (J first).
int ChanEngine::fock_device(const double* dPm, int ldp, double* dJm, int ldj, double* dKm, int ldk, cudaStream_t st) {
  if (!have_k) { set_error("Primitive teis have not been computed!"); return HFB200_ENOTEI; }
  if (!st2) {
    HFB_CUDA_OK(cudaStreamCreateWithFlags(&st2, cudaStreamNonBlocking));
    HFB_CUDA_OK(cudaEventCreateWithFlags(&evFork, cudaEventDisableTiming));
    HFB_CUDA_OK(cudaEventCreateWithFlags(&evJoin, cudaEventDisableTiming));
  }
  // the exchange pipeline computes the block mask of P first; the Coulomb kernels fork off once it exists and use its exact
  // non-zero flags to skip the zero blocks in their pack pass (1 777 of 80 089 blocks survive for an N2 density)
  bool forked = false;
  on_mask_ready = [&]() -> int {
    HFB_CUDA_OK(cudaEventRecord(evFork, st));
    HFB_CUDA_OK(cudaStreamWaitEvent(st2, evFork, 0));
    const int rc_ = coulomb_device(dPm, ldp, dJm, ldj, st2, dNzX);
    if (rc_) return rc_;
    HFB_CUDA_OK(cudaEventRecord(evJoin, st2));
    forked = true;
    return 0;
  };
  const int rc = exchange_device(dPm, ldp, dKm, ldk, st);
  on_mask_ready = nullptr;
  if (forked) HFB_CUDA_OK(cudaStreamWaitEvent(st, evJoin, 0));
  return rc;
}

int ChanEngine::exchange_device(const double* dPm, int ldpm, double* dKm, int ldkm, cudaStream_t st) {
  if (!have_k) { set_error("Primitive teis have not been computed!"); return HFB200_ENOTEI; }
  const int nel = prm.nel, np2 = prm.np2, nang = xprm.nang, ntot = nang * nrad, stack = prm.stack;
  if ((size_t)ldpm * (size_t)ntot >= ((size_t)1 << 32)) { set_error("exchange: density matrix too large for 32-bit block offsets"); return HFB200_EINVAL; }
  const size_t nang2 = (size_t)nang * nang;
  int rc;
  HFB_CUDA_OK(cudaMemset2DAsync(dKm, (size_t)ldkm * sizeof(double), 0, (size_t)ntot * sizeof(double), ntot, st));
  size_t ev_used = 0;
  std::vector<std::pair<int, size_t>> ev_marks;  // (stage, index of the start event; the end event follows)
  auto tick = [&](int stage, bool begin) {
    if (!profile_stages) return;
    if (ev_used == ev_pool.size()) { cudaEvent_t e; cudaEventCreate(&e); ev_pool.push_back(e); }
    if (begin) ev_marks.push_back({stage, ev_used});
    cudaEventRecord(ev_pool[ev_used++], st);
  };
  // block mask of P and its exact-symmetry flag; the per-pattern lists are rebuilt only when either changes
  static const bool allow_sym = getenv("HFB200_X_NOSYM") == nullptr;
  tick(0, true);
  if (comm && comm_size > 1) {  // every rank scans its share of the blocks; the flags are summed over the ranks
    if ((rc = dHdr.reserve(2 * nang2 + 2))) return rc;
    HFB_CUDA_OK(cudaMemsetAsync(dHdr.p, 0, (2 * nang2 + 1) * sizeof(double), st));
    chan_x0_mask_kernel<<<(unsigned)((nang2 + comm_size - 1) / comm_size), 128, 0, st>>>(dPm, ldpm, nrad, nang, nullptr, nullptr, comm_rank, comm_size, dHdr.p);
    HFB_CUDA_OK(cudaGetLastError());
    if ((rc = comm_allreduce_sum(comm, dHdr.p, 2 * nang2 + 1, st))) return rc;
    chan_x0_flags_kernel<<<(unsigned)((nang2 + 1 + 255) / 256), 256, 0, st>>>(dHdr.p, (int)nang2, dNzP, allow_sym ? 1 : 0, dNzX);
  } else {
    HFB_CUDA_OK(cudaMemsetAsync(dNzP + nang2, allow_sym ? 1 : 0, 1, st));
    chan_x0_mask_kernel<<<(unsigned)nang2, 128, 0, st>>>(dPm, ldpm, nrad, nang, dNzP, allow_sym ? dNzP + nang2 : nullptr, 0, 1, nullptr, dNzX);
  }
  tick(0, false);
  HFB_CUDA_OK(cudaGetLastError());
  if (on_mask_ready) {  // fock_device: the Coulomb part of the same build starts here, with the exact block mask (dNzX) in hand
    if ((rc = on_mask_ready())) return rc;
  }
  HFB_CUDA_OK(cudaMemcpyAsync(hNzP, dNzP, nang2 + 1, cudaMemcpyDeviceToHost, st));
  HFB_CUDA_OK(cudaStreamSynchronize(st));
  if (mask_cached.size() != nang2 + 1 || memcmp(mask_cached.data(), hNzP, nang2 + 1) != 0) {
    if ((rc = build_items(hNzP, hNzP[nang2] != 0))) return rc;
    mask_cached.assign(hNzP, hNzP + nang2 + 1);
  }
  last_sym = hNzP[nang2] != 0;
  last_items = total_items;
  if (total_items > 0 || comm) {
    if ((rc = dKblocks.reserve((size_t)max_batch_items * nel * nel * np2)) ||
        (rc = dW.reserve((size_t)std::max<long long>(max_wblocks, 1) * stack * xprm.WB)))
      return rc;
    if (x1tab_ldp != ldpm) {  // W index (element pair, i, l) -> offset inside an angular block of the dense P
      std::vector<unsigned> tab((size_t)xprm.WB, 0xffffffffu);
      const int np = prm.np;
      for (int ie = 0; ie < nel; ie++)
        for (int je = 0; je < nel; je++)
          for (int l = 0; l <= last[je] - first[je]; l++)
            for (int i = 0; i <= last[ie] - first[ie]; i++)
              tab[(size_t)(ie * nel + je) * xprm.PB + i + l * np] = (unsigned)(first[ie] + i) + (unsigned)(first[je] + l) * (unsigned)ldpm;
      if ((rc = up_list(dX1tab, tab))) return rc;
      x1tab_ldp = ldpm;
    }
    ChanXParams q = xprm;
    q.P = dPm; q.ldp = ldpm; q.nzP = dNzP; q.W = dW.p; q.blocks = dKblocks.p; q.K = dKm; q.ldk = ldkm;
    q.x1tab = (const unsigned*)dX1tab.p;
    q.fam_c0 = prm.fam_c0; q.fam_fac = dFamFac; q.items = (const int2*)dItems.p; q.item_fam = (const int*)dItemFam.p;
    q.wbase = (const long long*)dWbase.p; q.wbits = (const unsigned long long*)dWbits.p; q.nw = (nang + 63) / 64;
    q.tile_u = (const int2*)dTileU.p; q.slot_wblk = (const long long*)dSlotWblk.p; q.slot_coef = (const double*)dSlotCoef.p;
    q.sym = last_sym ? 1 : 0;
    const int ntl = (prm.np + 7) / 8;
    if (((size_t)ni_same * q.LDK + chxf_inline_ring_doubles(q.PB)) * sizeof(double) > 227 * 1024) { set_error("exchange: element block too large for the shared-memory tile"); return HFB200_EINVAL; }
    const size_t smem1 = (size_t)2 * (CHX_UMAX + 1) * CHX_LDC * sizeof(double) + 8 * CHX_T * sizeof(long long) +
                         (size_t)(2 * nang + 16 * (CHX_UMAX + 4)) * sizeof(int) + (size_t)2 * nang;
    if (smem1 > 227 * 1024) { set_error("exchange: angular basis too large for the shared-memory coupling tables"); return HFB200_EINVAL; }
    HFB_CUDA_OK(cudaFuncSetAttribute(chan_x1_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(smem1, 48 * 1024)));
    const int2* tiles = (const int2*)dTiles2.p;
    int nsm = 148, dev = 0;
    HFB_CUDA_OK(cudaGetDevice(&dev));
    HFB_CUDA_OK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
    for (const XBatch& B : batches) {
      q.item0 = B.item0;
      const int nt1 = B.t1_1 - B.t1_0, ngs = B.gs1 - B.gs0, ngp = B.gp1 - B.gp0;
      if (nt1 > 0 && B.wblocks > 0) {
        tick(1, true);
        chan_x1_mma_kernel<<<dim3(nt1, std::min(nang, 7)), 256, smem1, st>>>(q, dTiles1 + (size_t)B.t1_0 * (2 + CHX_T));
        HFB_CUDA_OK(cudaGetLastError());
        tick(1, false);
      }
      if (ngs == 0) continue;
      tick(2, true);
      const int4* gS = (const int4*)dGroupsS.p + B.gs0;
      if (ni_same == 32) rc = launch_xf_same<32>(q, gS, ngs, tiles, nsm, st);
      else if (ni_same == 16) rc = launch_xf_same<16>(q, gS, ngs, tiles, nsm, st);
      else rc = launch_xf_same<8>(q, gS, ngs, tiles, nsm, st);
      if (rc) return rc;
      tick(2, false);
      tick(3, true);
      const int4* gP = (const int4*)dGroupsP.p + B.gp0;
      switch (ntl) {
        case 1: rc = launch_xf_pair<8, 1>(q, gP, ngp, tiles, nsm, st); break;
        case 2: rc = launch_xf_pair<8, 2>(q, gP, ngp, tiles, nsm, st); break;
        case 3: rc = launch_xf_pair<4, 3>(q, gP, ngp, tiles, nsm, st); break;
        default: rc = launch_xf_pair<4, 4>(q, gP, ngp, tiles, nsm, st); break;
      }
      if (rc) return rc;
      tick(3, false);
      tick(4, true);
      {
        const size_t smem4 = (size_t)nrad * (nrad + 1) * sizeof(double);
        static const bool x4_tiled = getenv("HFB200_X4_TILED") != nullptr;
        if (!x4_tiled && prm.np2 <= 256 && smem4 <= 100 * 1024) {
          HFB_CUDA_OK(cudaFuncSetAttribute(chan_x4_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(smem4, 48 * 1024)));
          const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (size_t)220 * 1024 / (smem4 + 1024)));
          chan_x4_smem_kernel<<<(unsigned)std::min(B.nout, nsm * per_sm), 256, smem4, st>>>(q, (const int*)dOutPtr.p + B.out0, (const int*)dOutItem.p, B.nout);
        } else {
          chan_x4_kernel<<<(unsigned)std::min(B.nout, 148 * 8), 256, 0, st>>>(q, (const int*)dOutPtr.p + B.out0, (const int*)dOutItem.p, B.nout);
        }
      }
      HFB_CUDA_OK(cudaGetLastError());
      tick(4, false);
    }
    if (comm && !red_blocks.empty()) {  // sum the blocks of K that can be non-zero over the ranks, packed; with a symmetric P
      // only j <= k travels and the scatter writes both triangles
      const size_t nb = red_blocks.size();
      if ((rc = dRed.reserve(nb * nrad * nrad))) return rc;
      tick(5, true);
      chan_blocks_gather_kernel<<<(unsigned)nb, 256, 0, st>>>(dKm, ldkm, nrad, (const int2*)dRedBlocks.p, dRed.p);
      HFB_CUDA_OK(cudaGetLastError());
      if ((rc = comm_allreduce_sum(comm, dRed.p, nb * nrad * nrad, st))) return rc;
      chan_blocks_scatter_kernel<<<(unsigned)nb, 256, 0, st>>>(dKm, ldkm, nrad, (const int2*)dRedBlocks.p, dRed.p, last_sym ? 1 : 0);
      HFB_CUDA_OK(cudaGetLastError());
      tick(5, false);
    }
    if (profile_stages) {
      HFB_CUDA_OK(cudaStreamSynchronize(st));
      for (float& v : stage_ms) v = 0.f;
      for (const auto& mk : ev_marks) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, ev_pool[mk.second], ev_pool[mk.second + 1]);
        stage_ms[mk.first] += ms;
      }
    }
  }
  // algorithmic flops of the radial contraction per output block and family: same-element products 2 np_e^4 and two
  // (np x np x np) products per ordered pair of different elements, per stack combination.  The count is that of the
  // FULL item list (both (j,k) and (k,j)): with a symmetric P the engine computes half of them and mirrors the rest,
  // so last_items is the number of blocks actually contracted and last_x3_flops the work actually executed.
  {
    double per = 0.0, per_same = 0.0;
    for (int a = 0; a < nel; a++)
      for (int b = 0; b < nel; b++) {
        const double na = last[a] - first[a] + 1, nb = last[b] - first[b] + 1;
        per += (a == b) ? 2.0 * na * na * na * na : 2.0 * na * na * nb + 2.0 * na * nb * nb;
        if (a == b) per_same += 2.0 * na * na * na * na;
      }
    last_x3_flops = per * stack * stack * (double)last_items;
    last_same_flops = per_same * stack * stack * (double)last_items;
  }
  return 0;
}

void ChanEngine::release() {
  for (void* p : owned) cudaFree(p);
  owned.clear();
  for (DevBuf* b : {&dP, &dPc, &dTr, &dJc, &dJ, &dPdense, &dKdense, &dKblocks, &dW, &dRed, &dHdr}) b->release();
  if (st2) { cudaStreamDestroy(st2); cudaEventDestroy(evFork); cudaEventDestroy(evJoin); st2 = nullptr; evFork = evJoin = nullptr; }
  for (DevList* L : {&dItems, &dItemFam, &dTiles2, &dGroupsS, &dGroupsP, &dTileU, &dSlotWblk, &dSlotCoef, &dWbase, &dWbits, &dOutPtr, &dOutItem, &dX1tab, &dRedBlocks, &dIoBlocks}) {
    if (L->p) cudaFree(L->p);
    L->p = nullptr;
    L->cap = 0;
  }
  if (hNzP) cudaFreeHost(hNzP);
  hNzP = nullptr;
  for (cudaEvent_t e : ev_pool) cudaEventDestroy(e);
  ev_pool.clear();
  hIn.release();
  hOut.release();
  hX.release();
}

}  // namespace hfb

// ---- C ABI: device-resident operator calls of the atomic / diatomic contexts (include/helfem_b200.h) ----
extern "C" {
int hfb200_chan_dim(hfb200_ctx* c, int* ndim) {
  if (!c || !c->chan || !ndim) { hfb::set_error("hfb200_chan_dim: no channel engine (call set_tei first)"); return HFB200_EINVAL; }
  *ndim = (int)c->lval.size() * c->chan->nrad;
  return HFB200_OK;
}
int hfb200_chan_coulomb_device(hfb200_ctx* c, const double* dP, int ldp, double* dJ, int ldj, void* stream) {
  if (!c || !c->chan || !dP || !dJ) { hfb::set_error("hfb200_chan_coulomb_device: invalid argument"); return HFB200_EINVAL; }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  return c->chan->coulomb_device(dP, ldp, dJ, ldj, stream ? (cudaStream_t)stream : c->stream);
}
int hfb200_chan_exchange_device(hfb200_ctx* c, const double* dP, int ldp, double* dK, int ldk, void* stream) {
  if (!c || !c->chan || !dP || !dK) { hfb::set_error("hfb200_chan_exchange_device: invalid argument"); return HFB200_EINVAL; }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  return c->chan->exchange_device(dP, ldp, dK, ldk, stream ? (cudaStream_t)stream : c->stream);
}
int hfb200_chan_fock_device(hfb200_ctx* c, const double* dP, int ldp, double* dJ, int ldj, double* dK, int ldk, void* stream) {
  if (!c || !c->chan || !dP || !dJ || !dK) { hfb::set_error("hfb200_chan_fock_device: invalid argument"); return HFB200_EINVAL; }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  return c->chan->fock_device(dP, ldp, dJ, ldj, dK, ldk, stream ? (cudaStream_t)stream : c->stream);
}
int hfb200_chan_exchange_stats(hfb200_ctx* c, long long* items, double* x3_flops, double* stage_ms) {
  if (!c || !c->chan) { hfb::set_error("hfb200_chan_exchange_stats: invalid argument"); return HFB200_EINVAL; }
  if (items) *items = c->chan->last_items;
  if (x3_flops) *x3_flops = c->chan->last_x3_flops;
  if (stage_ms)
    for (int q = 0; q < 5; q++) stage_ms[q] = c->chan->stage_ms[q];
  return HFB200_OK;
}
int hfb200_chan_exchange_counts(hfb200_ctx* c, long long* out, int n) {
  if (!c || !c->chan || !out || n < 1) { hfb::set_error("hfb200_chan_exchange_counts: invalid argument"); return HFB200_EINVAL; }
  for (int q = 0; q < n; q++) out[q] = q < 12 ? c->chan->counts[q] : 0;
  return HFB200_OK;
}
int hfb200_comm_set_root(hfb200_ctx* c, int root) {
  if (!c || !c->chan || !c->comm || root >= c->shard_n) { hfb::set_error("hfb200_comm_set_root: needs a context with a communicator (hfb200_comm_init) and root < nranks"); return HFB200_EINVAL; }
  c->chan->comm_root = root < 0 ? -1 : root;
  return HFB200_OK;
}
int hfb200_chan_serve(hfb200_ctx* c, int exchange) {
  if (!c || !c->chan || !c->comm) { hfb::set_error("hfb200_chan_serve: needs a context with a communicator"); return HFB200_EINVAL; }
  if (c->chan->comm_root < 0 || c->chan->comm_rank == c->chan->comm_root) { hfb::set_error("hfb200_chan_serve: only the workers of a root-driven communicator (hfb200_comm_set_root) serve"); return HFB200_EINVAL; }
  if (cudaSetDevice(c->device) != cudaSuccess) { hfb::set_error("cudaSetDevice failed"); return HFB200_ECUDA; }
  return c->chan->dense_apply(exchange != 0, nullptr, 0, nullptr, 0, c->stream);
}
int hfb200_chan_allreduce_ms(hfb200_ctx* c, double* ms) {
  if (!c || !c->chan || !ms) { hfb::set_error("hfb200_chan_allreduce_ms: invalid argument"); return HFB200_EINVAL; }
  *ms = c->chan->stage_ms[5];
  return HFB200_OK;
}
int hfb200_chan_exchange_batches(hfb200_ctx* c, int* nbatches) {
  if (!c || !c->chan || !nbatches) { hfb::set_error("hfb200_chan_exchange_batches: invalid argument"); return HFB200_EINVAL; }
  *nbatches = c->chan->nbatches();
  return HFB200_OK;
}
int hfb200_chan_set_profile(hfb200_ctx* c, int on) {
  if (!c || !c->chan) { hfb::set_error("hfb200_chan_set_profile: invalid argument"); return HFB200_EINVAL; }
  c->chan->profile_stages = on != 0;
  return HFB200_OK;
}
}

#ifdef HFB_XF_TRACE
// debug only: copies the phase trace of the fused same-element kernel (chan_exchange_fused.cuh) to the host
extern "C" int hfb200_debug_trace(unsigned long long* out, int* counts) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, hfb::g_trace, sizeof(unsigned long long) * 148 * 4096);
  cudaMemcpyFromSymbol(counts, hfb::g_trace_n, sizeof(int) * 148);
  return 0;
}
#endif
