// Host-side driver of the channel-coupled Coulomb / exchange kernels; see chan_engine.h.
#include "chan_engine.h"

#include <algorithm>
#include <cstring>
#include <map>
#include "../../include/helfem_b200.h"
#include "chan_exchange.cuh"

namespace hfb {

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
      for (const auto& cp : S.pair_coupl[pr]) by_chan[cp.chan].push_back({pr, &cp});
    std::vector<int> pre_ptr(nchan + 1, 0), pre_pair, post_ptr(npair + 1, 0), post_chan;
    std::vector<double> pre_coef, post_coef;
    for (int c = 0; c < nchan; c++) {
      for (auto& en : by_chan[c]) {
        pre_pair.push_back(en.first);
        pre_coef.push_back(en.second->coef[0]);
        pre_coef.push_back(en.second->coef[1]);
      }
      pre_ptr[c + 1] = (int)pre_pair.size();
    }
    // bra contraction of block (i,j) uses the couplings of the pair (j,i)
    for (int i = 0; i < nang; i++)
      for (int j = 0; j < nang; j++) {
        for (const auto& cp : S.pair_coupl[j * nang + i]) {
          post_chan.push_back(cp.chan);
          post_coef.push_back(cp.coef[0]);
          post_coef.push_back(cp.coef[1]);
        }
        post_ptr[i * nang + j + 1] = (int)post_chan.size();
      }
    if ((rc = up(this, &prm.chan_r, S.chan_r)) || (rc = up(this, &prm.fam_c0, fam_c0)) || (rc = up(this, &prm.pre_ptr, pre_ptr)) ||
        (rc = up(this, &prm.pre_pair, pre_pair)) || (rc = up(this, &prm.pre_coef, pre_coef)) || (rc = up(this, &prm.post_ptr, post_ptr)) ||
        (rc = up(this, &prm.post_chan, post_chan)) || (rc = up(this, &prm.post_coef, post_coef)))
      return rc;
    fam_c0_host = fam_c0;
  }
  have_k = false;
  if (with_exchange && S.ktei) return build_exchange(S);
  return 0;
}

// Exchange tables: K-ordered in-element integrals, full-square disjoint factors, and per output angular
// pair (j,k) the list of (input pair (i,l), family, coefficient[a][b]) with
//   coef_ab = fac_r * (-1)^{a != b} * c^a_c(j,i) * c^b_c(k,l),   c = channel shared by both pairs
// (atomic/TwoDBasis.cpp:1023-1059, diatomic/basis.cpp:1604-1650).
int ChanEngine::build_exchange(const ChanSpec& S) {
  const int nel = S.nel, nang = S.nang, st = S.stack, nr = S.nr, np = prm.np, np2 = prm.np2;
  xprm.nel = nel; xprm.np = np; xprm.np2 = np2; xprm.stack = st; xprm.nrad = nrad; xprm.nang = nang;
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
  // K-ordered integrals, zero padded, one (family, element, a, b) block at a time
  {
    const size_t blk = (size_t)np2 * np2;
    double* d = nullptr;
    HFB_CUDA_OK(cudaMalloc(&d, (size_t)nr * nel * st * st * blk * sizeof(double)));
    owned.push_back(d);
    xprm.ktei = d;
    std::vector<double> stage(blk);
    for (int r = 0; r < nr; r++)
      for (int e = 0; e < nel; e++)
        for (int a = 0; a < st; a++)
          for (int b = 0; b < st; b++) {
            double sg = 1.0;
            const double* t = S.ktei(r, e, a, b, &sg);
            double* dst = d + ((((size_t)r * nel + e) * st + a) * st + b) * blk;
            if (!t) { HFB_CUDA_OK(cudaMemset(dst, 0, blk * sizeof(double))); continue; }
            const int n = npe[e];
            if (n == np) {
              HFB_CUDA_OK(cudaMemcpy(dst, t, blk * sizeof(double), cudaMemcpyHostToDevice));
            } else {
              std::fill(stage.begin(), stage.end(), 0.0);
              const size_t n2 = (size_t)n * n;
              for (int cs = 0; cs < n; cs++)
                for (int cf = 0; cf < n; cf++)
                  for (int rs = 0; rs < n; rs++)
                    for (int rf = 0; rf < n; rf++)
                      stage[(size_t)(rf + rs * np) + (size_t)(cf + cs * np) * np2] = t[(size_t)(rf + rs * n) + (size_t)(cf + cs * n) * n2];
              HFB_CUDA_OK(cudaMemcpy(dst, stage.data(), blk * sizeof(double), cudaMemcpyHostToDevice));
            }
          }
  }
  // coupling lists
  {
    struct Ent { int in, r; double c[4]; };
    std::vector<int> xptr(1, 0), xin, xr, out_list;
    std::vector<double> xcoef;
    for (int j = 0; j < nang; j++)
      for (int k = 0; k < nang; k++) {
        std::vector<Ent> ents;
        for (int i = 0; i < nang; i++) {
          const auto& cji = S.pair_coupl[j * nang + i];
          if (cji.empty()) continue;
          for (int l = 0; l < nang; l++) {
            const auto& ckl = S.pair_coupl[k * nang + l];
            for (const auto& A : cji)
              for (const auto& B : ckl) {
                if (A.chan != B.chan) continue;
                const int r = S.chan_r[A.chan];
                Ent en;
                en.in = i * nang + l;
                en.r = r;
                bool any = false;
                for (int a = 0; a < 2; a++)
                  for (int b = 0; b < 2; b++) {
                    double v = 0.0;
                    if (a < st && b < st) v = S.fac[r] * (a != b ? -1.0 : 1.0) * A.coef[a] * B.coef[b];
                    en.c[2 * a + b] = v;
                    any = any || v != 0.0;
                  }
                if (any) ents.push_back(en);
              }
          }
        }
        if (ents.empty()) continue;
        std::stable_sort(ents.begin(), ents.end(), [](const Ent& x, const Ent& y) { return x.r < y.r; });
        out_list.push_back(j * nang + k);
        for (const auto& en : ents) {
          xin.push_back(en.in);
          xr.push_back(en.r);
          // kernel reads coefficient (a,b) at a*stack + b
          double cc[4] = {0, 0, 0, 0};
          for (int a = 0; a < st; a++)
            for (int b = 0; b < st; b++) cc[a * st + b] = en.c[2 * a + b];
          xcoef.insert(xcoef.end(), cc, cc + 4);
        }
        xptr.push_back((int)xin.size());
      }
    nout = (int)out_list.size();
    out_pairs_host = out_list;
    if ((rc = up(this, &xprm.xptr, xptr)) || (rc = up(this, &xprm.xin, xin)) || (rc = up(this, &xprm.xr, xr)) ||
        (rc = up(this, &xprm.xcoef, xcoef)) || (rc = up(this, &xprm.out_list, out_list)))
      return rc;
  }
  have_k = true;
  return 0;
}

// J = coulomb(P) with dense host matrices ((nang*nrad)^2, column-major, block (iang,jang) at rows
// iang*nrad, cols jang*nrad: atomic/TwoDBasis.cpp:290-300).  Only the element-diagonal radial blocks of
// each angular pair enter or leave the operator, so that is all that crosses PCIe.
int ChanEngine::coulomb_dense(const double* P, int ldp, double* J, int ldj, cudaStream_t st) {
  const int nel = prm.nel, np = prm.np, np2 = prm.np2, nang = (int)std::lround(std::sqrt((double)npair));
  const size_t nband = (size_t)npair * nel * np2;
  const int KP = 4 * prm.KS, MP = 8 * prm.MT;
  int rc;
  if ((rc = hIn.reserve(nband)) || (rc = hOut.reserve(nband)) || (rc = dP.reserve(nband)) || (rc = dJ.reserve(nband)) ||
      (rc = dPc.reserve((size_t)nchan * nel * KP)) || (rc = dTr.reserve((size_t)nchan * nel * 2)) ||
      (rc = dJc.reserve((size_t)nchan * nel * MP)))
    return rc;
  // pack: [pair = k*nang + l][element][np2]
#pragma omp parallel for collapse(2) schedule(static)
  for (int k = 0; k < nang; k++)
    for (int l = 0; l < nang; l++) {
      double* dst = hIn.p + ((size_t)(k * nang + l) * nel) * np2;
      for (int e = 0; e < nel; e++) {
        const int f = first[e], npe = last[e] - f + 1;
        double* blk = dst + (size_t)e * np2;
        if (npe < np) memset(blk, 0, sizeof(double) * np2);
        for (int j = 0; j < npe; j++)
          memcpy(blk + (size_t)j * np, P + (size_t)(k * nrad + f) + (size_t)(l * nrad + f + j) * ldp, sizeof(double) * npe);
      }
    }
  HFB_CUDA_OK(cudaMemcpyAsync(dP.p, hIn.p, nband * sizeof(double), cudaMemcpyHostToDevice, st));
  ChanParams q = prm;
  q.nb = 1; q.P = dP.p; q.pc = dPc.p; q.traces = dTr.p; q.jc = dJc.p; q.J = dJ.p;
  chan_pre_kernel<<<dim3(nel, nchan, 1), 128, 0, st>>>(q);
  HFB_CUDA_OK(cudaGetLastError());
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
  chan_post_kernel<<<dim3(nel, npair, 1), 128, 0, st>>>(q);
  HFB_CUDA_OK(cudaGetLastError());
  HFB_CUDA_OK(cudaMemcpyAsync(hOut.p, dJ.p, nband * sizeof(double), cudaMemcpyDeviceToHost, st));
  const int ntot = nang * nrad;
#pragma omp parallel for schedule(static)
  for (int c = 0; c < ntot; c++) memset(J + (size_t)c * ldj, 0, sizeof(double) * ntot);
  HFB_CUDA_OK(cudaStreamSynchronize(st));
#pragma omp parallel for collapse(2) schedule(static)
  for (int i = 0; i < nang; i++)
    for (int j = 0; j < nang; j++) {
      const double* src = hOut.p + ((size_t)(i * nang + j) * nel) * np2;
      for (int e = 0; e < nel; e++) {
        const int f = first[e], npe = last[e] - f + 1;
        for (int b = 0; b < npe; b++)
          for (int a = 0; a < npe; a++)
            J[(size_t)(i * nrad + f + a) + (size_t)(j * nrad + f + b) * ldj] += src[(size_t)e * np2 + a + (size_t)b * np];
      }
    }
  return 0;
}

// -K = exchange(P), dense host matrices.
int ChanEngine::exchange_dense(const double* P, int ldp, double* K, int ldk, cudaStream_t st) {
  if (!have_k) { set_error("Primitive teis have not been computed!"); return HFB200_ENOTEI; }
  const int nel = prm.nel, np2 = prm.np2, nang = xprm.nang, ntot = nang * nrad;
  int rc;
  if ((rc = dPdense.reserve((size_t)ntot * ntot)) || (rc = dKdense.reserve((size_t)ntot * ntot)) ||
      (rc = dKblocks.reserve((size_t)std::max(nout, 1) * nel * nel * np2)))
    return rc;
  HFB_CUDA_OK(cudaMemcpy2DAsync(dPdense.p, ntot * sizeof(double), P, ldp * sizeof(double), ntot * sizeof(double), ntot,
                                cudaMemcpyHostToDevice, st));
  HFB_CUDA_OK(cudaMemsetAsync(dKdense.p, 0, (size_t)ntot * ntot * sizeof(double), st));
  if (nout > 0) {
    ChanXParams q = xprm;
    q.P = dPdense.p; q.ldp = ntot; q.blocks = dKblocks.p;
    const size_t smem = (size_t)(prm.stack * prm.stack + 2) * np2 * sizeof(double);
    HFB_CUDA_OK(cudaFuncSetAttribute(chan_exchange_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    chan_exchange_block_kernel<<<dim3(nel * nel, nout), 256, smem, st>>>(q);
    HFB_CUDA_OK(cudaGetLastError());
    chan_exchange_assemble_kernel<<<148 * 8, 256, 0, st>>>(q, nout, dKdense.p, ntot);
    HFB_CUDA_OK(cudaGetLastError());
  }
  HFB_CUDA_OK(cudaMemcpy2DAsync(K, ldk * sizeof(double), dKdense.p, ntot * sizeof(double), ntot * sizeof(double), ntot,
                                cudaMemcpyDeviceToHost, st));
  HFB_CUDA_OK(cudaStreamSynchronize(st));
  return 0;
}

void ChanEngine::release() {
  for (void* p : owned) cudaFree(p);
  owned.clear();
  for (DevBuf* b : {&dP, &dPc, &dTr, &dJc, &dJ, &dPdense, &dKdense, &dKblocks}) b->release();
  hIn.release();
  hOut.release();
}

}  // namespace hfb
