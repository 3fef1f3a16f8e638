// C ABI, spherically averaged atom: context creation, integral upload, and the operators
// coulomb / eval_Fxc in dense-host, band-host and batched-device form (exchange lives in
// capi_sadatom_exchange.cu).
#include <cstring>
#include "../../include/helfem_b200.h"
#include "context.h"
#include "sadatom_coulomb_mma.cuh"
#include "sadatom_grid_mma.cuh"

using namespace hfb;

namespace hfb {  // sadatom_grid_{lda,gga,mgga}.cu
int launch_grid_lda(int nspin, const SadGridParams& prm, cudaStream_t st);
int launch_grid_gga(int nspin, const SadGridParams& prm, cudaStream_t st);
int launch_grid_mgga(int nspin, const SadGridParams& prm, cudaStream_t st);
}  // namespace hfb

namespace {

template <class T>
int upload(T** dst, const std::vector<T>& src) {
  HFB_CUDA_OK(cudaMalloc(dst, src.size() * sizeof(T)));
  HFB_CUDA_OK(cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}

bool valid_pair(int x_func, int c_func, int* fam) {
  const int fx = xc::xc_family(x_func), fc = xc::xc_family(c_func);
  if (x_func != 0 && (fx == 0 || !xc::xc_is_exchange(x_func))) return false;
  if (c_func != 0 && (fc == 0 || xc::xc_is_exchange(c_func))) return false;
  *fam = fx > fc ? fx : fc;
  return *fam != 0;
}

int launch_grid(int fam, int nspin, const SadGridParams& prm, cudaStream_t st) {
  if (fam == xc::FAM_LDA) return launch_grid_lda(nspin, prm, st);
  if (fam == xc::FAM_GGA) return launch_grid_gga(nspin, prm, st);
  return launch_grid_mgga(nspin, prm, st);
}

int launch_coulomb(const SadJParams& q, bool pack, cudaStream_t st) {
  if (pack) {  // otherwise the grid kernel of the same build has already written pc / traces
    sad_jpack_kernel<<<dim3(q.nel, (q.nb + 3) / 4), 128, 0, st>>>(q);
    HFB_CUDA_OK(cudaGetLastError());
  }
  const int rts = (q.MT + 3) / 4;
  if (q.nb <= 8) {
    sad_jmma_kernel<true, 1><<<dim3(q.nel * rts, 1), 256, 0, st>>>(q);
  } else if (q.nb <= 16) {
    sad_jmma_kernel<true, 2><<<dim3(q.nel * rts, 1), 256, 0, st>>>(q);
  } else {
#ifndef HFB_JMMA_NTW
#define HFB_JMMA_NTW 2
#endif
    const size_t smem = (size_t)4 * q.KS * 32 * sizeof(double);
    if (smem > 48 * 1024)
      HFB_CUDA_OK(cudaFuncSetAttribute(sad_jmma_smem_kernel<HFB_JMMA_NTW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    sad_jmma_smem_kernel<HFB_JMMA_NTW><<<dim3(q.nel * rts, (q.nb + 64 * HFB_JMMA_NTW - 1) / (64 * HFB_JMMA_NTW)), 256, smem, st>>>(q);
  }
  HFB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace

extern "C" {

int hfb200_sadatom_create(hfb200_ctx** out, int device, int lmax, int nel, int nq, const int* first,
                          const int* last, const double* r, const double* wrad, const double* bf,
                          const double* df, const double* lf) {
  (void)lf;  // no supported functional needs the Laplacian (reference: do_lapl arms)
  if (!out || nel < 1 || nq < 1 || !first || !last || !r || !wrad || !bf || !df || lmax < 0) {
    set_error("hfb200_sadatom_create: invalid argument");
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
  c->kind = CTX_SADATOM;
  c->device = device;
  c->lmax = lmax;
  c->nel = nel;
  c->nq = nq;
  c->nq8 = (nq + 7) & ~7;
  c->ldq = sad_grid_ldq(nq);  // = 4 (mod 16): conflict-free DMMA fragment loads (sadatom_grid_mma.cuh)
  c->first.assign(first, first + nel);
  c->last.assign(last, last + nel);
  c->np_max = 0;
  for (int e = 0; e < nel; e++) {
    const int np = last[e] - first[e] + 1;
    if (np < 1) { delete c; set_error("element %d has no functions", e); return HFB200_EINVAL; }
    if (np > c->np_max) c->np_max = np;
  }
  c->nrad = last[nel - 1] + 1;
  c->np_pad = (c->np_max + 7) & ~7;
  // tables: host (nq x np_e) column-major per element -> device [el][i][ldq], zero padded
  const int np = c->np_pad, ldq = c->ldq;
  std::vector<double> B((size_t)nel * np * ldq, 0.0), dB((size_t)nel * np * ldq, 0.0);
  size_t off = 0;
  for (int e = 0; e < nel; e++) {
    const int npe = last[e] - first[e] + 1;
    for (int i = 0; i < npe; i++)
      for (int p = 0; p < nq; p++) {
        B[((size_t)e * np + i) * ldq + p] = bf[off + (size_t)i * nq + p];
        dB[((size_t)e * np + i) * ldq + p] = df[off + (size_t)i * nq + p];
      }
    off += (size_t)npe * nq;
  }
  int rc;
  if ((rc = upload(&c->dBt, B)) || (rc = upload(&c->ddBt, dB)) ||
      (rc = upload(&c->dR, std::vector<double>(r, r + (size_t)nel * nq))) ||
      (rc = upload(&c->dW, std::vector<double>(wrad, wrad + (size_t)nel * nq)))) {
    hfb200_destroy(c);
    return rc;
  }
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

int hfb200_sadatom_set_tei(hfb200_ctx* c, int nL, const double* const* disjoint_L,
                           const double* const* disjoint_m1L, const double* const* prim_tei,
                           const double* const* prim_ktei) {
  if (!c || c->kind != CTX_SADATOM || nL < 1 || !disjoint_L || !disjoint_m1L || !prim_tei) {
    set_error("hfb200_sadatom_set_tei: invalid argument");
    return HFB200_EINVAL;
  }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  const int nel = c->nel, np = c->np_max, np2 = np * np;
  for (double** p : {&c->dDisL, &c->dDisM1L, &c->dTeiF, &c->dKtei, &c->dLc, &c->dm1Lc})
    if (*p) { cudaFree(*p); *p = nullptr; }
  if (c->dTri) { cudaFree(c->dTri); c->dTri = nullptr; }
  int rc;
  // ---- disjoint factors, all L, zero padded to np_max (exchange path) ----
  {
    std::vector<double> dl((size_t)nL * nel * np2, 0.0), dm((size_t)nL * nel * np2, 0.0);
    for (int L = 0; L < nL; L++)
      for (int e = 0; e < nel; e++) {
        const int npe = c->last[e] - c->first[e] + 1;
        const double* a = disjoint_L[L * nel + e];
        const double* bb = disjoint_m1L[L * nel + e];
        for (int j = 0; j < npe; j++)
          for (int i = 0; i < npe; i++) {
            dl[((size_t)L * nel + e) * np2 + i + j * np] = a[i + j * npe];
            dm[((size_t)L * nel + e) * np2 + i + j * np] = bb[i + j * npe];
          }
      }
    if ((rc = upload(&c->dDisL, dl)) || (rc = upload(&c->dDisM1L, dm))) return rc;
  }
  // ---- Coulomb path (L = 0 only, sadatom/basis.cpp:306): triangle-packed, fragment-ordered ----
  {
    const int ntri = np * (np + 1) / 2;
    const int MP = (ntri + 7) & ~7, KP = MP, MT = MP / 8, KS = KP / 4;  // KS even (16-byte operand loads)
    c->ntri = ntri; c->jMT = MT; c->jKS = KS;
    std::vector<int2> tri(ntri);
    for (int j = 0; j < np; j++)
      for (int i = 0; i <= j; i++) tri[i + j * (j + 1) / 2] = make_int2(i, j);
    std::vector<double> dlc((size_t)nel * MP, 0.0), dmc((size_t)nel * MP, 0.0);
    std::vector<double> F((size_t)nel * MT * KS * 32, 0.0);
    for (int e = 0; e < nel; e++) {
      const int npe = c->last[e] - c->first[e] + 1;
      const size_t n2 = (size_t)npe * npe;
      const double* t = prim_tei[e];          // L = 0
      const double* a = disjoint_L[e];
      const double* bb = disjoint_m1L[e];
      for (int m = 0; m < ntri; m++) {
        const int i = tri[m].x, j = tri[m].y;
        if (j >= npe) continue;
        dlc[(size_t)e * MP + m] = a[i + j * npe];
        dmc[(size_t)e * MP + m] = bb[i + j * npe];
        for (int kk = 0; kk < ntri; kk++) {
          const int k = tri[kk].x, l = tri[kk].y;
          if (l >= npe) continue;
          const double v = t[(size_t)(i + j * npe) + (size_t)(k + l * npe) * n2];
          const int mt = m >> 3, ks = kk % KS, lane = (m & 7) * 4 + kk / KS;  // column kk -> k-step kk % KS, slot kk / KS
          F[(((size_t)e * MT + mt) * KS + ks) * 32 + lane] = v;
        }
      }
    }
    if ((rc = upload(&c->dTeiF, F)) || (rc = upload(&c->dLc, dlc)) || (rc = upload(&c->dm1Lc, dmc)) ||
        (rc = upload(&c->dTri, tri)))
      return rc;
  }
  // ---- K-ordered in-element blocks, all L, zero padded (exchange path) ----
  if (prim_ktei) {
    const size_t blk = (size_t)np2 * np2;
    HFB_CUDA_OK(cudaMalloc(&c->dKtei, (size_t)nL * nel * blk * sizeof(double)));
    std::vector<double> stage(blk);
    for (int L = 0; L < nL; L++)
      for (int e = 0; e < nel; e++) {
        const int npe = c->last[e] - c->first[e] + 1;
        const double* t = prim_ktei[L * nel + e];
        double* dst = c->dKtei + ((size_t)L * nel + e) * blk;
        if (npe == np) {
          HFB_CUDA_OK(cudaMemcpy(dst, t, blk * sizeof(double), cudaMemcpyHostToDevice));
        } else {
          std::fill(stage.begin(), stage.end(), 0.0);
          const size_t n2 = (size_t)npe * npe;
          // both index pairs are (fast + slow*npe) -> (fast + slow*np)
          for (int cs = 0; cs < npe; cs++)
            for (int cf = 0; cf < npe; cf++)
              for (int rs = 0; rs < npe; rs++)
                for (int rf = 0; rf < npe; rf++)
                  stage[(size_t)(rf + rs * np) + (size_t)(cf + cs * np) * np2] = t[(size_t)(rf + rs * npe) + (size_t)(cf + cs * npe) * n2];
          HFB_CUDA_OK(cudaMemcpy(dst, stage.data(), blk * sizeof(double), cudaMemcpyHostToDevice));
        }
      }
  }
  c->nL = nL;
  c->have_ktei = prim_ktei != nullptr;
  return HFB200_OK;
}

int hfb200_sadatom_band_size(hfb200_ctx* c, int* nel, int* np_max) {
  if (!c) { set_error("null context"); return HFB200_EINVAL; }
  if (nel) *nel = c->nel;
  if (np_max) *np_max = c->np_max;
  return HFB200_OK;
}

int hfb200_sadatom_pack_band(hfb200_ctx* c, const double* M, int ld, double* band) {
  if (!c || !M || !band || ld < c->nrad) { set_error("pack_band: invalid argument"); return HFB200_EINVAL; }
  const int np = c->np_max, np2 = np * np;
  for (int e = 0; e < c->nel; e++) {
    const int f = c->first[e], npe = c->last[e] - f + 1;
    double* blk = band + (size_t)e * np2;
    if (npe < np) memset(blk, 0, sizeof(double) * np2);
    for (int j = 0; j < npe; j++)
      memcpy(blk + (size_t)j * np, M + (size_t)f + (size_t)(f + j) * ld, sizeof(double) * npe);
  }
  return HFB200_OK;
}

int hfb200_sadatom_unpack_band_add(hfb200_ctx* c, const double* band, double* M, int ld) {
  if (!c || !M || !band || ld < c->nrad) { set_error("unpack_band: invalid argument"); return HFB200_EINVAL; }
  const int np = c->np_max, np2 = np * np;
  for (int e = 0; e < c->nel; e++) {
    const int f = c->first[e], npe = c->last[e] - f + 1;
    for (int j = 0; j < npe; j++)
      for (int i = 0; i < npe; i++) M[(size_t)(f + i) + (size_t)(f + j) * ld] += band[(size_t)e * np2 + i + (size_t)j * np];
  }
  return HFB200_OK;
}

}  // extern "C"

// Kernels of `nbatch` builds whose per-build scratch (folded densities, traces, partial scalars) starts at build
// index `sb` of the context's scratch buffers, which the caller has reserved for `sb + nbatch` builds at least:
// chunks of one host batch run concurrently on different streams with disjoint scratch.
static int fock_batch_at(hfb200_ctx* c, size_t sb, int nbatch, int x_func, int c_func, int nspin, int nslices, double dens_thr,
                         const double* dP_band, double* dJ_band, double* dH_band, double* dScalars, cudaStream_t st);

static int fock_reserve(hfb200_ctx* c, size_t nbuilds, bool withJ, bool withH) {
  int rc;
  if (withJ) {
    if (!c->dTeiF) { set_error("Primitive teis have not been computed!"); return HFB200_ENOTEI; }
    if ((rc = c->dPtot.reserve(nbuilds * c->nel * 4 * c->jKS))) return rc;
    if ((rc = c->dTraces.reserve(nbuilds * c->nel * 2))) return rc;
  }
  if (withH && (rc = c->dPartial.reserve(nbuilds * c->nel * 4))) return rc;
  return 0;
}

extern "C" {

int hfb200_sadatom_fock_batch(hfb200_ctx* c, int nbatch, int x_func, int c_func, int nspin, int nslices,
                              double dens_thr, const double* dP_band, double* dJ_band, double* dH_band,
                              double* dScalars, void* stream) {
  if (!c || c->kind != CTX_SADATOM || nbatch < 1 || (nspin != 1 && nspin != 2) || nslices < 1 || !dP_band) {
    set_error("hfb200_sadatom_fock_batch: invalid argument");
    return HFB200_EINVAL;
  }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  int rc;
  if ((rc = fock_reserve(c, (size_t)nbatch, dJ_band != nullptr, dH_band != nullptr))) return rc;
  return fock_batch_at(c, 0, nbatch, x_func, c_func, nspin, nslices, dens_thr, dP_band, dJ_band, dH_band, dScalars, (cudaStream_t)stream);
}

}  // extern "C"

static int fock_batch_at(hfb200_ctx* c, size_t sb, int nbatch, int x_func, int c_func, int nspin, int nslices, double dens_thr,
                         const double* dP_band, double* dJ_band, double* dH_band, double* dScalars, cudaStream_t st) {
  int rc;
  double* sPtot = dJ_band ? c->dPtot.p + sb * c->nel * 4 * c->jKS : nullptr;
  double* sTraces = dJ_band ? c->dTraces.p + sb * c->nel * 2 : nullptr;
  double* sPartial = dH_band ? c->dPartial.p + sb * c->nel * 4 : nullptr;
  // J and Vxc in one build: the grid kernel emits the folded density + traces of the Coulomb kernel, P is read once
  const bool fused_pack = dJ_band && dH_band;
  if (dH_band) {
    int fam = 0;
    if (!valid_pair(x_func, c_func, &fam)) {
      set_error("functional pair (%d,%d) has no device implementation (no CPU fallback)", x_func, c_func);
      return HFB200_EFUNC;
    }
    SadGridParams g;
    g.Bt = c->dBt; g.dBt = c->ddBt; g.r = c->dR; g.wrad = c->dW;
    g.nel = c->nel; g.nq = c->nq; g.nq8 = c->nq8; g.ldq = c->ldq; g.np_max = c->np_max; g.np_pad = c->np_pad;
    g.nslices = nslices; g.nb = nbatch;
    g.x_func = x_func; g.c_func = c_func; g.dens_thr = dens_thr;
    g.P = dP_band; g.H = dH_band; g.partial = sPartial;
    g.pc = fused_pack ? sPtot : nullptr;
    g.traces = sTraces; g.dLc = c->dLc; g.dm1Lc = c->dm1Lc; g.tri = c->dTri;
    g.ntri = c->ntri; g.jKP = fused_pack ? 4 * c->jKS : 0; g.jMP = 8 * c->jMT;
    if ((rc = launch_grid(fam, nspin, g, st))) return rc;
    if (dScalars) {
      sad_reduce_scalars_kernel<<<nbatch, 32, 0, st>>>(sPartial, c->nel, dScalars);
      HFB_CUDA_OK(cudaGetLastError());
    }
  }
  if (dJ_band) {
    SadJParams q;
    q.teiF = c->dTeiF; q.dLc = c->dLc; q.dm1Lc = c->dm1Lc; q.tri = c->dTri;
    q.nel = c->nel; q.np_max = c->np_max; q.ntri = c->ntri; q.MT = c->jMT; q.KS = c->jKS;
    q.nslices = nslices; q.nspin = nspin; q.nb = nbatch;
    q.P = dP_band; q.pc = sPtot; q.traces = sTraces; q.J = dJ_band;
    if ((rc = launch_coulomb(q, !fused_pack, st))) return rc;
  }
  return HFB200_OK;
}

extern "C" {

int hfb200_sadatom_coulomb(hfb200_ctx* c, const double* P, int ldp, double* J, int ldj) {
  if (!c || c->kind != CTX_SADATOM || !P || !J || ldp < c->nrad || ldj < c->nrad) {
    set_error("hfb200_sadatom_coulomb: density matrix does not match basis set");
    return HFB200_EINVAL;
  }
  if (!c->dTeiF) { set_error("Primitive teis have not been computed!"); return HFB200_ENOTEI; }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  const size_t nband = (size_t)c->nel * c->np_max * c->np_max;
  int rc;
  if ((rc = c->hIn.reserve(nband)) || (rc = c->hOut.reserve(nband)) || (rc = c->dP.reserve(nband)) || (rc = c->dJ.reserve(nband))) return rc;
  hfb200_sadatom_pack_band(c, P, ldp, c->hIn.p);
  HFB_CUDA_OK(cudaMemcpyAsync(c->dP.p, c->hIn.p, nband * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  if ((rc = hfb200_sadatom_fock_batch(c, 1, 0, 0, 1, 1, 0.0, c->dP.p, c->dJ.p, nullptr, nullptr, c->stream))) return rc;
  HFB_CUDA_OK(cudaMemcpyAsync(c->hOut.p, c->dJ.p, nband * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  for (int j = 0; j < c->nrad; j++) memset(J + (size_t)j * ldj, 0, sizeof(double) * c->nrad);
  HFB_CUDA_OK(cudaStreamSynchronize(c->stream));
  return hfb200_sadatom_unpack_band_add(c, c->hOut.p, J, ldj);
}

int hfb200_sadatom_eval_fxc(hfb200_ctx* c, int x_func, int c_func, int nspin, const double* Pa,
                            const double* Pb, int ldp, int nslices, double dens_thr, double* Ha,
                            double* Hb, int ldh, int beta, double* Exc, double* Nel) {
  if (!c || c->kind != CTX_SADATOM || !Pa || !Ha || (nspin != 1 && nspin != 2) || ldp < c->nrad || ldh < c->nrad ||
      nslices < 1 || (nspin == 2 && (!Pb || (beta && !Hb)))) {
    set_error("hfb200_sadatom_eval_fxc: invalid argument / empty density matrix");
    return HFB200_EINVAL;
  }
  int fam = 0;
  if (!valid_pair(x_func, c_func, &fam)) {
    set_error("functional pair (%d,%d) has no device implementation (no CPU fallback)", x_func, c_func);
    return HFB200_EFUNC;
  }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  const size_t nblk = (size_t)c->nel * c->np_max * c->np_max;
  const size_t nband = nblk * nslices * nspin;
  int rc;
  if ((rc = c->hIn.reserve(nband)) || (rc = c->hOut.reserve(nband + 4)) || (rc = c->dP.reserve(nband)) ||
      (rc = c->dH.reserve(nband)) || (rc = c->dScalars.reserve(4)))
    return rc;
  for (int s = 0; s < nspin; s++)
    for (int l = 0; l < nslices; l++)
      hfb200_sadatom_pack_band(c, (s ? Pb : Pa) + (size_t)l * ldp * c->nrad, ldp, c->hIn.p + ((size_t)s * nslices + l) * nblk);
  HFB_CUDA_OK(cudaMemcpyAsync(c->dP.p, c->hIn.p, nband * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  if ((rc = hfb200_sadatom_fock_batch(c, 1, x_func, c_func, nspin, nslices, dens_thr, c->dP.p, nullptr, c->dH.p, c->dScalars.p, c->stream)))
    return rc;
  HFB_CUDA_OK(cudaMemcpyAsync(c->hOut.p, c->dH.p, nband * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  HFB_CUDA_OK(cudaMemcpyAsync(c->hOut.p + nband, c->dScalars.p, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  // zero the dense outputs while the device works (H.zeros() in the reference)
  for (int s = 0; s < nspin; s++) {
    double* H = s ? Hb : Ha;
    if (!H) continue;
    for (int l = 0; l < nslices; l++)
      for (int j = 0; j < c->nrad; j++) memset(H + (size_t)l * ldh * c->nrad + (size_t)j * ldh, 0, sizeof(double) * c->nrad);
  }
  HFB_CUDA_OK(cudaStreamSynchronize(c->stream));
  for (int s = 0; s < nspin; s++) {
    double* H = s ? Hb : Ha;
    if (s == 1 && !beta) continue;
    for (int l = 0; l < nslices; l++)
      hfb200_sadatom_unpack_band_add(c, c->hOut.p + ((size_t)s * nslices + l) * nblk, H + (size_t)l * ldh * c->nrad, ldh);
  }
  if (Exc) *Exc = c->hOut.p[nband + 0];
  if (Nel) *Nel = c->hOut.p[nband + 1];
  return HFB200_OK;
}

/* Batched Fock builds with HOST element-band buffers (pinned or pageable): densities down, J / H bands and
 * scalars back, pipelined in chunks over three streams (H2D, kernels and D2H of different chunks overlap);
 * returns after the results are on the host. */
int hfb200_sadatom_fock_host(hfb200_ctx* c, int nbatch, int x_func, int c_func, int nspin, int nslices,
                             double dens_thr, const double* hP_band, double* hJ_band, double* hH_band,
                             double* hScalars) {
  if (!c || c->kind != CTX_SADATOM || nbatch < 1 || !hP_band || (nspin != 1 && nspin != 2) || nslices < 1) {
    set_error("hfb200_sadatom_fock_host: invalid argument");
    return HFB200_EINVAL;
  }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  const size_t nblk = (size_t)c->nel * c->np_max * c->np_max;
  const size_t nP = nblk * nslices * nspin * nbatch, nJ = nblk * nbatch;
  int rc;
  if ((rc = c->dP.reserve(nP)) || (hH_band && (rc = c->dH.reserve(nP))) || (hJ_band && (rc = c->dJ.reserve(nJ))) ||
      (rc = c->dScalars.reserve((size_t)4 * nbatch)))
    return rc;
  if ((rc = fock_reserve(c, (size_t)nbatch, hJ_band != nullptr, hH_band != nullptr))) return rc;
  if (hH_band) {
    int fam = 0;
    if (!valid_pair(x_func, c_func, &fam)) {
      set_error("functional pair (%d,%d) has no device implementation (no CPU fallback)", x_func, c_func);
      return HFB200_EFUNC;
    }
  }
  // Chunks of builds round-robin over three streams: the H2D of one chunk, the kernels of another and the D2H of a
  // third overlap, so the batch costs max(H2D, D2H) instead of their sum (PCIe is full duplex).
  for (int q = 0; q < 3; q++)
    if (!c->xs[q]) HFB_CUDA_OK(cudaStreamCreateWithFlags(&c->xs[q], cudaStreamNonBlocking));
  // (chunks stay above 16 builds: the Coulomb kernel picks a split-K variant below that, and results must not
  // depend on how a batch is chunked -- the host path reproduces the device-resident path bit for bit)
  const int chunk = std::max(32, (nbatch + 11) / 12);
  const size_t perP = nblk * nslices * nspin;
  int ci = 0;
  for (int b0 = 0, nb = 0; b0 < nbatch; b0 += nb, ci++) {
    nb = std::min(chunk, nbatch - b0);
    if (nbatch - b0 - nb > 0 && nbatch - b0 - nb <= 16) nb = nbatch - b0;
    cudaStream_t st = c->xs[ci % 3];
    HFB_CUDA_OK(cudaMemcpyAsync(c->dP.p + b0 * perP, hP_band + b0 * perP, nb * perP * sizeof(double), cudaMemcpyHostToDevice, st));
    if ((rc = fock_batch_at(c, (size_t)b0, nb, x_func, c_func, nspin, nslices, dens_thr, c->dP.p + b0 * perP,
                            hJ_band ? c->dJ.p + b0 * nblk : nullptr, hH_band ? c->dH.p + b0 * perP : nullptr,
                            (hH_band && hScalars) ? c->dScalars.p + (size_t)4 * b0 : nullptr, st)))
      return rc;
    if (hJ_band) HFB_CUDA_OK(cudaMemcpyAsync(hJ_band + b0 * nblk, c->dJ.p + b0 * nblk, nb * nblk * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (hH_band) HFB_CUDA_OK(cudaMemcpyAsync(hH_band + b0 * perP, c->dH.p + b0 * perP, nb * perP * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (hH_band && hScalars)
      HFB_CUDA_OK(cudaMemcpyAsync(hScalars + (size_t)4 * b0, c->dScalars.p + (size_t)4 * b0, (size_t)4 * nb * sizeof(double), cudaMemcpyDeviceToHost, st));
  }
  for (int q = 0; q < 3; q++) HFB_CUDA_OK(cudaStreamSynchronize(c->xs[q]));
  return HFB200_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------------
// Allocation-free variant of hfb200_sadatom_fock_batch (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that
// forbid it).  The per-build scratch of the batch kernels (triangle-folded densities `pc`, disjoint traces, partial
// XC scalars) lives in ONE file-scope device array of HFB200_STATIC_SCRATCH_BYTES instead of the context's growable
// cudaMalloc'ed buffers.  Every GPU that loads the library holds the array until the process exits.  All calls (of all
// contexts on one device) share it and it keeps the last call's contents: the range a call uses is cleared on the
// call's stream first, and calls that overlap on the GPU (different streams, or different host threads) race.
// ------------------------------------------------------------------------------------------------------------------
#include <mutex>

static_assert(HFB200_STATIC_SCRATCH_BYTES % 256 == 0, "scratch size must keep the 256-byte alignment of its parts");
static __device__ __align__(256) double g_static_scratch[HFB200_STATIC_SCRATCH_BYTES / sizeof(double)];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
double* g_static_ptr[kMaxDevices] = {};  // DEVICE address of g_static_scratch on each device, resolved once per device

// In host code `g_static_scratch` is only the host-side stand-in of the symbol, not a device address: it must never be
// handed to a kernel.  Resolve the device address once per device, with that device current.
int static_scratch_address(int device, double** out) {
  if (device < 0 || device >= kMaxDevices) { set_error("device %d out of range", device); return HFB200_EINVAL; }
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[device]) {
    int cur = -1;
    HFB_CUDA_OK(cudaGetDevice(&cur));
    if (cur != device) HFB_CUDA_OK(cudaSetDevice(device));
    double* p = nullptr;
    HFB_CUDA_OK(cudaGetSymbolAddress((void**)&p, g_static_scratch));
    if (!p) { set_error("cudaGetSymbolAddress returned a null address for the static scratch"); return HFB200_ECUDA; }
    g_static_ptr[device] = p;
  }
  *out = g_static_ptr[device];
  return 0;
}

size_t round32(size_t n) { return (n + 31) & ~(size_t)31; }  // 32 doubles = 256 bytes

}  // namespace

extern "C" int hfb200_sadatom_fock_batch_static(hfb200_ctx* c, int nbatch, int x_func, int c_func, int nspin, int nslices,
                                                double dens_thr, const double* dP_band, double* dJ_band, double* dH_band,
                                                double* dScalars, void* stream) {
  if (!c || c->kind != CTX_SADATOM || nbatch < 1 || (nspin != 1 && nspin != 2) || nslices < 1 || !dP_band) {
    set_error("hfb200_sadatom_fock_batch_static: invalid argument");
    return HFB200_EINVAL;
  }
  if (dJ_band && !c->dTeiF) { set_error("Primitive teis have not been computed!"); return HFB200_ENOTEI; }
  int fam = 0;
  if (dH_band && !valid_pair(x_func, c_func, &fam)) {
    set_error("functional pair (%d,%d) has no device implementation (no CPU fallback)", x_func, c_func);
    return HFB200_EFUNC;
  }
  // layout inside the static array: [pc | traces | partial], each part starting on a 256-byte boundary
  const size_t nPc = dJ_band ? round32((size_t)nbatch * c->nel * 4 * c->jKS) : 0;
  const size_t nTr = dJ_band ? round32((size_t)nbatch * c->nel * 2) : 0;
  const size_t nPa = dH_band ? round32((size_t)nbatch * c->nel * 4) : 0;
  const size_t need = (nPc + nTr + nPa) * sizeof(double);
  if (need > (size_t)HFB200_STATIC_SCRATCH_BYTES) {
    set_error("hfb200_sadatom_fock_batch_static: %d builds need %zu bytes of scratch, the static array holds %zu "
              "(HFB200_STATIC_SCRATCH_BYTES); split the batch or use hfb200_sadatom_fock_batch",
              nbatch, need, (size_t)HFB200_STATIC_SCRATCH_BYTES);
    return HFB200_EINVAL;
  }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  double* scratch = nullptr;
  int rc;
  if ((rc = static_scratch_address(c->device, &scratch))) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // the array keeps the previous call's contents; every part is rewritten in full before it is read, the clear makes
  // that independent of the kernels' write patterns
  if (need) HFB_CUDA_OK(cudaMemsetAsync(scratch, 0, need, st));
  double* sPtot = dJ_band ? scratch : nullptr;
  double* sTraces = dJ_band ? scratch + nPc : nullptr;
  double* sPartial = dH_band ? scratch + nPc + nTr : nullptr;
  const bool fused_pack = dJ_band && dH_band;
  if (dH_band) {
    SadGridParams g;
    g.Bt = c->dBt; g.dBt = c->ddBt; g.r = c->dR; g.wrad = c->dW;
    g.nel = c->nel; g.nq = c->nq; g.nq8 = c->nq8; g.ldq = c->ldq; g.np_max = c->np_max; g.np_pad = c->np_pad;
    g.nslices = nslices; g.nb = nbatch;
    g.x_func = x_func; g.c_func = c_func; g.dens_thr = dens_thr;
    g.P = dP_band; g.H = dH_band; g.partial = sPartial;
    g.pc = fused_pack ? sPtot : nullptr;
    g.traces = sTraces; g.dLc = c->dLc; g.dm1Lc = c->dm1Lc; g.tri = c->dTri;
    g.ntri = c->ntri; g.jKP = fused_pack ? 4 * c->jKS : 0; g.jMP = 8 * c->jMT;
    if ((rc = launch_grid(fam, nspin, g, st))) return rc;
    if (dScalars) {
      sad_reduce_scalars_kernel<<<nbatch, 32, 0, st>>>(sPartial, c->nel, dScalars);
      HFB_CUDA_OK(cudaGetLastError());
    }
  }
  if (dJ_band) {
    SadJParams q;
    q.teiF = c->dTeiF; q.dLc = c->dLc; q.dm1Lc = c->dm1Lc; q.tri = c->dTri;
    q.nel = c->nel; q.np_max = c->np_max; q.ntri = c->ntri; q.MT = c->jMT; q.KS = c->jKS;
    q.nslices = nslices; q.nspin = nspin; q.nb = nbatch;
    q.P = dP_band; q.pc = sPtot; q.traces = sTraces; q.J = dJ_band;
    if ((rc = launch_coulomb(q, !fused_pack, st))) return rc;
  }
  return HFB200_OK;
}
