// C ABI: exchange() of the spherically averaged atom (include/helfem_b200.h).
#include <cmath>
#include "../../include/helfem_b200.h"
#include "../host/angular.h"
#include "context.h"
#include "exchange_kernels.cuh"

using namespace hfb;

namespace {

// Constant tables of the exchange path, built on first use.
int exchange_setup(hfb200_ctx* c) {
  if (c->dCpl) return 0;
  const int nel = c->nel, nrad = c->nrad, nsl = c->lmax + 1, nL = 2 * c->lmax + 1;
  std::vector<int> npe(nel), elem_of(2 * (size_t)nrad, -1);
  for (int e = 0; e < nel; e++) {
    npe[e] = c->last[e] - c->first[e] + 1;
    for (int a = c->first[e]; a <= c->last[e]; a++) {
      if (elem_of[2 * a] < 0) elem_of[2 * a] = e;
      else if (elem_of[2 * a + 1] < 0) elem_of[2 * a + 1] = e;
      else { set_error("radial function %d belongs to more than two elements", a); return HFB200_EINVAL; }
    }
  }
  // sadatom/basis.cpp:428-456: totcoup(L) = sum_{mout,min} G(lout mout; L M; lin min)^2 / (2 lout + 1),
  // density prefactor Lfac*totcoup with Lfac = 4 pi/(2L+1)
  std::vector<double> cpl((size_t)nsl * nsl * nL, 0.0);
  for (int lout = 0; lout < nsl; lout++)
    for (int lin = 0; lin < nsl; lin++)
      for (int L = std::abs(lin - lout); L <= lin + lout; L++) {
        double tot = 0.0;
        for (int mout = -lout; mout <= lout; mout++)
          for (int min = -lin; min <= lin; min++) {
            const double g = angular::gaunt_coefficient(lout, mout, L, mout - min, lin, min);
            tot += g * g;
          }
        tot /= 2 * lout + 1;
        if (tot == 0.0) continue;
        cpl[((size_t)lout * nsl + lin) * nL + L] = (4.0 * M_PI / (2 * L + 1)) * tot;
      }
  HFB_CUDA_OK(cudaMalloc(&c->dFirst, nel * sizeof(int)));
  HFB_CUDA_OK(cudaMalloc(&c->dNpe, nel * sizeof(int)));
  HFB_CUDA_OK(cudaMalloc(&c->dElemOf, elem_of.size() * sizeof(int)));
  HFB_CUDA_OK(cudaMemcpy(c->dFirst, c->first.data(), nel * sizeof(int), cudaMemcpyHostToDevice));
  HFB_CUDA_OK(cudaMemcpy(c->dNpe, npe.data(), nel * sizeof(int), cudaMemcpyHostToDevice));
  HFB_CUDA_OK(cudaMemcpy(c->dElemOf, elem_of.data(), elem_of.size() * sizeof(int), cudaMemcpyHostToDevice));
  HFB_CUDA_OK(cudaMalloc(&c->dCpl, cpl.size() * sizeof(double)));
  HFB_CUDA_OK(cudaMemcpy(c->dCpl, cpl.data(), cpl.size() * sizeof(double), cudaMemcpyHostToDevice));
  return 0;
}

}  // namespace

extern "C" int hfb200_sadatom_exchange(hfb200_ctx* c, const double* P, int ldp, int nslices, double* K, int ldk) {
  if (!c || c->kind != CTX_SADATOM || !P || !K) { set_error("hfb200_sadatom_exchange: invalid argument"); return HFB200_EINVAL; }
  if (nslices != c->lmax + 1) { set_error("Density matrix am does not match basis set!"); return HFB200_EINVAL; }
  if (ldp < c->nrad || ldk < c->nrad) { set_error("Density matrix does not match basis set!"); return HFB200_EINVAL; }
  if (!c->dKtei || c->nL < 2 * c->lmax + 1) { set_error("Primitive teis have not been computed!"); return HFB200_ENOTEI; }
  HFB_CUDA_OK(cudaSetDevice(c->device));
  int rc;
  if ((rc = exchange_setup(c))) return rc;
  const int nrad = c->nrad, np2 = c->np_max * c->np_max;
  const size_t ncube = (size_t)nrad * nrad * nslices;
  if ((rc = c->dP.reserve(ncube)) || (rc = c->dK.reserve(ncube)) || (rc = c->dH.reserve((size_t)nslices * c->nel * c->nel * np2)))
    return rc;
  // dense cube down (compact ld = nrad on the device)
  for (int l = 0; l < nslices; l++)
    HFB_CUDA_OK(cudaMemcpy2DAsync(c->dP.p + (size_t)l * nrad * nrad, nrad * sizeof(double), P + (size_t)l * ldp * nrad,
                                  ldp * sizeof(double), nrad * sizeof(double), nrad, cudaMemcpyHostToDevice, c->stream));
  SadExchParams q;
  q.P = c->dP.p; q.ldp = nrad; q.nrad = nrad; q.nslices = nslices; q.nL = 2 * c->lmax + 1; q.nel = c->nel; q.np_max = c->np_max;
  q.first = c->dFirst; q.npe = c->dNpe; q.elem_of = c->dElemOf; q.cpl = c->dCpl;
  q.dL = c->dDisL; q.dm1L = c->dDisM1L; q.ktei = c->dKtei; q.blocks = c->dH.p; q.K = c->dK.p; q.ldk = nrad;
  dim3 grid(c->nel * c->nel, nslices);
  sad_exchange_block_kernel<<<grid, 256, 3 * np2 * sizeof(double), c->stream>>>(q);
  HFB_CUDA_OK(cudaGetLastError());
  sad_exchange_assemble_kernel<<<148 * 4, 256, 0, c->stream>>>(q);
  HFB_CUDA_OK(cudaGetLastError());
  for (int l = 0; l < nslices; l++)
    HFB_CUDA_OK(cudaMemcpy2DAsync(K + (size_t)l * ldk * nrad, ldk * sizeof(double), c->dK.p + (size_t)l * nrad * nrad,
                                  nrad * sizeof(double), nrad * sizeof(double), nrad, cudaMemcpyDeviceToHost, c->stream));
  HFB_CUDA_OK(cudaStreamSynchronize(c->stream));
  return HFB200_OK;
}
