// Launch dispatch of sad_grid_mma_kernel for ONE functional family; instantiated by sadatom_grid_{lda,gga,mgga}.cu
// so that the three families compile in parallel.
#pragma once
#include "context.h"
#include "sadatom_grid_mma.cuh"

namespace hfb {

#ifndef HFB_GRID_GPC
#define HFB_GRID_GPC 1
#endif

template <int FAM, int NSPIN, int NT, int PAIR>
int launch_grid_inst(const SadGridParams& prm_in, int nwarps, size_t smem, cudaStream_t st) {
  SadGridParams prm = prm_in;
  // a CTA can keep its element's tables for up to HFB_GRID_GPC groups of builds, as long as the grid still has
  // several waves of CTAs for the hardware scheduler to balance (elements differ in screened points).  Measured on
  // B200 (cfg2, 1024 builds): 1 group per CTA 0.176 ms, 2: 0.181, 3: 0.185 -- dynamic balance wins, default 1.
  const int ngroups = (prm.nb + nwarps - 1) / nwarps;
  int gpc = HFB_GRID_GPC;
  while (gpc > 1 && (long)prm.nel * ((ngroups + gpc - 1) / gpc) < 4L * 148) gpc--;
  prm.gpc = gpc;
  dim3 grid(prm.nel, (ngroups + gpc - 1) / gpc);
  HFB_CUDA_OK(cudaFuncSetAttribute(sad_grid_mma_kernel<FAM, NSPIN, NT, PAIR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  sad_grid_mma_kernel<FAM, NSPIN, NT, PAIR><<<grid, 32 * nwarps, smem, st>>>(prm);
  HFB_CUDA_OK(cudaGetLastError());
  return 0;
}

template <int FAM, int NSPIN, int PAIR>
int launch_grid_nt(const SadGridParams& prm, int nwarps, size_t smem, cudaStream_t st) {
  switch (prm.np_pad / 8) {
    case 1: return launch_grid_inst<FAM, NSPIN, 1, PAIR>(prm, nwarps, smem, st);
    case 2: return launch_grid_inst<FAM, NSPIN, 2, PAIR>(prm, nwarps, smem, st);
    case 3: return launch_grid_inst<FAM, NSPIN, 3, PAIR>(prm, nwarps, smem, st);
    case 4: return launch_grid_inst<FAM, NSPIN, 4, PAIR>(prm, nwarps, smem, st);
    default:
      set_error("more than 32 functions per element are not supported by the grid kernel");
      return HFB200_EINVAL;
  }
}

template <int FAM>
int launch_grid_family(int nspin, const SadGridParams& prm, cudaStream_t st) {
  // one warp per build; up to HFB_GRID_WARPS builds share a CTA's staged tables
  int nwarps = prm.nb < HFB_GRID_WARPS ? prm.nb : HFB_GRID_WARPS;
  size_t smem = sad_grid_mma_smem_bytes(FAM, nspin, prm.np_pad, prm.nq8, prm.ldq, nwarps, prm.jKP);
  while (smem > 227 * 1024 && nwarps > 1) {
    nwarps >>= 1;
    smem = sad_grid_mma_smem_bytes(FAM, nspin, prm.np_pad, prm.nq8, prm.ldq, nwarps, prm.jKP);
  }
  if (smem > 227 * 1024) { set_error("element tables need %zu bytes of shared memory (>227 KB)", smem); return HFB200_EINVAL; }
  const bool canonical = prm.x_func == CanonicalPair<FAM>::x && prm.c_func == CanonicalPair<FAM>::c;
  if (nspin == 1) return canonical ? launch_grid_nt<FAM, 1, 1>(prm, nwarps, smem, st) : launch_grid_nt<FAM, 1, 0>(prm, nwarps, smem, st);
  return canonical ? launch_grid_nt<FAM, 2, 1>(prm, nwarps, smem, st) : launch_grid_nt<FAM, 2, 0>(prm, nwarps, smem, st);
}

}  // namespace hfb
