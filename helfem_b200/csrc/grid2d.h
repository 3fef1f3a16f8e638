// Product-grid XC engine of the atomic / diatomic programs (grid2d.cu).
#pragma once
#include <cuda_runtime.h>
#include <vector>
#include "context.h"

namespace hfb {

struct Grid2D {
  bool ready = false;
  int kind = 0;  // 0 atomic (spherical polar metric), 1 diatomic (prolate spheroidal metric)
  double Rhalf = 0.0;
  int nang = 0, npair = 0, nel = 0, nq = 0, np_max = 0, nrad = 0, npts = 0;
  int launches = 0;  // kernels launched by the last eval_fxc
  std::vector<int> first, npe;
  double *dChi = nullptr, *dDChi = nullptr;  // [e][np_max][nq], q fastest
  double *dR = nullptr, *dWr = nullptr;      // [e][nq]
  double *dCth = nullptr, *dWang = nullptr;  // [npts]
  double* dWp = nullptr;                     // [point][5][pair]   density stage
  double* dWt = nullptr;                     // [5][pair][point]   Fock stage (tables 1, 2 symmetrised)
  int *dFirst = nullptr, *dNpe = nullptr;
  DevBuf P, H, Rq, F, VGT, partial, scal;

  int setup(int kind, double Rhalf, const std::vector<int>& lval, const std::vector<int>& mval, int nel,
            const std::vector<int>& first, const std::vector<int>& last, int nq, const double* r, const double* wrad,
            const double* bf, const double* df, int npts, const double* cth, const double* wang, const double* Y,
            const double* dY);
  int eval_fxc(int fam, int x_func, int c_func, int nspin, const double* Pa, const double* Pb, int ldp, double dens_thr,
               double* Ha, double* Hb, int ldh, int beta, double* scalars3, cudaStream_t st);
  void release();
};

}  // namespace hfb
