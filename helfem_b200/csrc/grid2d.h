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
  std::vector<int> first, npe, mval;
  double *dChi = nullptr, *dDChi = nullptr;  // [e][np_max][nq], q fastest
  double *dR = nullptr, *dWr = nullptr;      // [e][nq]
  double *dCth = nullptr, *dWang = nullptr;  // [npts]
  int *dFirst = nullptr, *dNpe = nullptr;
  // ---- general angular rule: real pair tables over all points ----
  double* dWp = nullptr;                     // [point][5][pair]   density stage
  double* dWt = nullptr;                     // [5][pair][point]   Fock stage (tables 1, 2 symmetrised)
  // ---- product rule theta x uniform phi (what angular::angular_chebyshev builds): Delta-m factorisation ----
  bool product = false;
  int nth = 0, nph = 0, nD = 0, dmax = 0;    // nD = 2 dmax + 1 classes Delta = m_a - m_b
  double *dTh = nullptr, *dDTh = nullptr;    // [a][t]  Theta_a = Y_a(theta_t, 0), d/dtheta
  double *dCosT = nullptr, *dSinT = nullptr; // [class][j]  cos / sin (Delta phi_j)
  int* dM = nullptr;                         // [a]
  int* dClsPtr = nullptr;                    // [nD+1]
  int2* dClsPair = nullptr;                  // pairs (a, b) sorted by class
  std::vector<int> cls_ptr;
  std::vector<int2> cls_pair;
  // ---- per call ----
  DevBuf Pp, Hp, Rq, F, VGT, partial, scal, Aq, Fh;
  PinBuf hX;
  int *dPidx = nullptr, *dOidx = nullptr;    // [npair] angular block -> packed slot of P / H (-1: not held)
  int2* dOutBlk = nullptr;                   // [npair] packed output slot -> (a, b)
  int4* dChunks = nullptr;                   // [npair] stage C2 work: (class, first slot, count, -)
  unsigned char* dNzR = nullptr;             // [nspin][nel][npair] the radial density of the pair is non-zero
  size_t nzr_cap = 0;
  int* dClsNz = nullptr;                     // [nspin*nel][nD] class has a non-zero pair (product rule)
  size_t clsnz_cap = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;  // around the kernels of the last eval_fxc
  float device_ms = 0.f;
  double h2d_bytes = 0.0, d2h_bytes = 0.0;    // packed payload of the last eval_fxc

  int setup(int kind, double Rhalf, const std::vector<int>& lval, const std::vector<int>& mval, int nel,
            const std::vector<int>& first, const std::vector<int>& last, int nq, const double* r, const double* wrad,
            const double* bf, const double* df, int npts, const double* cth, const double* wang, const double* Y,
            const double* dY);
  // Pa/Pb, Ha/Hb: HOST matrices of order nout with leading dimensions ldp / ldh.  skip (optional, [nang]): number of leading
  // radial functions of each angular function that are absent from the caller's matrices (the diatomic boundary functions).
  int eval_fxc(int fam, int x_func, int c_func, int nspin, const double* Pa, const double* Pb, int ldp, double dens_thr,
               double* Ha, double* Hb, int ldh, int beta, double* scalars3, cudaStream_t st, const int* skip);
  void release();
};

}  // namespace hfb
