// Associated Legendre functions of the first and second kind for argument xi = cosh(mu) > 1,
// the radial factors of the Neumann expansion of 1/r12 in prolate spheroidal coordinates.
//
// Behavioural specification (reference, read-only): the Fortran library under
// /root/reference/src/legendre/ (calc_Plm_arr / calc_Qlm_arr, Legendre_Wrapper.f90:135-190) as
// consumed through /root/reference/src/general/legendretable.cpp:60-97.  No Fortran compiler exists
// in this image, so the functions are restated from their recurrences:
//   P_L^M(xi) = (xi^2-1)^{M/2} d^M P_L / d xi^M     (positive for xi > 1; upward recursion in L)
//   Q_L^M(xi) = (xi^2-1)^{M/2} d^M Q_L / d xi^M     (sign (-1)^M; Q_L^0 by Miller's backward
//                recursion in L normalised with Q_0 = 1/2 ln((xi+1)/(xi-1)), then upward in M)
// The conventions are pinned by the reference's own identity test (Neumann expansion = 1/r12,
// /root/reference/src/legendre/legendre_test.cpp:52-96) in tests/test_diatomic_setup.py.
#pragma once
#include "mat.h"

namespace hfb {
namespace legendre {

/// P(L, M) and Q(L, M) for 0 <= M <= Mmax, M <= L <= Lmax at xi = cosh(mu), mu > 0.
/// Layout out[M * (Lmax+1) + L]; entries with L < M are zero.
void plm_array(int Lmax, int Mmax, double mu, double* out);
void qlm_array(int Lmax, int Mmax, double mu, double* out);

}  // namespace legendre
}  // namespace hfb
