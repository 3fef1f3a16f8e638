// Angular tables of the Fock-build path: Gaunt coefficients, spherical harmonics,
// angular quadrature.  All of these are constants of the hot path (computed once).
//
// Behavioural specification (reference files, read-only):
//   Gaunt coefficient           /root/reference/src/general/gaunt.cpp:35-53  (GSL 3j symbols there)
//   modified (cos^2) coefficient /root/reference/src/general/gaunt.cpp:55-71,167-180
//   Y_lm                        /root/reference/src/general/spherical_harmonics.cpp:25-41 (GSL sphPlm there)
//   theta x phi compound rule   /root/reference/src/general/angular.cpp:22-71
//   angular basis ordering      /root/reference/src/atomic/basis.cpp:174-198
// GSL is not part of this image; the same quantities are obtained from the stable recursion for
// the normalised associated Legendre functions and an exact Gauss-Legendre quadrature of the
// triple product (the integrand is a polynomial), which agrees with the 3j formula to rounding.
#pragma once
#include <complex>
#include "mat.h"

namespace hfb {
namespace angular {

/// sqrt((2l+1)/(4 pi) (l-m)!/(l+m)!) P_l^m(x) with Condon-Shortley phase, m>=0  (= gsl_sf_legendre_sphPlm)
double sph_plm(int l, int m, double x);
/// all l = m..lmax at once: out[l-m]
void sph_plm_array(int lmax, int m, double x, double* out);
std::complex<double> spherical_harmonics(int l, int m, double cth, double phi);

/// G = int conj(Y_L^M) Y_l^m Y_lp^mp dOmega
double gaunt_coefficient(int L, int M, int l, int m, int lp, int mp);
double modified_gaunt_coefficient(int lj, int mj, int L, int M, int li, int mi);

/// Gauss-Legendre rule on [-1,1]
void gauss_legendre(int n, Vec& x, Vec& w);
/// Chebyshev(theta) x uniform(phi) compound rule, point index i*nphi+j
void angular_chebyshev(int l, int m, Vec& cth, Vec& phi, Vec& w);
/// (l,m) list of the atomic programs: m=0, then +-1, ... with +m before -m per l
void angular_basis(int lmax, int mmax, std::vector<int>& lval, std::vector<int>& mval);

}  // namespace angular
}  // namespace hfb
