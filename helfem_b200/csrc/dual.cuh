// Forward-mode dual numbers for the device XC functionals.
//
// The reference obtains exc, vrho, vsigma, vtau from libxc
// (/root/reference/src/sadatom/dftgrid.cpp:349-398, xc_{lda,gga,mgga}_exc_vxc).  The device
// implementation writes every functional once as an energy expression templated on the
// scalar type and differentiates it by carrying N partial derivatives alongside the value.
// Everything is inlined; with N<=4 the partials live in registers.
#pragma once
#include <cmath>

#if defined(__CUDACC__)
#define HFB_HD __host__ __device__ __forceinline__
#else
#define HFB_HD inline
#endif

namespace hfb {
namespace xc {

template <int N>
struct Dual {
  double v;
  double d[N];
  HFB_HD Dual() {}
  HFB_HD Dual(double c) : v(c) {
#pragma unroll
    for (int i = 0; i < N; i++) d[i] = 0.0;
  }
  /// independent variable number k
  HFB_HD static Dual var(double x, int k) {
    Dual r(x);
    r.d[k] = 1.0;
    return r;
  }
};

template <int N> HFB_HD double val(const Dual<N>& a) { return a.v; }
HFB_HD double val(double a) { return a; }

// ---- arithmetic ----
template <int N> HFB_HD Dual<N> operator+(const Dual<N>& a, const Dual<N>& b) {
  Dual<N> r; r.v = a.v + b.v;
#pragma unroll
  for (int i = 0; i < N; i++) r.d[i] = a.d[i] + b.d[i];
  return r;
}
template <int N> HFB_HD Dual<N> operator+(const Dual<N>& a, double b) { Dual<N> r = a; r.v += b; return r; }
template <int N> HFB_HD Dual<N> operator+(double b, const Dual<N>& a) { Dual<N> r = a; r.v += b; return r; }
template <int N> HFB_HD Dual<N> operator-(const Dual<N>& a) {
  Dual<N> r; r.v = -a.v;
#pragma unroll
  for (int i = 0; i < N; i++) r.d[i] = -a.d[i];
  return r;
}
template <int N> HFB_HD Dual<N> operator-(const Dual<N>& a, const Dual<N>& b) {
  Dual<N> r; r.v = a.v - b.v;
#pragma unroll
  for (int i = 0; i < N; i++) r.d[i] = a.d[i] - b.d[i];
  return r;
}
template <int N> HFB_HD Dual<N> operator-(const Dual<N>& a, double b) { Dual<N> r = a; r.v -= b; return r; }
template <int N> HFB_HD Dual<N> operator-(double b, const Dual<N>& a) { Dual<N> r = -a; r.v += b; return r; }
template <int N> HFB_HD Dual<N> operator*(const Dual<N>& a, const Dual<N>& b) {
  Dual<N> r; r.v = a.v * b.v;
#pragma unroll
  for (int i = 0; i < N; i++) r.d[i] = a.d[i] * b.v + a.v * b.d[i];
  return r;
}
template <int N> HFB_HD Dual<N> operator*(const Dual<N>& a, double b) {
  Dual<N> r; r.v = a.v * b;
#pragma unroll
  for (int i = 0; i < N; i++) r.d[i] = a.d[i] * b;
  return r;
}
template <int N> HFB_HD Dual<N> operator*(double b, const Dual<N>& a) { return a * b; }
template <int N> HFB_HD Dual<N> operator/(const Dual<N>& a, const Dual<N>& b) {
  Dual<N> r; const double ib = 1.0 / b.v; r.v = a.v * ib;
#pragma unroll
  for (int i = 0; i < N; i++) r.d[i] = (a.d[i] - r.v * b.d[i]) * ib;
  return r;
}
template <int N> HFB_HD Dual<N> operator/(const Dual<N>& a, double b) { return a * (1.0 / b); }
template <int N> HFB_HD Dual<N> operator/(double a, const Dual<N>& b) {
  Dual<N> r; const double ib = 1.0 / b.v; r.v = a * ib;
#pragma unroll
  for (int i = 0; i < N; i++) r.d[i] = -r.v * b.d[i] * ib;
  return r;
}

// ---- elementary functions: f(a) with derivative fp ----
template <int N> HFB_HD Dual<N> chain(const Dual<N>& a, double f, double fp) {
  Dual<N> r; r.v = f;
#pragma unroll
  for (int i = 0; i < N; i++) r.d[i] = fp * a.d[i];
  return r;
}
template <int N> HFB_HD Dual<N> sqrt(const Dual<N>& a) { const double s = ::sqrt(a.v); return chain(a, s, 0.5 / s); }
template <int N> HFB_HD Dual<N> cbrt(const Dual<N>& a) { const double c = ::cbrt(a.v); return chain(a, c, c / (3.0 * a.v)); }
template <int N> HFB_HD Dual<N> exp(const Dual<N>& a) { const double e = ::exp(a.v); return chain(a, e, e); }
template <int N> HFB_HD Dual<N> log(const Dual<N>& a) { return chain(a, ::log(a.v), 1.0 / a.v); }
template <int N> HFB_HD Dual<N> log1p(const Dual<N>& a) { return chain(a, ::log1p(a.v), 1.0 / (1.0 + a.v)); }
template <int N> HFB_HD Dual<N> expm1(const Dual<N>& a) { const double e = ::expm1(a.v); return chain(a, e, e + 1.0); }
template <int N> HFB_HD Dual<N> pow(const Dual<N>& a, double p) {
  const double f = ::pow(a.v, p);
  return chain(a, f, p * f / a.v);
}
HFB_HD double sqrt(double a) { return ::sqrt(a); }
HFB_HD double cbrt(double a) { return ::cbrt(a); }
HFB_HD double exp(double a) { return ::exp(a); }
HFB_HD double log(double a) { return ::log(a); }
HFB_HD double log1p(double a) { return ::log1p(a); }
HFB_HD double expm1(double a) { return ::expm1(a); }
HFB_HD double pow(double a, double p) { return ::pow(a, p); }

/// Replace by a constant (derivatives vanish): what a clamped input does in libxc's
/// generated code, where d/dx max(thr, x) = 0 on the clamped branch.
template <int N> HFB_HD Dual<N> constant_like(const Dual<N>&, double c) { return Dual<N>(c); }
HFB_HD double constant_like(double, double c) { return c; }

}  // namespace xc
}  // namespace hfb
