// Minimal column-major FP64 containers used by the host-side mirror.
//
// The reference passes Armadillo objects (arma::mat / arma::cube, column-major,
// see /root/reference/src/sadatom/basis.h:91-93).  Armadillo is not part of this
// image, so the host mirror uses these PODs, which have the same memory layout:
// element (i,j) of a Mat sits at d[i + j*nr]; slice s of a Cube is a Mat at
// offset s*nr*nc.  arma_adaptor.h converts when Armadillo is available.
#pragma once
#include <cstddef>
#include <cstring>
#include <stdexcept>
#include <vector>

namespace hfb {

typedef std::vector<double> Vec;

struct Mat {
  size_t nr = 0, nc = 0;
  std::vector<double> d;
  Mat() {}
  Mat(size_t r, size_t c) : nr(r), nc(c), d(r * c, 0.0) {}
  void zeros(size_t r, size_t c) { nr = r; nc = c; d.assign(r * c, 0.0); }
  double& operator()(size_t i, size_t j) { return d[i + j * nr]; }
  double operator()(size_t i, size_t j) const { return d[i + j * nr]; }
  double* ptr() { return d.data(); }
  const double* ptr() const { return d.data(); }
  size_t size() const { return d.size(); }
};

struct Cube {
  size_t nr = 0, nc = 0, ns = 0;
  std::vector<double> d;
  Cube() {}
  Cube(size_t r, size_t c, size_t s) : nr(r), nc(c), ns(s), d(r * c * s, 0.0) {}
  void zeros(size_t r, size_t c, size_t s) { nr = r; nc = c; ns = s; d.assign(r * c * s, 0.0); }
  double& operator()(size_t i, size_t j, size_t s) { return d[i + j * nr + s * nr * nc]; }
  double operator()(size_t i, size_t j, size_t s) const { return d[i + j * nr + s * nr * nc]; }
  double* slice(size_t s) { return d.data() + s * nr * nc; }
  const double* slice(size_t s) const { return d.data() + s * nr * nc; }
  double* ptr() { return d.data(); }
  const double* ptr() const { return d.data(); }
};

// C = A^T * diag(w) * B, A: (n x p), B: (n x q)  ->  (p x q)
inline Mat atwb(const Mat& A, const Vec& w, const Mat& B) {
  if (A.nr != B.nr || w.size() != A.nr) throw std::logic_error("atwb: size mismatch");
  Mat C(A.nc, B.nc);
  for (size_t j = 0; j < B.nc; j++)
    for (size_t i = 0; i < A.nc; i++) {
      double s = 0.0;
      for (size_t k = 0; k < A.nr; k++) s += A(k, i) * w[k] * B(k, j);
      C(i, j) = s;
    }
  return C;
}

}  // namespace hfb
