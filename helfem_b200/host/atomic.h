// Host-side mirror of the reference's `atomic` operator class.
//
//   helfem::atomic::basis::TwoDBasis   /root/reference/src/atomic/TwoDBasis.h:38-226
//
// Same method names, argument meaning, index conventions (global index iang*Nrad + irad,
// atomic/TwoDBasis.cpp:290-300), sign conventions (exchange returns -K) and error behaviour.
// coulomb / exchange run on the GPU through the C ABI; there is no host implementation here.
#pragma once
#include <stdexcept>
#include "fem.h"

struct hfb200_ctx;

namespace hfb {
namespace atomic {

class TwoDBasis {
 public:
  TwoDBasis() {}
  /// Reference ctor: atomic/TwoDBasis.cpp:38-60 (point nucleus only).
  TwoDBasis(int Z, const fem::PolyBasis& poly, bool zeroder, int n_quad, const Vec& bval, int taylor_order,
            const std::vector<int>& lval, const std::vector<int>& mval, int device = 0);
  ~TwoDBasis();
  TwoDBasis(const TwoDBasis&) = delete;
  TwoDBasis& operator=(const TwoDBasis&) = delete;

  size_t Nbf() const { return lval.size() * radial.Nbf(); }
  size_t Nrad() const { return radial.Nbf(); }
  size_t Nang() const { return lval.size(); }
  const std::vector<int>& get_lval() const { return lval; }
  const std::vector<int>& get_mval() const { return mval; }
  const fem::RadialBasis& get_radial() const { return radial; }

  /// one-electron matrices in the product basis (atomic/TwoDBasis.cpp:300-420): block diagonal in (l,m)
  Mat overlap() const;
  Mat kinetic() const;
  Mat nuclear() const;

  /// atomic/TwoDBasis.cpp:666-739
  void compute_tei(bool exchange);
  /// J = coulomb(P), atomic/TwoDBasis.cpp:817-955
  Mat coulomb(const Mat& P) const;
  /// -K = exchange(P), atomic/TwoDBasis.cpp:957-1140
  Mat exchange(const Mat& P) const;
  void coulomb(const double* P, double* J) const;   // raw n x n column-major buffers, n = Nbf()
  void exchange(const double* P, double* K) const;

  const std::vector<Mat>& get_disjoint_L() const { return disjoint_L; }
  const std::vector<Mat>& get_disjoint_m1L() const { return disjoint_m1L; }
  const std::vector<Mat>& get_prim_tei() const { return prim_tei; }
  const std::vector<Mat>& get_prim_ktei() const { return prim_ktei; }
  int get_nL() const { return nL_stored; }
  /// multi-GPU: this object contracts only the integral families r % nranks == rank (see hfb200_set_shard)
  void set_shard(int rank, int nranks);  // throws std::logic_error once the device context exists (it is sharded at creation)
  hfb200_ctx* context() const;

 private:
  Mat radial_assemble(int which) const;
  int Z = 0, device = 0;
  std::vector<int> lval, mval;
  fem::RadialBasis radial;
  int nL_stored = 0;
  std::vector<Mat> disjoint_L, disjoint_m1L, prim_tei, prim_ktei;  // index [L*Nel+iel]
  int shard_rank = 0, shard_n = 1;
  mutable hfb200_ctx* ctx = nullptr;
  mutable bool tei_uploaded = false;
};

}  // namespace atomic
}  // namespace hfb
