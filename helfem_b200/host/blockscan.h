// Host-side helpers of the dense-matrix operator boundary (ChanEngine::dense_apply, Grid2D::eval_fxc): the reference's operators
// take and return dense arma::mat; here only the angular blocks that can be non-zero cross PCIe, so every call scans P for its
// non-zero blocks and zero-fills the result.  Both are plain memory streaming; this file makes them run at DRAM speed
// (column-order streaming, vector bit tests, non-temporal stores) since they are the larger part of an end-to-end call.
#pragma once
#include <cstddef>

namespace hfb {

// nz[i * nang + l] = 1 iff block (i, l) of any of the nmat matrices has an entry that is not +-0.  Rows of angular function i:
// r0[i] .. r0[i] + cnt[i] - 1 (same for columns).  Column-major, leading dimension ld.
void scan_nonzero_blocks(const double* const* mats, int nmat, int ld, int nang, const int* r0, const int* cnt, unsigned char* nz);

// M(0:n, 0:n) = 0 with non-temporal stores (column-major, leading dimension ld)
void zero_fill(double* M, int ld, int n);

}  // namespace hfb
