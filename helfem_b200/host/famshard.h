// Which rank holds which integral family of a sharded atomic / diatomic context (hfb200_set_shard).
//
// The operators are sums over integral families (atomic: L of the Legendre expansion, atomic/TwoDBasis.cpp:973-1073; diatomic:
// (L,|M|) of the Neumann expansion in lm_map order, diatomic/basis.cpp:1570-1650).  Family r lives on rank r mod N: neighbours in
// that order have similar cost, so the interleave balances well (the first rank ends up ~3.6 % above the mean at 8 ranks of
// configs[3]).  Dealing the families by a PREDICTED cost, longest first, was tried in round 2 (cost model: exchange output blocks
// of the family for an m-diagonal density) and measured worse -- 34.0 instead of 31.6 ms per build at N = 4: the model misses the
// slot / W-block structure that sets the real cost -- so the map stays the plain interleave.  Both the host classes (which
// integrals to compute) and the C ABI (which to upload and contract) call this one function.
#pragma once
#include <vector>

namespace hfb {

inline std::vector<int> family_owner(int nfam, int nranks) {
  std::vector<int> owner((size_t)(nfam > 0 ? nfam : 0), 0);
  if (nranks > 1)
    for (int r = 0; r < nfam; r++) owner[(size_t)r] = r % nranks;
  return owner;
}

}  // namespace hfb
