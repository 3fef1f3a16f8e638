// Which rank holds which integral family of a sharded atomic / diatomic context (hfb200_set_shard).
//
// The operators are sums over integral families (atomic: L of the Legendre expansion, atomic/TwoDBasis.cpp:973-1073; diatomic:
// (L,|M|) of the Neumann expansion, diatomic/basis.cpp:1570-1650) and the families differ in cost by more than an order of
// magnitude, so they are dealt to the ranks by predicted cost (longest processing time first), not round-robin: r mod N leaves
// the first rank 3.6 % above the mean at 8 ranks of configs[3], which is 3.6 % of the whole strong-scaling step.
// Cost model: the number of exchange output blocks (j <= k) the family contributes for an m-diagonal density (what an SCF
// produces): per channel (L, M) the angular functions j that have a coupling partner i (m_j - m_i = M, triangle rule widened by
// `reach`, l_j + l_i + L even) form, per m, n(m) functions and n (n + 1) / 2 blocks.  Both the host classes (which integrals to
// compute) and the C ABI (which to upload and contract) call this one function, so every rank derives the same map.
#pragma once
#include <algorithm>
#include <cstdlib>
#include <numeric>
#include <vector>

namespace hfb {

// famL[r], famM[r]: L and |M| of family r; famM[r] < 0: the family holds every M with |M| <= L (atomic).  reach: 0 (Gaunt
// coefficients) or 2 (the diatomic cos^2-modified coefficients couple L to l_j + l_i +- 2 as well).
inline std::vector<int> family_owner(const std::vector<int>& lval, const std::vector<int>& mval, const std::vector<int>& famL,
                                     const std::vector<int>& famM, int reach, int nranks) {
  const int nfam = (int)famL.size(), nang = (int)lval.size();
  std::vector<int> owner((size_t)nfam, 0);
  if (nranks <= 1 || nfam == 0) return owner;
  int mlo = 0, mhi = 0;
  for (int a = 0; a < nang; a++) { mlo = std::min(mlo, mval[a]); mhi = std::max(mhi, mval[a]); }
  std::vector<double> cost((size_t)nfam, 0.0);
#pragma omp parallel for schedule(dynamic)
  for (int r = 0; r < nfam; r++) {
    const int L = famL[r];
    std::vector<int> cnt((size_t)(mhi - mlo + 1));
    double c = 0.0;
    const int Mfirst = famM[r] < 0 ? -L : -famM[r], Mlast = famM[r] < 0 ? L : famM[r], Mstep = famM[r] < 0 ? 1 : std::max(1, 2 * famM[r]);
    for (int M = Mfirst; M <= Mlast; M += Mstep) {
      std::fill(cnt.begin(), cnt.end(), 0);
      for (int j = 0; j < nang; j++) {
        bool has = false;
        for (int i = 0; i < nang && !has; i++)
          has = mval[j] - mval[i] == M && std::abs(lval[j] - lval[i]) - reach <= L && L <= lval[j] + lval[i] + reach &&
                ((lval[j] + lval[i] + L) & 1) == 0;
        if (has) cnt[(size_t)(mval[j] - mlo)]++;
      }
      for (int n : cnt) c += 0.5 * n * (n + 1);
    }
    cost[r] = c;
  }
  std::vector<int> order((size_t)nfam);
  std::iota(order.begin(), order.end(), 0);
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cost[a] > cost[b]; });
  std::vector<double> load((size_t)nranks, 0.0);
  for (int r : order) {
    const int k = (int)(std::min_element(load.begin(), load.end()) - load.begin());  // ties: the lowest rank
    owner[r] = k;
    load[k] += cost[r];
  }
  return owner;
}

}  // namespace hfb
