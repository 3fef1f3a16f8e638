#include "blockscan.h"
#include <cstdint>
#include <cstring>
#include <vector>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace hfb {

namespace {

// any entry whose bit pattern is not +-0 (NaN and denormals count as non-zero, like v != 0.0 does)
#if defined(__x86_64__)
__attribute__((target_clones("avx512f", "avx2", "default")))
#endif
bool any_nonzero(const double* p, size_t n) {
  const uint64_t* q = reinterpret_cast<const uint64_t*>(p);
  uint64_t acc = 0;
  for (size_t i = 0; i < n; i++) acc |= q[i];
  return (acc << 1) != 0;
}

#if defined(__x86_64__)
__attribute__((target("avx2"))) void stream_zero_avx2(double* p, size_t n) {
  size_t i = 0;
  while (i < n && (reinterpret_cast<uintptr_t>(p + i) & 31)) p[i++] = 0.0;
  const __m256d z = _mm256_setzero_pd();
  for (; i + 4 <= n; i += 4) _mm256_stream_pd(p + i, z);
  for (; i < n; i++) p[i] = 0.0;
  _mm_sfence();  // make the write-combining stores of this thread globally visible before it moves on
}
#endif

void stream_zero(double* p, size_t n) {
#if defined(__x86_64__)
  static const bool avx2 = __builtin_cpu_supports("avx2");
  if (avx2) { stream_zero_avx2(p, n); return; }
#endif
  memset(p, 0, n * sizeof(double));
}

}  // namespace

void scan_nonzero_blocks(const double* const* mats, int nmat, int ld, int nang, const int* r0, const int* cnt, unsigned char* nz) {
  // one angular COLUMN block per task: its columns are streamed top to bottom (contiguous memory), a block row is skipped as
  // soon as it is known to be non-zero
#pragma omp parallel
  {
    std::vector<unsigned char> loc((size_t)nang);
#pragma omp for schedule(dynamic, 1)
    for (int l = 0; l < nang; l++) {
      std::fill(loc.begin(), loc.end(), 0);
      for (int s = 0; s < nmat; s++)
        for (int b = 0; b < cnt[l]; b++) {
          const double* col = mats[s] + (size_t)(r0[l] + b) * ld;
          for (int i = 0; i < nang; i++)
            if (!loc[i] && any_nonzero(col + r0[i], (size_t)cnt[i])) loc[i] = 1;
        }
      for (int i = 0; i < nang; i++) nz[(size_t)i * nang + l] = loc[i];
    }
  }
}

void zero_fill(double* M, int ld, int n) {
#pragma omp parallel for schedule(static)
  for (int c = 0; c < n; c++) stream_zero(M + (size_t)c * ld, (size_t)n);
}

}  // namespace hfb
