/*
 * k_perm_common.cuh -- helpers shared by the data-movement kernels (k_perm.cu, k_perm_vec.cu).
 */
#ifndef NNSB_K_PERM_COMMON_CUH
#define NNSB_K_PERM_COMMON_CUH

#include <cstdint>

namespace nnsb {

/* exact q / d for q < 2^31 with one 64-bit multiply: m = ceil (2^(31+s) / d), s = ceil (log2 d) */
struct FastDiv
{
  uint32_t d, m, s;
};

static inline FastDiv
make_fastdiv (uint32_t d)
{
  FastDiv f;
  f.d = d ? d : 1;
  f.s = 0;
  while ((1ull << f.s) < f.d)
    f.s++;
  f.m = (uint32_t) (((1ull << (31 + f.s)) + f.d - 1) / f.d);
  return f;
}

__device__ __forceinline__ void
divmod (uint32_t q, const FastDiv &f, uint32_t &quot, uint32_t &rem)
{
  quot = (uint32_t) (((uint64_t) q * f.m) >> (31 + f.s));
  rem = q - quot * f.d;
}

}  // namespace nnsb
#endif
