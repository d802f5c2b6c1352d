/*
 * k_perm_common.cuh -- helpers shared by the data-movement kernels (k_perm.cu, k_perm_vec.cu).
 */
#ifndef NNSB_K_PERM_COMMON_CUH
#define NNSB_K_PERM_COMMON_CUH

#include <cstdint>

#include <cuda_runtime.h>

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

/* One dimension of a canonical permutation problem: extent, input stride, output stride
 * (strides in elements; the output is dense, dims are listed in output order). */
struct CDim
{
  uint64_t size, in, out;
};

/* Canonical form of a permutation: extent-1 dims dropped, the batch appended (batch > 1),
 * dims that are adjacent on both sides merged.  If the innermost dim is the same on both
 * sides (a pixel / a channel group that moves as a block), a power-of-two part of it is folded
 * into the element: es grows up to max_es as far as `alignment` (OR of the base addresses and
 * the frame size in bytes) allows, and all strides shrink accordingly.
 * Returns the dim count; *es_io is updated. */
static inline int
canonicalise_perm (const uint32_t *size, const uint64_t *in_stride, int rank, uint64_t frame_elems,
    uint32_t batch, int max_es, uint64_t alignment, CDim *dims, int *es_io)
{
  int n = 0;
  uint64_t ostride = 1;
  for (int i = 0; i < rank; i++) {
    if (size[i] != 1) {
      dims[n].size = size[i];
      dims[n].in = in_stride[i];
      dims[n].out = ostride;
      n++;
    }
    ostride *= size[i];
  }
  if (batch > 1) {
    dims[n].size = batch;
    dims[n].in = frame_elems;
    dims[n].out = frame_elems;
    n++;
  }
  int m = 0;
  for (int i = 0; i < n; i++) {
    if (m > 0 && dims[m - 1].in * dims[m - 1].size == dims[i].in
        && dims[m - 1].out * dims[m - 1].size == dims[i].out)
      dims[m - 1].size *= dims[i].size;
    else
      dims[m++] = dims[i];
  }
  n = m;
  if (n >= 2 && dims[0].in == 1 && dims[0].out == 1) {
    int es = *es_io;
    uint64_t f = 1;
    while (dims[0].size % (2 * f) == 0 && (int) (2 * f) * es <= max_es
        && alignment % (2 * f * es) == 0)
      f *= 2;
    if (f > 1) {
      dims[0].size /= f;
      for (int i = 1; i < n; i++) {
        dims[i].in /= f;
        dims[i].out /= f;
      }
      *es_io = es * (int) f;
      if (dims[0].size == 1) {
        for (int i = 1; i < n; i++)
          dims[i - 1] = dims[i];
        n--;
      }
    }
  }
  return n;
}

/* k_perm_slab.cu: -1 = not suited (short runs, sparse input, ...), else the launch result;
 * dry_run only answers the question */
struct LaunchCfg;
int launch_permute_slab (const CDim *dims, int n, int es, const void *in, void *out,
    uint64_t total_elems, const LaunchCfg &cfg, cudaStream_t stream, bool dry_run = false);

/* k_perm_pix.cu: transpose 0:2:1:3 of packed 3-byte pixels; -1 = not that shape */
int launch_pixel_transpose (const CDim *dims, int n, int es, const void *in, void *out,
    cudaStream_t stream, bool dry_run = false);

}  // namespace nnsb
#endif
