/*
 * k_clamp.cu -- K5: clamp.  Replaces gst_tensor_transform_clamp
 * (gsttensor_transform.c:1632-1658): x -> double -> CLAMP (x, min, max) -> same type, with
 * the C conversion rules of tensor_data.c (double -> unsigned goes through the signed type).
 * The comparison is done in fp64 for every type so the result is bit-identical to the
 * reference, including int64/uint64 above 2^53 (which lose precision there too).
 * Algorithmic bytes per element: 2 * sizeof(T).
 */
#include "nnsb_device.cuh"
#include "nnsb_kernels.h"

namespace nnsb {

constexpr int CLAMP_THREADS = 256;

template <typename T> __device__ __forceinline__ T
clamp_one (T x, double lo, double hi)
{
  double d = CCast<double, T>::run (x);
  /* glib CLAMP: (x > high) ? high : ((x < low) ? low : x); NaN falls through unchanged */
  d = (d > hi) ? hi : ((d < lo) ? lo : d);
  return CCast<T, double>::run (d);
}

template <typename T, int U>
__global__ void __launch_bounds__ (CLAMP_THREADS)
clamp_kernel (const T *__restrict__ in, T *__restrict__ out, size_t n, size_t n_vec, double lo,
    double hi)
{
  constexpr int V = 32 / sizeof (T) > 16 ? 16 : 32 / sizeof (T);
  const size_t tid = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
  const size_t nthreads = (size_t) gridDim.x * blockDim.x;
  for (size_t v = tid; v < n_vec; v += nthreads * U) {
    Vec<T, V> x[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t vu = v + (size_t) u * nthreads;
      if (vu < n_vec)
        load_vec (x[u], in + vu * V);
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t vu = v + (size_t) u * nthreads;
      if (vu < n_vec) {
#pragma unroll
        for (int k = 0; k < V; k++)
          x[u].e[k] = clamp_one<T> (x[u].e[k], lo, hi);
        store_vec (out + vu * V, x[u]);
      }
    }
  }
  for (size_t e = n_vec * V + tid; e < n; e += nthreads) {
    T xv;
    ld_stream<sizeof (T)> (&xv, in + e);
    xv = clamp_one<T> (xv, lo, hi);
    st_stream<sizeof (T)> (out + e, &xv);
  }
}

template <typename T> static int
launch_clamp_t (const void *in, void *out, size_t n, double lo, double hi, const LaunchCfg &cfg,
    cudaStream_t stream)
{
  constexpr int V = 32 / sizeof (T) > 16 ? 16 : 32 / sizeof (T);
  constexpr int U = 4;
  if (n == 0)
    return 0;
  const bool aligned = ((uintptr_t) in % (V * sizeof (T)) == 0)
      && ((uintptr_t) out % (V * sizeof (T)) == 0);
  const size_t n_vec = aligned ? n / V : 0;
  const size_t work = n_vec ? n_vec : n;
  const int bps = cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : 256;   /* see k_chain.cuh */
  size_t blocks = (work + (size_t) CLAMP_THREADS * U - 1) / ((size_t) CLAMP_THREADS * U);
  const size_t cap = (size_t) cfg.sm_count * bps;
  if (blocks > cap)
    blocks = cap;
  if (blocks == 0)
    blocks = 1;
  clamp_kernel<T, U><<<(unsigned int) blocks, CLAMP_THREADS, 0, stream>>> ((const T *) in,
      (T *) out, n, n_vec, lo, hi);
  return (int) cudaGetLastError ();
}

int launch_clamp (int type, const void *in, void *out, size_t n, double lo, double hi,
    const LaunchCfg &cfg, cudaStream_t stream)
{
  switch (type) {
    case NNSB_INT32: return launch_clamp_t<int32_t> (in, out, n, lo, hi, cfg, stream);
    case NNSB_UINT32: return launch_clamp_t<uint32_t> (in, out, n, lo, hi, cfg, stream);
    case NNSB_INT16: return launch_clamp_t<int16_t> (in, out, n, lo, hi, cfg, stream);
    case NNSB_UINT16: return launch_clamp_t<uint16_t> (in, out, n, lo, hi, cfg, stream);
    case NNSB_INT8: return launch_clamp_t<int8_t> (in, out, n, lo, hi, cfg, stream);
    case NNSB_UINT8: return launch_clamp_t<uint8_t> (in, out, n, lo, hi, cfg, stream);
    case NNSB_FLOAT64: return launch_clamp_t<double> (in, out, n, lo, hi, cfg, stream);
    case NNSB_FLOAT32: return launch_clamp_t<float> (in, out, n, lo, hi, cfg, stream);
    case NNSB_INT64: return launch_clamp_t<int64_t> (in, out, n, lo, hi, cfg, stream);
    case NNSB_UINT64: return launch_clamp_t<uint64_t> (in, out, n, lo, hi, cfg, stream);
    case NNSB_FLOAT16: return launch_clamp_t<__half> (in, out, n, lo, hi, cfg, stream);
    default: return (int) cudaErrorInvalidValue;
  }
}

}  // namespace nnsb
