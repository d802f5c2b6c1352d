/*
 * k_clamp.cu -- K5: clamp.  Replaces gst_tensor_transform_clamp
 * (gsttensor_transform.c:1632-1658): x -> double -> CLAMP (x, min, max) -> same type, with
 * the C conversion rules of tensor_data.c (double -> unsigned goes through the signed type).
 * The comparison is done in fp64 for every type so the result is bit-identical to the
 * reference, including int64/uint64 above 2^53 (which lose precision there too).
 * Algorithmic bytes per element: 2 * sizeof(T).
 */
#include <climits>

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

/* 8/16/32-bit integers: every value is exact in fp64, so `x < lo` is `x < ceil (lo)` and
 * `x > hi` is `x > floor (hi)` -- two integer compares and two selects per element instead of
 * two conversions and two fp64 compares (uint8 1080p: 3.0 -> 5+ TB/s).  The replacement values
 * are the reference's own casts of the bounds (a fractional or out-of-range bound wraps the
 * way tensor_data.c does it). */
template <typename T> struct ClampInt
{
  /* not uint32: an unchanged value >= 2^31 still passes through the reference's
   * double -> int32 -> uint32 cast and comes out as 0x80000000 */
  static constexpr bool fast = !IsFloat<T>::v && sizeof (T) <= 4 && TypeId<T>::id != NNSB_FLOAT16
      && TypeId<T>::id != NNSB_UINT32;
  typedef long long W;
  W a, b;                       /* x < a -> lo_v ; x > b -> hi_v */
  T lo_v, hi_v;
  __device__ __forceinline__ void init (double lo, double hi)
  {
    const double big = 9.0e18;
    a = (lo != lo || lo <= -big) ? LLONG_MIN : (lo >= big ? LLONG_MAX : (W) ceil (lo));
    b = (hi != hi || hi >= big) ? LLONG_MAX : (hi <= -big ? LLONG_MIN : (W) floor (hi));
    lo_v = CCast<T, double>::run (lo);
    hi_v = CCast<T, double>::run (hi);
  }
  __device__ __forceinline__ T run (T x) const
  {
    if constexpr (sizeof (T) == 4) {
      const W xi = (W) x;
      return (xi > b) ? hi_v : ((xi < a) ? lo_v : x);
    } else {
      /* 8/16-bit: thresholds saturated to int never change the outcome */
      const int ai = a < (W) INT_MIN ? INT_MIN : (a > (W) INT_MAX ? INT_MAX : (int) a);
      const int bi = b < (W) INT_MIN ? INT_MIN : (b > (W) INT_MAX ? INT_MAX : (int) b);
      const int xi = (int) x;
      return (xi > bi) ? hi_v : ((xi < ai) ? lo_v : x);
    }
  }
};

/* 1-byte integers, four at a time with the byte-SIMD compares: uniform thresholds, three
 * cases per side (no byte / every byte / compare) decided once per thread */
template <typename T> struct ClampBytes
{
  uint32_t a4, b4, lo4, hi4;    /* thresholds and replacement values, replicated */
  uint32_t lt_const, gt_const;  /* 0: compare; 1: never; 2: always */
  __device__ __forceinline__ void init (const ClampInt<T> &ci)
  {
    constexpr long long tmin = IsSigned<T>::v ? -128 : 0, tmax = IsSigned<T>::v ? 127 : 255;
    lt_const = ci.a <= tmin ? 1u : (ci.a > tmax ? 2u : 0u);     /* x < a */
    gt_const = ci.b >= tmax ? 1u : (ci.b < tmin ? 2u : 0u);     /* x > b */
    a4 = 0x01010101u * (uint32_t) (uint8_t) (lt_const ? 0 : ci.a);
    b4 = 0x01010101u * (uint32_t) (uint8_t) (gt_const ? 0 : ci.b);
    lo4 = 0x01010101u * (uint32_t) (uint8_t) ci.lo_v;
    hi4 = 0x01010101u * (uint32_t) (uint8_t) ci.hi_v;
  }
  __device__ __forceinline__ uint32_t run (uint32_t w) const
  {
    uint32_t gt, lt;
    if constexpr (IsSigned<T>::v) {
      gt = gt_const ? (gt_const == 2u ? 0xffffffffu : 0u) : __vcmpgts4 (w, b4);
      lt = lt_const ? (lt_const == 2u ? 0xffffffffu : 0u) : __vcmplts4 (w, a4);
    } else {
      gt = gt_const ? (gt_const == 2u ? 0xffffffffu : 0u) : __vcmpgtu4 (w, b4);
      lt = lt_const ? (lt_const == 2u ? 0xffffffffu : 0u) : __vcmpltu4 (w, a4);
    }
    lt &= ~gt;                  /* CLAMP tests the upper bound first */
    return (w & ~(gt | lt)) | (gt & hi4) | (lt & lo4);
  }
};

template <typename T, int U>
__global__ void __launch_bounds__ (CLAMP_THREADS)
clamp_kernel (const T *__restrict__ in, T *__restrict__ out, size_t n, size_t n_vec, double lo,
    double hi)
{
  constexpr int V = 32 / sizeof (T) > 16 ? 16 : 32 / sizeof (T);
  const size_t tid = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
  const size_t nthreads = (size_t) gridDim.x * blockDim.x;
  ClampInt<T> ci;
  if constexpr (ClampInt<T>::fast)
    ci.init (lo, hi);
  ClampBytes<T> cb;
  if constexpr (sizeof (T) == 1)
    cb.init (ci);
  auto one = [&] (T x) {
    if constexpr (ClampInt<T>::fast) return ci.run (x);
    else return clamp_one<T> (x, lo, hi);
  };
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
        if constexpr (sizeof (T) == 1) {
          uint32_t *w = reinterpret_cast<uint32_t *> (x[u].e);
#pragma unroll
          for (int k = 0; k < V / 4; k++)
            w[k] = cb.run (w[k]);
        } else {
#pragma unroll
          for (int k = 0; k < V; k++)
            x[u].e[k] = one (x[u].e[k]);
        }
        store_vec (out + vu * V, x[u]);
      }
    }
  }
  for (size_t e = n_vec * V + tid; e < n; e += nthreads) {
    T xv;
    ld_stream<sizeof (T)> (&xv, in + e);
    xv = one (xv);
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
