/*
 * k_stand.cuh -- K6/K7: stand (standardisation).
 *
 * Replaces gst_tensor_transform_stand (gsttensor_transform.c:1512-1619) and the statistics
 * helpers gst_tensor_data_raw_average[_per_channel] / raw_std[_per_channel]
 * (tensor_data.c:314-493).  Reference semantics, all in fp64 whatever the input type:
 *     mean  = running mean over the tensor (or over each channel c = e % dim[0])
 *     std   = sqrt (sum ((x - mean)^2) / N), or 1e-10 when that sum is exactly 0
 *     default    : out = (Out) fabs ((x - mean) / std)
 *     dc-average : out = (Out) (x - mean)
 * The reference makes three sequential passes; a parallel reduction cannot reproduce its
 * summation order bit for bit, so stand is the one mode with a tolerance (tests: 1e-6 rel).
 *
 * GPU scheme (3 launches per call, independent of the batch size):
 *   1. stats   : one read of the input.  Every thread accumulates sum (x-K) and sum (x-K)^2
 *                in fp64 with the shift K = first element of the channel (kills the
 *                cancellation of the one-pass variance formula); block partials go to a
 *                scratch array in global memory, in a fixed order (bit-reproducible).
 *   2. finalize: one warp per (tensor, channel) folds the block partials -> {mean, std}.
 *   3. apply   : second read (L2-resident when a tensor fits), fp64 math, one write.
 *                (x - mean) * (1 / std) with the correctly rounded reciprocal: <= 1.5 ulp of
 *                fp64 from the reference's division, 1 DMUL instead of a ~30 instruction DDIV.
 * For 1..4 channels (gray / RGB / RGBA, and the not-per-channel case) the kernels are
 * specialised on the channel count (stand_*_kernel_n): ~40-60 registers instead of ~100.
 * Measured (profiles/): 4.0-4.4 TB/s algorithmic on 1080p f32 x 64, i.e. the 12 B/element
 * two-read scheme runs at the HBM roof; a per-frame fused stats+apply pipeline relying on L2
 * residency was tried and was slower (18 us per 25 MB frame launch), so it is not kept.
 * Thread/channel mapping: the per-tensor thread stride is rounded down to a multiple of
 * nch, so the channel of every element a thread touches is fixed for the whole kernel and
 * lives in registers.  Algorithmic bytes per element: sizeof(In) + sizeof(Out).
 */
#ifndef NNSB_K_STAND_CUH
#define NNSB_K_STAND_CUH

#include "nnsb_device.cuh"
#include "nnsb_kernels.h"

namespace nnsb {

constexpr int STAND_THREADS = 256;
constexpr uint32_t STAND_NCH_FAST = 8;    /* vectorised statistics path */

template <typename T> struct StandVec
{
  static constexpr int V = (32 / sizeof (T)) > 16 ? 16 : (32 / sizeof (T));
};

/* how the statistics of one launch are laid out in scratch */
struct StandGeom
{
  uint32_t blocks_x;            /* blocks per tensor */
  uint32_t nch;
  uint32_t batch;
  int path;                     /* 0 = A (nch <= 8, vector), 1 = B (strided scalar), 2 = C (column) */
};

__host__ __device__ inline size_t
stand_partial_index (const StandGeom &g, uint32_t f, uint32_t b, uint32_t c)
{
  return (((size_t) f * g.blocks_x + b) * g.nch + c) * 2;
}

/* ---------------------------------------------------------------------------------- */
/* 1. statistics                                                                        */
/* ---------------------------------------------------------------------------------- */
template <typename In, bool NEED_STD>
__global__ void __launch_bounds__ (STAND_THREADS)
stand_stats_kernel (const In *__restrict__ in, size_t frame_elems, StandGeom g,
    double *__restrict__ partial)
{
  constexpr int V = StandVec<In>::V;
  const uint32_t f = blockIdx.y;
  const In *base = in + (size_t) f * frame_elems;
  const uint32_t nch = g.nch;
  const size_t t_total = (size_t) gridDim.x * STAND_THREADS;
  const size_t tid_f = (size_t) blockIdx.x * STAND_THREADS + threadIdx.x;

  if (g.path == 2) {
    /* C: more channels than threads: one thread walks whole columns; the result is final,
     * stored as partial of block 0 (blocks_x == 1 in this geometry) */
    const size_t rows = frame_elems / nch;
    for (size_t c = tid_f; c < nch; c += t_total) {
      const double k0 = CCast<double, In>::run (base[c]);
      double s1 = 0.0, s2 = 0.0;
      for (size_t i = 0; i < rows; i++) {
        const double d = __dsub_rn (CCast<double, In>::run (base[i * nch + c]), k0);
        s1 += d;
        if (NEED_STD) s2 = fma (d, d, s2);
      }
      double *p = partial + stand_partial_index (g, f, 0, (uint32_t) c);
      p[0] = s1;
      p[1] = s2;
    }
    return;
  }

  const size_t t_eff = (t_total / nch) * nch;   /* stride: a multiple of nch */
  const bool active = tid_f < t_eff;

  __shared__ double sm[2][STAND_NCH_FAST][STAND_THREADS];

  if (g.path == 0) {
    /* A: vector loads; lane k of every vector of this thread has channel (c0 + k) % nch */
    const bool aligned = ((uintptr_t) base % (V * sizeof (In))) == 0;
    const size_t n_vec = aligned ? frame_elems / V : 0;
    double a1[V], a2[V], kk[V];
#pragma unroll
    for (int k = 0; k < V; k++) {
      a1[k] = 0.0;
      a2[k] = 0.0;
      kk[k] = 0.0;
    }
    const uint32_t c0 = (uint32_t) ((tid_f * V) % nch);
    if (active) {
#pragma unroll
      for (int k = 0; k < V; k++)
        kk[k] = CCast<double, In>::run (base[(c0 + k) % nch]);
      for (size_t v = tid_f; v < n_vec; v += t_eff * 2) {
        Vec<In, V> x0, x1;
        const size_t v1 = v + t_eff;
        load_vec (x0, base + v * V);
        if (v1 < n_vec)
          load_vec (x1, base + v1 * V);
#pragma unroll
        for (int k = 0; k < V; k++) {
          const double d = __dsub_rn (CCast<double, In>::run (x0.e[k]), kk[k]);
          a1[k] += d;
          if (NEED_STD) a2[k] = fma (d, d, a2[k]);
        }
        if (v1 < n_vec) {
#pragma unroll
          for (int k = 0; k < V; k++) {
            const double d = __dsub_rn (CCast<double, In>::run (x1.e[k]), kk[k]);
            a1[k] += d;
            if (NEED_STD) a2[k] = fma (d, d, a2[k]);
          }
        }
      }
    }
    /* rotate lane accumulators into per-channel columns owned by this thread */
    for (uint32_t c = 0; c < nch; c++) {
      sm[0][c][threadIdx.x] = 0.0;
      sm[1][c][threadIdx.x] = 0.0;
    }
    if (active) {
#pragma unroll
      for (int k = 0; k < V; k++) {
        const uint32_t c = (c0 + k) % nch;
        sm[0][c][threadIdx.x] += a1[k];
        sm[1][c][threadIdx.x] += a2[k];
      }
      /* scalar remainder (and everything when the tensor base is not vector aligned) */
      const size_t e0 = n_vec * V + tid_f;
      if (e0 < frame_elems) {
        const uint32_t c = (uint32_t) (e0 % nch);
        const double k0 = CCast<double, In>::run (base[c]);
        double s1 = 0.0, s2 = 0.0;
        for (size_t e = e0; e < frame_elems; e += t_eff) {
          const double d = __dsub_rn (CCast<double, In>::run (base[e]), k0);
          s1 += d;
          if (NEED_STD) s2 = fma (d, d, s2);
        }
        sm[0][c][threadIdx.x] += s1;
        sm[1][c][threadIdx.x] += s2;
      }
    }
    __syncthreads ();
    /* fixed-order tree per channel */
    for (int s = STAND_THREADS / 2; s > 0; s >>= 1) {
      if ((int) threadIdx.x < s) {
        for (uint32_t c = 0; c < nch; c++) {
          sm[0][c][threadIdx.x] += sm[0][c][threadIdx.x + s];
          sm[1][c][threadIdx.x] += sm[1][c][threadIdx.x + s];
        }
      }
      __syncthreads ();
    }
    if (threadIdx.x < nch) {
      double *p = partial + stand_partial_index (g, f, blockIdx.x, threadIdx.x);
      p[0] = sm[0][threadIdx.x][0];
      p[1] = sm[1][threadIdx.x][0];
    }
    return;
  }

  /* B: every thread reads elements tid_f, tid_f + t_eff, ... : one fixed channel */
  {
    double (*col)[STAND_THREADS] = sm[0];       /* col[0] = s1, col[1] = s2 per thread */
    double s1 = 0.0, s2 = 0.0;
    if (active && tid_f < frame_elems) {
      const uint32_t c = (uint32_t) (tid_f % nch);
      const double k0 = CCast<double, In>::run (base[c]);
      for (size_t e = tid_f; e < frame_elems; e += t_eff) {
        const double d = __dsub_rn (CCast<double, In>::run (base[e]), k0);
        s1 += d;
        if (NEED_STD) s2 = fma (d, d, s2);
      }
    }
    col[0][threadIdx.x] = s1;
    col[1][threadIdx.x] = s2;
    __syncthreads ();
    /* thread j folds, in thread order, the threads of this block that own channel c */
    const size_t block_first = (size_t) blockIdx.x * STAND_THREADS;
    for (uint32_t c = threadIdx.x; c < nch; c += STAND_THREADS) {
      /* smallest t in [0, 256) with (block_first + t) % nch == c */
      const uint32_t off = (uint32_t) (block_first % nch);
      uint32_t t = (c >= off) ? (c - off) : (c + nch - off);
      double r1 = 0.0, r2 = 0.0;
      for (; t < STAND_THREADS; t += nch) {
        r1 += col[0][t];
        r2 += col[1][t];
      }
      double *p = partial + stand_partial_index (g, f, blockIdx.x, c);
      p[0] = r1;
      p[1] = r2;
    }
  }
}

/* ---------------------------------------------------------------------------------- */
/* 2. finalize: one warp per (tensor, channel)                                          */
/* ---------------------------------------------------------------------------------- */
template <typename In>
__global__ void __launch_bounds__ (STAND_THREADS)
stand_finalize_kernel (const In *__restrict__ in, size_t frame_elems, StandGeom g,
    const double *__restrict__ partial, double *__restrict__ stats, int need_std)
{
  const size_t warp = ((size_t) blockIdx.x * STAND_THREADS + threadIdx.x) / 32;
  const uint32_t lane = threadIdx.x & 31;
  const size_t total = (size_t) g.batch * g.nch;
  if (warp >= total)
    return;
  const uint32_t f = (uint32_t) (warp / g.nch), c = (uint32_t) (warp % g.nch);
  double s1 = 0.0, s2 = 0.0;
  for (uint32_t b = lane; b < g.blocks_x; b += 32) {
    const double *p = partial + stand_partial_index (g, f, b, c);
    s1 += p[0];
    s2 += p[1];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s1 += __shfl_xor_sync (0xffffffffu, s1, o);
    s2 += __shfl_xor_sync (0xffffffffu, s2, o);
  }
  if (lane == 0) {
    const double n = (double) (frame_elems / g.nch);
    const double k0 = CCast<double, In>::run (in[(size_t) f * frame_elems + c]);
    const double mean = k0 + s1 / n;
    double sd = 0.0;
    if (need_std) {
      const double var = (s2 - s1 * (s1 / n)) / n;
      /* tensor_data.c:438,488: std = (sum != 0) ? sqrt (sum) : 1e-10 */
      sd = (var > 0.0) ? sqrt (var) : 1e-10;
    }
    stats[((size_t) f * g.nch + c) * 2 + 0] = mean;
    stats[((size_t) f * g.nch + c) * 2 + 1] = sd;
  }
}

/* The same fold with a whole block per (tensor, channel): one tensor of 400 MB leaves ~9500
 * block partials, which a single warp walks for 39 us (296 dependent rounds); 256 threads take
 * 37 of them each, then a fixed tree -- still bit-reproducible, 4 us. */
template <typename In>
__global__ void __launch_bounds__ (STAND_THREADS)
stand_finalize_block_kernel (const In *__restrict__ in, size_t frame_elems, StandGeom g,
    const double *__restrict__ partial, double *__restrict__ stats, int need_std)
{
  const uint32_t f = blockIdx.x / g.nch, c = blockIdx.x % g.nch;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double s1 = 0.0, s2 = 0.0;
  for (uint32_t b = threadIdx.x; b < g.blocks_x; b += STAND_THREADS) {
    const double *p = partial + stand_partial_index (g, f, b, c);
    s1 += p[0];
    s2 += p[1];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s1 += __shfl_xor_sync (0xffffffffu, s1, o);
    s2 += __shfl_xor_sync (0xffffffffu, s2, o);
  }
  __shared__ double w1[STAND_THREADS / 32], w2[STAND_THREADS / 32];
  if (lane == 0) {
    w1[warp] = s1;
    w2[warp] = s2;
  }
  __syncthreads ();
  if (threadIdx.x == 0) {
    s1 = s2 = 0.0;
#pragma unroll
    for (int w = 0; w < STAND_THREADS / 32; w++) {
      s1 += w1[w];
      s2 += w2[w];
    }
    const double n = (double) (frame_elems / g.nch);
    const double k0 = CCast<double, In>::run (in[(size_t) f * frame_elems + c]);
    const double mean = k0 + s1 / n;
    double sd = 0.0;
    if (need_std) {
      const double var = (s2 - s1 * (s1 / n)) / n;
      sd = (var > 0.0) ? sqrt (var) : 1e-10;
    }
    stats[((size_t) f * g.nch + c) * 2 + 0] = mean;
    stats[((size_t) f * g.nch + c) * 2 + 1] = sd;
  }
}

/* ---------------------------------------------------------------------------------- */
/* 3. apply                                                                             */
/* ---------------------------------------------------------------------------------- */
/* One element.  default mode: |x - mean| / std, formed as (x - mean) * (1 / std) with the
 * correctly rounded reciprocal: at most ~1.5 ulp (3e-16 relative) away from the reference's
 * fp64 division -- far inside the 1e-6 tolerance stand carries anyway because the mean itself
 * is reduced in another order -- and 1 fp64 multiply instead of a ~30-instruction DDIV.
 * inf / NaN inputs and overflow behave like the division (inf * r = inf, NaN stays NaN). */
template <bool DEFAULT_MODE> __device__ __forceinline__ double
stand_one (double x, double mean, double sd, double rsd)
{
  (void) sd;
  const double d = __dsub_rn (x, mean);
  if (!DEFAULT_MODE)
    return d;
  return fabs (__dmul_rn (d, rsd));
}

template <typename In, typename Out>
__global__ void __launch_bounds__ (STAND_THREADS)
stand_apply_kernel (const In *__restrict__ in, Out *__restrict__ out, size_t frame_elems,
    StandGeom g, const double *__restrict__ stats, int default_mode)
{
  constexpr int VI = StandVec<In>::V, VO = StandVec<Out>::V;
  constexpr int V = VI < VO ? VI : VO;
  const uint32_t f = blockIdx.y;
  const In *ibase = in + (size_t) f * frame_elems;
  Out *obase = out + (size_t) f * frame_elems;
  const uint32_t nch = g.nch;
  const double *st = stats + (size_t) f * nch * 2;
  const size_t t_total = (size_t) gridDim.x * STAND_THREADS;
  const size_t tid_f = (size_t) blockIdx.x * STAND_THREADS + threadIdx.x;

  if (g.path == 2) {
    /* C: walk rows; consecutive threads take consecutive channels of a row */
    const size_t rows = frame_elems / nch;
    for (size_t c = tid_f; c < nch; c += t_total) {
      const double mean = st[c * 2], sd = st[c * 2 + 1];
      const double rsd = default_mode ? __drcp_rn (sd) : 0.0;
      for (size_t i = 0; i < rows; i++) {
        const double x = CCast<double, In>::run (ibase[i * nch + c]);
        const double y = default_mode ? stand_one<true> (x, mean, sd, rsd)
            : stand_one<false> (x, mean, sd, rsd);
        obase[i * nch + c] = CCast<Out, double>::run (y);
      }
    }
    return;
  }

  const size_t t_eff = (t_total / nch) * nch;
  if (tid_f >= t_eff)
    return;
  const bool aligned = ((uintptr_t) ibase % (V * sizeof (In))) == 0
      && ((uintptr_t) obase % (V * sizeof (Out))) == 0;
  const size_t n_vec = aligned ? frame_elems / V : 0;

  if (n_vec) {
    const uint32_t c0 = (uint32_t) ((tid_f * V) % nch);
    double mean[V], sd[V], rsd[V];
#pragma unroll
    for (int k = 0; k < V; k++) {
      const uint32_t c = (c0 + k) % nch;
      mean[k] = st[c * 2];
      sd[k] = st[c * 2 + 1];
      rsd[k] = default_mode ? __drcp_rn (sd[k]) : 0.0;
    }
    for (size_t v = tid_f; v < n_vec; v += t_eff * 2) {
      Vec<In, V> x0, x1;
      const size_t v1 = v + t_eff;
      load_vec (x0, ibase + v * V);
      if (v1 < n_vec)
        load_vec (x1, ibase + v1 * V);
      Vec<Out, V> y;
      if (default_mode) {
#pragma unroll
        for (int k = 0; k < V; k++)
          y.e[k] = CCast<Out, double>::run (stand_one<true> (CCast<double, In>::run (x0.e[k]),
                  mean[k], sd[k], rsd[k]));
      } else {
#pragma unroll
        for (int k = 0; k < V; k++)
          y.e[k] = CCast<Out, double>::run (stand_one<false> (CCast<double, In>::run (x0.e[k]),
                  mean[k], sd[k], rsd[k]));
      }
      store_vec (obase + v * V, y);
      if (v1 < n_vec) {
        if (default_mode) {
#pragma unroll
          for (int k = 0; k < V; k++)
            y.e[k] = CCast<Out, double>::run (stand_one<true> (CCast<double, In>::run (x1.e[k]),
                    mean[k], sd[k], rsd[k]));
        } else {
#pragma unroll
          for (int k = 0; k < V; k++)
            y.e[k] = CCast<Out, double>::run (stand_one<false> (CCast<double, In>::run (x1.e[k]),
                    mean[k], sd[k], rsd[k]));
        }
        store_vec (obase + v1 * V, y);
      }
    }
  }
  /* scalar remainder / unaligned tensors: fixed channel per thread */
  const size_t e0 = n_vec * V + tid_f;
  if (e0 < frame_elems) {
    const uint32_t c = (uint32_t) (e0 % nch);
    const double mean = st[c * 2], sd = st[c * 2 + 1];
    const double rsd = default_mode ? __drcp_rn (sd) : 0.0;
    for (size_t e = e0; e < frame_elems; e += t_eff) {
      const double x = CCast<double, In>::run (ibase[e]);
      const double y = default_mode ? stand_one<true> (x, mean, sd, rsd)
          : stand_one<false> (x, mean, sd, rsd);
      obase[e] = CCast<Out, double>::run (y);
    }
  }
}

/* ---------------------------------------------------------------------------------- */
/* Specialisations for 1..4 channels (gray, gray+alpha, RGB, RGBA -- and "not per-channel"):  */
/* with the channel count a compile-time constant, lane k of a vector belongs to the      */
/* RELATIVE channel k % NCH (actual channel (c0 + k % NCH) % NCH, c0 fixed per thread), so  */
/* a thread needs NCH accumulator pairs / NCH {mean, 1/std} pairs instead of one per lane:   */
/* ~40 registers instead of ~100, i.e. 2-3x the warps in flight.                            */
/* ---------------------------------------------------------------------------------- */
/* statistics of one tensor: block `bid` of `nblocks`; writes NCH {sum, sum of squares} pairs */
template <typename In, bool NEED_STD, int NCH>
__device__ __forceinline__ void
stand_stats_body (const In *__restrict__ base, size_t frame_elems, uint32_t bid, uint32_t nblocks,
    double *__restrict__ my_partial)
{
  constexpr int V = StandVec<In>::V;
  constexpr int U = 2;
  const size_t t_total = (size_t) nblocks * STAND_THREADS;
  const size_t t_eff = (t_total / NCH) * NCH;
  const size_t tid_f = (size_t) bid * STAND_THREADS + threadIdx.x;
  const bool active = tid_f < t_eff;
  const bool aligned = ((uintptr_t) base % (V * sizeof (In))) == 0;
  const size_t n_vec = aligned ? frame_elems / V : 0;
  const uint32_t c0 = (uint32_t) ((tid_f * V) % NCH);

  double a1[NCH], a2[NCH], kk[NCH];
#pragma unroll
  for (int r = 0; r < NCH; r++) {
    a1[r] = 0.0;
    a2[r] = 0.0;
    kk[r] = CCast<double, In>::run (base[(c0 + r) % NCH]);
  }
  if (active) {
    for (size_t v = tid_f; v < n_vec; v += t_eff * U) {
      Vec<In, V> x[U];
#pragma unroll
      for (int u = 0; u < U; u++)
        if (v + (size_t) u * t_eff < n_vec)
          load_vec (x[u], base + (v + (size_t) u * t_eff) * V);
#pragma unroll
      for (int u = 0; u < U; u++) {
        if (v + (size_t) u * t_eff < n_vec) {
#pragma unroll
          for (int k = 0; k < V; k++) {
            const double d = __dsub_rn (CCast<double, In>::run (x[u].e[k]), kk[k % NCH]);
            a1[k % NCH] += d;
            if (NEED_STD) a2[k % NCH] = fma (d, d, a2[k % NCH]);
          }
        }
      }
    }
  }
  /* actual-channel order */
  double s1[NCH], s2[NCH];
#pragma unroll
  for (int c = 0; c < NCH; c++) {
    s1[c] = 0.0;
    s2[c] = 0.0;
#pragma unroll
    for (int r = 0; r < NCH; r++)
      if ((c0 + r) % NCH == (uint32_t) c) {
        s1[c] = a1[r];
        s2[c] = a2[r];
      }
  }
  if (!active) {
#pragma unroll
    for (int c = 0; c < NCH; c++)
      s1[c] = s2[c] = 0.0;
  }
  /* scalar remainder (and everything when the tensor base is not vector aligned) */
  if (active) {
    const size_t e0 = n_vec * V + tid_f;
    if (e0 < frame_elems) {
      const uint32_t c = (uint32_t) (e0 % NCH);
      const double k0 = CCast<double, In>::run (base[c]);
      double r1 = 0.0, r2 = 0.0;
      for (size_t e = e0; e < frame_elems; e += t_eff) {
        const double d = __dsub_rn (CCast<double, In>::run (base[e]), k0);
        r1 += d;
        if (NEED_STD) r2 = fma (d, d, r2);
      }
#pragma unroll
      for (int cc = 0; cc < NCH; cc++)
        if (c == (uint32_t) cc) {
          s1[cc] += r1;
          s2[cc] += r2;
        }
    }
  }
  /* block reduction in a fixed order: warp shuffles, then 8 warp results */
  __shared__ double wsum[2][NCH][STAND_THREADS / 32];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int c = 0; c < NCH; c++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s1[c] += __shfl_xor_sync (0xffffffffu, s1[c], o);
      s2[c] += __shfl_xor_sync (0xffffffffu, s2[c], o);
    }
    if (lane == 0) {
      wsum[0][c][warp] = s1[c];
      wsum[1][c][warp] = s2[c];
    }
  }
  __syncthreads ();
  if (threadIdx.x < NCH) {
    double t1 = 0.0, t2 = 0.0;
#pragma unroll
    for (int w = 0; w < STAND_THREADS / 32; w++) {
      t1 += wsum[0][threadIdx.x][w];
      t2 += wsum[1][threadIdx.x][w];
    }
    my_partial[threadIdx.x * 2 + 0] = t1;
    my_partial[threadIdx.x * 2 + 1] = t2;
  }
}

template <typename In, bool NEED_STD, int NCH>
__global__ void __launch_bounds__ (STAND_THREADS)
stand_stats_kernel_n (const In *__restrict__ in, size_t frame_elems, StandGeom g,
    double *__restrict__ partial)
{
  const uint32_t f = blockIdx.y;
  stand_stats_body<In, NEED_STD, NCH> (in + (size_t) f * frame_elems, frame_elems, blockIdx.x,
      gridDim.x, partial + stand_partial_index (g, f, blockIdx.x, 0));
}

/* standardise one tensor: block `bid` of `nblocks`; st = its NCH {mean, std} pairs */
template <typename In, typename Out, int NCH>
__device__ __forceinline__ void
stand_apply_body (const In *__restrict__ ibase, Out *__restrict__ obase, size_t frame_elems,
    uint32_t bid, uint32_t nblocks, const double *__restrict__ st, int default_mode)
{
  constexpr int VI = StandVec<In>::V, VO = StandVec<Out>::V;
  constexpr int V = VI < VO ? VI : VO;
  constexpr int U = 2;
  const size_t t_total = (size_t) nblocks * STAND_THREADS;
  const size_t t_eff = (t_total / NCH) * NCH;
  const size_t tid_f = (size_t) bid * STAND_THREADS + threadIdx.x;
  if (tid_f >= t_eff)
    return;
  const bool aligned = ((uintptr_t) ibase % (V * sizeof (In))) == 0
      && ((uintptr_t) obase % (V * sizeof (Out))) == 0;
  const size_t n_vec = aligned ? frame_elems / V : 0;
  const uint32_t c0 = (uint32_t) ((tid_f * V) % NCH);
  double mean[NCH], rsd[NCH];
#pragma unroll
  for (int r = 0; r < NCH; r++) {
    const uint32_t c = (c0 + r) % NCH;
    mean[r] = st[c * 2];
    rsd[r] = default_mode ? __drcp_rn (st[c * 2 + 1]) : 0.0;
  }
  for (size_t v = tid_f; v < n_vec; v += t_eff * U) {
    Vec<In, V> x[U];
#pragma unroll
    for (int u = 0; u < U; u++)
      if (v + (size_t) u * t_eff < n_vec)
        load_vec (x[u], ibase + (v + (size_t) u * t_eff) * V);
#pragma unroll
    for (int u = 0; u < U; u++) {
      if (v + (size_t) u * t_eff < n_vec) {
        Vec<Out, V> y;
        if (default_mode) {
#pragma unroll
          for (int k = 0; k < V; k++)
            y.e[k] = CCast<Out, double>::run (stand_one<true> (CCast<double, In>::run (x[u].e[k]),
                    mean[k % NCH], 0.0, rsd[k % NCH]));
        } else {
#pragma unroll
          for (int k = 0; k < V; k++)
            y.e[k] = CCast<Out, double>::run (stand_one<false> (CCast<double, In>::run (x[u].e[k]),
                    mean[k % NCH], 0.0, 0.0));
        }
        store_vec (obase + (v + (size_t) u * t_eff) * V, y);
      }
    }
  }
  const size_t e0 = n_vec * V + tid_f;
  if (e0 < frame_elems) {
    const uint32_t c = (uint32_t) (e0 % NCH);
    const double m = st[c * 2];
    const double r_ = default_mode ? __drcp_rn (st[c * 2 + 1]) : 0.0;
    for (size_t e = e0; e < frame_elems; e += t_eff) {
      const double x = CCast<double, In>::run (ibase[e]);
      const double y = default_mode ? stand_one<true> (x, m, 0.0, r_) : stand_one<false> (x, m, 0.0, r_);
      obase[e] = CCast<Out, double>::run (y);
    }
  }
}

template <typename In, typename Out, int NCH>
__global__ void __launch_bounds__ (STAND_THREADS)
stand_apply_kernel_n (const In *__restrict__ in, Out *__restrict__ out, size_t frame_elems,
    StandGeom g, const double *__restrict__ stats, int default_mode)
{
  const uint32_t f = blockIdx.y;
  (void) g;
  stand_apply_body<In, Out, NCH> (in + (size_t) f * frame_elems, out + (size_t) f * frame_elems,
      frame_elems, blockIdx.x, gridDim.x, stats + (size_t) f * NCH * 2, default_mode);
}

/* geometry shared by the three launches (host) */
inline StandGeom
stand_geometry (uint32_t batch, uint32_t nch, size_t frame_elems, const LaunchCfg &cfg)
{
  StandGeom g;
  /* oversubscribed like the chain kernels (sweep on f32 1080p x64: 4 blocks/SM 4.06 TB/s, 16: 4.23,
   * 32: 4.39, 64: 4.47, 128: 4.48; one 400 MB tensor peaks at 32: 4.20) */
  const int bps = cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : (batch > 1 ? 64 : 32);
  uint64_t cap = (uint64_t) cfg.sm_count * bps;
  /* split the SMs between the tensors of a batch, but keep at least 8 blocks per tensor */
  uint64_t bx = cap / (batch ? batch : 1);
  if (bx < 8)
    bx = 8;
  if (bx > cap)
    bx = cap;
  /* no more blocks than there is work for */
  uint64_t need = (frame_elems + (uint64_t) STAND_THREADS * 32 - 1) / ((uint64_t) STAND_THREADS * 32);
  if (need < 1)
    need = 1;
  if (bx > need)
    bx = need;
  g.nch = nch;
  g.batch = batch;
  if (nch <= STAND_NCH_FAST) {
    g.path = 0;
  } else if ((uint64_t) nch <= bx * STAND_THREADS) {
    g.path = 1;
    /* bound the partial array: blocks * nch * 16 B <= ~4 MB per tensor */
    uint64_t lim = (4u << 20) / ((uint64_t) nch * 16);
    if (lim < 1)
      lim = 1;
    if (bx > lim)
      bx = lim;
    if ((uint64_t) nch > bx * STAND_THREADS)
      g.path = 2;
  } else {
    g.path = 2;
  }
  if (g.path == 2) {
    /* column walk: result written as the partial of block 0 */
    g.blocks_x = 1;
  } else {
    g.blocks_x = (uint32_t) bx;
  }
  return g;
}

typedef int (*stand_stats_fn) (const void *in, size_t frame_elems, const StandGeom &g,
    uint32_t grid_x, double *partial, double *stats, int need_std, cudaStream_t stream);
typedef int (*stand_apply_fn) (const void *in, void *out, size_t frame_elems, const StandGeom &g,
    uint32_t grid_x, const double *stats, int default_mode, cudaStream_t stream);

template <typename In> int
stand_stats_t (const void *in, size_t frame_elems, const StandGeom &g, uint32_t grid_x,
    double *partial, double *stats, int need_std, cudaStream_t stream)
{
  dim3 grid (grid_x, g.batch);
#define NNSB_STATS_N(N) \
  if (need_std) stand_stats_kernel_n<In, true, N><<<grid, STAND_THREADS, 0, stream>>> ( \
      (const In *) in, frame_elems, g, partial); \
  else stand_stats_kernel_n<In, false, N><<<grid, STAND_THREADS, 0, stream>>> ( \
      (const In *) in, frame_elems, g, partial);
  if (g.path == 0 && g.nch == 1) { NNSB_STATS_N (1) }
  else if (g.path == 0 && g.nch == 2) { NNSB_STATS_N (2) }
  else if (g.path == 0 && g.nch == 3) { NNSB_STATS_N (3) }
  else if (g.path == 0 && g.nch == 4) { NNSB_STATS_N (4) }
  else if (need_std)
    stand_stats_kernel<In, true><<<grid, STAND_THREADS, 0, stream>>> ((const In *) in,
        frame_elems, g, partial);
  else
    stand_stats_kernel<In, false><<<grid, STAND_THREADS, 0, stream>>> ((const In *) in,
        frame_elems, g, partial);
#undef NNSB_STATS_N
  int err = (int) cudaGetLastError ();
  if (err)
    return err;
  const size_t warps = (size_t) g.batch * g.nch;
  if (g.blocks_x > 512 && warps <= 65535) {
    stand_finalize_block_kernel<In><<<(unsigned int) warps, STAND_THREADS, 0, stream>>> (
        (const In *) in, frame_elems, g, partial, stats, need_std);
    return (int) cudaGetLastError ();
  }
  const unsigned int fb = (unsigned int) ((warps * 32 + STAND_THREADS - 1) / STAND_THREADS);
  stand_finalize_kernel<In><<<fb, STAND_THREADS, 0, stream>>> ((const In *) in, frame_elems, g,
      partial, stats, need_std);
  return (int) cudaGetLastError ();
}

template <typename In, typename Out> int
stand_apply_t (const void *in, void *out, size_t frame_elems, const StandGeom &g, uint32_t grid_x,
    const double *stats, int default_mode, cudaStream_t stream)
{
  dim3 grid (grid_x, g.batch);
#define NNSB_APPLY_N(N) \
  stand_apply_kernel_n<In, Out, N><<<grid, STAND_THREADS, 0, stream>>> ((const In *) in, \
      (Out *) out, frame_elems, g, stats, default_mode);
  if (g.path == 0 && g.nch == 1) { NNSB_APPLY_N (1) }
  else if (g.path == 0 && g.nch == 2) { NNSB_APPLY_N (2) }
  else if (g.path == 0 && g.nch == 3) { NNSB_APPLY_N (3) }
  else if (g.path == 0 && g.nch == 4) { NNSB_APPLY_N (4) }
  else
    stand_apply_kernel<In, Out><<<grid, STAND_THREADS, 0, stream>>> ((const In *) in, (Out *) out,
        frame_elems, g, stats, default_mode);
#undef NNSB_APPLY_N
  return (int) cudaGetLastError ();
}

template <typename In> struct StandRow
{
  static stand_stats_fn stats () { return &stand_stats_t<In>; }
  static stand_apply_fn apply (int out_type)
  {
    switch (out_type) {
#define NNSB_CASE(T) case TypeId<T>::id: return &stand_apply_t<In, T>;
      NNSB_CASE (int32_t) NNSB_CASE (uint32_t) NNSB_CASE (int16_t) NNSB_CASE (uint16_t)
      NNSB_CASE (int8_t) NNSB_CASE (uint8_t) NNSB_CASE (double) NNSB_CASE (float)
      NNSB_CASE (int64_t) NNSB_CASE (uint64_t) NNSB_CASE (__half)
#undef NNSB_CASE
      default: return nullptr;
    }
  }
};

}  // namespace nnsb
#endif
