/*
 * k_perm_slab.cu -- K3/K4 slab path: permutations whose rows are NOT multiples of 16 bytes
 * (u16 225x113x3 frames, 227x227x3 floats, frames of 152 550 bytes back to back ...), where
 * the table-driven vector kernel (k_perm_vec.cu) falls back to 8- / 4-byte vectors or to the
 * scalar kernel because a 16-byte access would straddle a row.
 *
 * Same job as the other permute kernels: gst_tensor_transform_transpose
 * (gsttensor_transform.c:1417-1500) / _dimchg (:1120-1191).  Idea: bytes do not care about
 * rows.
 *   load    a SLAB = k consecutive slices of the dimension that is outermost on the input side
 *           (that fits: <= ~100 KB) is one contiguous byte range of the input, whatever its
 *           alignment: the 16-byte aligned superset of it is copied to shared memory with
 *           cp.async.cg 16 (the element at slab byte b sits at smem[mis + b], mis = start & 15).
 *   store   the slab's elements form contiguous RUNS in the output (the dims below the slab
 *           dimension in output order, full, times k).  A thread owns 16-byte aligned output
 *           chunks of a run: it gathers the chunk's 16 / es elements from shared memory
 *           (coordinates advance incrementally, one divide per chunk) and writes one
 *           st.global.v4; the at most two chunks per run that are cut by the run's ends are
 *           written element by element.
 * Two CTAs per SM alternate load and store phases.  Algorithmic bytes: 2 * es per element.
 * Used when the vector plan is not 16 bytes wide and the runs are >= 128 bytes long.
 */
#include <cstdlib>
#include <cstring>

#include "nnsb_device.cuh"
#include "nnsb_kernels.h"
#include "k_perm_common.cuh"

namespace nnsb {

namespace {

constexpr int SLAB_THREADS = 256;
constexpr uint32_t SLAB_MAX_BYTES = 72 * 1024;        /* three CTAs per SM overlap their load and store phases */
constexpr int SLAB_DIMS = 4;

struct SlabDev
{
  int es;
  /* dims inside a run, output order; the last one may be the slab dimension (extent k) */
  int n_run;
  uint32_t run_size[SLAB_DIMS];         /* extents (k for the slab dimension) */
  FastDiv run_div[SLAB_DIMS];           /* the same as divisors: only full-size dims are ever divided by */
  uint32_t run_in[SLAB_DIMS];           /* input strides, elements, slab relative */
  int run_slab;                         /* index of the slab dimension among the run dims, -1 if none */
  /* dims that enumerate the runs of a slab (the slab dimension is one of them unless it is in the run) */
  int n_idx;
  FastDiv idx_div[SLAB_DIMS];
  uint32_t idx_in[SLAB_DIMS];           /* elements, slab relative */
  uint64_t idx_out[SLAB_DIMS];          /* elements */
  int idx_slab;                         /* index of the slab dimension among them, -1 if none */
  /* dims outside (fixed per slab) */
  int n_outer;
  FastDiv outer_div[SLAB_DIMS];
  uint64_t outer_in[SLAB_DIMS], outer_out[SLAB_DIMS];
  /* the slab dimension */
  uint32_t s_size, k;
  FastDiv slabs_per_outer;
  uint64_t s_in, s_out;                 /* elements */
  uint32_t n_slabs;
  uint64_t total_bytes;
  /* divisors that depend on the slab height, for full slabs [0] and for the last, shorter one [1] */
  FastDiv cpr_div[2], keff_div[2];
  /* lane-per-element store (store_lanes): a run is cut into segments of `seg` elements (a multiple
   * of lane_w), a group of lane_w lanes (a warp, or a half / quarter of one for short runs) owns a
   * segment, lane l its elements l, l + lane_w, ...  The step of lane_w elements in the mixed
   * radix of the run dims, and what it adds to the source offset: */
  int lanes;                            /* use store_lanes */
  uint32_t lane_w;                      /* 32, 16 or 8 */
  uint32_t seg;
  FastDiv nseg_div[2];                  /* segments per run: full slab, last slab */
  uint32_t step_c0, step_c1;            /* lane_w = (.. * x1 + step_c1) * x0 + step_c0 */
  uint32_t step_s;                      /* source elements a step adds when nothing wraps */
  uint32_t wrap0_s, wrap1_s;            /* added (mod 2^32) when c0 / c1 wraps */
};

__device__ __forceinline__ uint32_t
lds_elem (const unsigned char *p, int es, uint32_t *hi)
{
  *hi = 0;
  if (es == 1)
    return *p;
  if (es == 2)
    return *reinterpret_cast<const uint16_t *> (p);
  if (es == 4)
    return *reinterpret_cast<const uint32_t *> (p);
  const uint2 v = *reinterpret_cast<const uint2 *> (p);
  *hi = v.y;
  return v.x;
}

template <int ES> __device__ __forceinline__ void
copy_elem (unsigned char *dst, const unsigned char *src)
{
  if constexpr (ES == 1) {
    const unsigned char v = *src;
    asm volatile ("st.global.L1::no_allocate.u8 [%0], %1;" :: "l" (dst), "r" ((uint32_t) v) : "memory");
  } else if constexpr (ES == 2) {
    const uint16_t v = *reinterpret_cast<const uint16_t *> (src);
    asm volatile ("st.global.L1::no_allocate.u16 [%0], %1;" :: "l" (dst), "h" (v) : "memory");
  } else if constexpr (ES == 4) {
    const uint32_t v = *reinterpret_cast<const uint32_t *> (src);
    asm volatile ("st.global.L1::no_allocate.u32 [%0], %1;" :: "l" (dst), "r" (v) : "memory");
  } else if constexpr (ES == 8) {
    const uint2 v = *reinterpret_cast<const uint2 *> (src);
    asm volatile ("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};" :: "l" (dst), "r" (v.x), "r" (v.y) : "memory");
  } else {
    /* 3, 6, 12: a pixel of three 1-, 2- or 4-byte channels that stays together (see the fold in
     * launch_permute_slab): three accesses of the channel's size, which is all its alignment allows */
    constexpr int C = ES / 3;
    copy_elem<C> (dst, src);
    copy_elem<C> (dst + C, src + C);
    copy_elem<C> (dst + 2 * C, src + 2 * C);
  }
}

/* Store phase, one element per lane: consecutive lanes write consecutive elements of a run (a
 * warp store is 32 * ES contiguous bytes, whatever the alignment of the run), and read their
 * sources from shared memory at the stride the permutation gives them.  Nothing here depends on
 * alignment, and a lane spends a handful of instructions per element: its coordinates inside the
 * run advance by 32 with at most one carry per dim.  NRUN = dims inside a run. */
template <int ES, int NRUN> __device__ __forceinline__ void
store_lanes (const SlabDev &p, const unsigned char *src0, unsigned char *out, uint64_t out_base,
    uint32_t k_eff, int short_slab, uint32_t L, uint32_t n_runs)
{
  /* runs shorter than a warp: 2 or 4 of them side by side (lane_w = 16 or 8 lanes each) */
  const uint32_t W = p.lane_w, G = 32u / W;
  const uint32_t lane = threadIdx.x & (W - 1);
  const uint32_t group = (threadIdx.x >> 5) * G + ((threadIdx.x & 31) / W);
  const FastDiv nseg_div = p.nseg_div[short_slab], keff_div = p.keff_div[short_slab];
  const uint32_t x0 = p.run_slab == 0 ? k_eff : p.run_size[0];
  const uint32_t x1 = p.run_slab == 1 ? k_eff : p.run_size[1];
  const uint32_t n0 = p.n_idx > 0 ? (p.idx_slab == 0 ? k_eff : p.idx_div[0].d) : 1;
  const uint32_t n_outer_runs = p.n_idx > 0 ? n_runs / n0 : 1;
  const uint32_t items = n0 * nseg_div.d;
  const uint32_t stride = (SLAB_THREADS / 32) * G;
  /* where lane l starts inside a run (element l): the same for every run of the slab */
  uint32_t c0i = lane, c1i = 0, c2i = 0;
  if (NRUN > 1) {
    divmod (lane, p.run_div[0], c1i, c0i);
    if (NRUN > 2) {
      const uint32_t t = c1i;
      divmod (t, p.run_div[1], c2i, c1i);
    }
  }
  uint32_t s_lane = c0i * p.run_in[0];
  if (NRUN > 1)
    s_lane += c1i * p.run_in[1];
  if (NRUN > 2)
    s_lane += c2i * p.run_in[2];
  const uint32_t in_step = p.n_idx > 0 ? stride * p.idx_in[0] : 0;
  const uint64_t out_step = p.n_idx > 0 ? (uint64_t) stride * p.idx_out[0] * ES : 0;
  for (uint32_t o = 0; o < n_outer_runs; o++) {
    /* the run-index dims above the first: the same for every warp */
    uint64_t out_o = out_base;
    uint32_t in_o = 0, r = o;
#pragma unroll
    for (int d = 1; d < SLAB_DIMS; d++)
      if (d < p.n_idx) {
        uint32_t q, c;
        if (d == p.idx_slab)
          divmod (r, keff_div, q, c);
        else
          divmod (r, p.idx_div[d], q, c);
        r = q;
        out_o += (uint64_t) c * p.idx_out[d];
        in_o += c * p.idx_in[d];
      }
    if (nseg_div.d == 1) {
      /* one segment per run (runs of at most `seg` elements): from run to run only the bases
       * move, by constants */
      uint32_t sbase = in_o + s_lane;
      unsigned char *dbase = out + (out_o + lane) * ES;
      if (p.n_idx > 0) {
        sbase += group * p.idx_in[0];
        dbase += (uint64_t) group * p.idx_out[0] * ES;
      }
      for (uint32_t m = group; m < n0; m += stride, sbase += in_step, dbase += out_step) {
        uint32_t sidx = sbase, c0 = c0i, c1 = c1i;
        unsigned char *dst = dbase;
        for (uint32_t j = lane; j < L; j += W) {
          copy_elem<ES> (dst, src0 + (size_t) sidx * ES);
          dst += W * ES;
          sidx += p.step_s;
          if (NRUN > 1) {
            c0 += p.step_c0;
            if (c0 >= x0) {
              c0 -= x0;
              sidx += p.wrap0_s;
              if (NRUN > 2)
                c1 += 1;
            }
            if (NRUN > 2) {
              c1 += p.step_c1;
              if (c1 >= x1) {
                c1 -= x1;
                sidx += p.wrap1_s;
              }
            }
          }
        }
      }
      continue;
    }
    for (uint32_t m = group; m < items; m += stride) {
      uint32_t i0, sg;
      divmod (m, nseg_div, i0, sg);
      uint32_t j = sg * p.seg + lane;
      uint32_t j_end = (sg + 1) * p.seg;
      if (j_end > L)
        j_end = L;
      uint32_t c0 = j, c1 = 0, c2 = 0;
      if (NRUN > 1) {
        divmod (j, p.run_div[0], c1, c0);
        if (NRUN > 2) {
          const uint32_t t = c1;
          divmod (t, p.run_div[1], c2, c1);
        }
      }
      uint32_t sidx = in_o + c0 * p.run_in[0];
      uint64_t out_r = out_o + j;
      if (p.n_idx > 0) {
        sidx += i0 * p.idx_in[0];
        out_r += (uint64_t) i0 * p.idx_out[0];
      }
      if (NRUN > 1)
        sidx += c1 * p.run_in[1];
      if (NRUN > 2)
        sidx += c2 * p.run_in[2];
      unsigned char *dst = out + out_r * ES;
      for (; j < j_end; j += W) {
        copy_elem<ES> (dst, src0 + (size_t) sidx * ES);
        dst += W * ES;
        sidx += p.step_s;
        if (NRUN > 1) {
          c0 += p.step_c0;
          if (c0 >= x0) {
            c0 -= x0;
            sidx += p.wrap0_s;
            if (NRUN > 2)
              c1 += 1;
          }
          if (NRUN > 2) {
            c1 += p.step_c1;
            if (c1 >= x1) {
              c1 -= x1;
              sidx += p.wrap1_s;
            }
          }
        }
      }
    }
  }
}

template <int ES, bool LANES>
__global__ void __launch_bounds__ (SLAB_THREADS, LANES ? 6 : 5)
permute_slab_kernel (const unsigned char *__restrict__ in, unsigned char *__restrict__ out,
    const __grid_constant__ SlabDev p)
{
  extern __shared__ __align__ (16) unsigned char buf[];
  const uintptr_t in_lo = (uintptr_t) in, in_hi = in_lo + p.total_bytes;

  for (uint32_t slab = blockIdx.x; slab < p.n_slabs; slab += gridDim.x) {
    /* which slab: (outer coordinates, position along the slab dimension) */
    uint32_t oi, si;
    divmod (slab, p.slabs_per_outer, oi, si);
    uint64_t in_base = 0, out_base = 0;         /* elements */
#pragma unroll
    for (int d = 0; d < SLAB_DIMS; d++)
      if (d < p.n_outer) {
        uint32_t q, c;
        divmod (oi, p.outer_div[d], q, c);
        oi = q;
        in_base += (uint64_t) c * p.outer_in[d];
        out_base += (uint64_t) c * p.outer_out[d];
      }
    const uint32_t o0 = si * p.k;
    const uint32_t k_eff = (p.s_size - o0 < p.k) ? p.s_size - o0 : p.k;
    in_base += (uint64_t) o0 * p.s_in;
    out_base += (uint64_t) o0 * p.s_out;
    const uint32_t slab_bytes = (uint32_t) (k_eff * p.s_in) * ES;

    /* ---- load: the aligned superset of [abs0, abs0 + slab_bytes) ---- */
    const uintptr_t abs0 = in_lo + in_base * ES;
    const uintptr_t ga = abs0 & ~(uintptr_t) 15;
    const uint32_t mis = (uint32_t) (abs0 - ga);
    const uint32_t n_chunks = (mis + slab_bytes + 15) / 16;
    for (uint32_t c = threadIdx.x; c < n_chunks; c += SLAB_THREADS) {
      const uintptr_t g = ga + (uintptr_t) c * 16;
      if (g >= in_lo && g + 16 <= in_hi) {
        asm volatile ("cp.async.cg.shared.global [%0], [%1], 16;"
            :: "r" ((uint32_t) __cvta_generic_to_shared (buf + c * 16)), "l" (g) : "memory");
      } else {
        /* first / last bytes of the whole tensor: never touch what is not ours */
        for (int b = 0; b < 16; b++)
          if (g + b >= in_lo && g + b < in_hi)
            buf[c * 16 + b] = *reinterpret_cast<const unsigned char *> (g + b);
      }
    }
    asm volatile ("cp.async.commit_group;" ::: "memory");
    asm volatile ("cp.async.wait_group 0;" ::: "memory");
    __syncthreads ();

    /* ---- store: runs of L elements, 16-byte chunks ---- */
    uint32_t L = 1;
#pragma unroll
    for (int d = 0; d < SLAB_DIMS; d++)
      if (d < p.n_run)
        L *= (d == p.run_slab) ? k_eff : p.run_size[d];
    uint32_t n_runs = 1;
#pragma unroll
    for (int d = 0; d < SLAB_DIMS; d++)
      if (d < p.n_idx)
        n_runs *= (d == p.idx_slab) ? k_eff : p.idx_div[d].d;
    const uint32_t run_bytes = L * ES;
    const int short_slab = k_eff != p.k;
    if constexpr (LANES) {
      if (p.n_run == 1)
        store_lanes<ES, 1> (p, buf + mis, out, out_base, k_eff, short_slab, L, n_runs);
      else if (p.n_run == 2)
        store_lanes<ES, 2> (p, buf + mis, out, out_base, k_eff, short_slab, L, n_runs);
      else
        store_lanes<ES, 3> (p, buf + mis, out, out_base, k_eff, short_slab, L, n_runs);
      __syncthreads ();         /* the buffer is about to be overwritten */
      continue;
    } else {
    const FastDiv cpr_div = p.cpr_div[short_slab], keff_div = p.keff_div[short_slab];
    const uint32_t cpr = cpr_div.d;                     /* chunks a run can touch, any alignment */
    const uint32_t items = n_runs * cpr;
    const unsigned char *src0 = buf + mis;
    for (uint32_t w = threadIdx.x; w < items; w += SLAB_THREADS) {
      uint32_t r, ci;
      divmod (w, cpr_div, r, ci);
      uint64_t out_off = out_base;
      uint32_t in_off = 0;
#pragma unroll
      for (int d = 0; d < SLAB_DIMS; d++)
        if (d < p.n_idx) {
          uint32_t q, c;
          if (d == p.idx_slab)
            divmod (r, keff_div, q, c);
          else
            divmod (r, p.idx_div[d], q, c);
          r = q;
          out_off += (uint64_t) c * p.idx_out[d];
          in_off += c * p.idx_in[d];
        }
      const uintptr_t A = (uintptr_t) out + out_off * ES;       /* first byte of the run */
      const uintptr_t g = (A & ~(uintptr_t) 15) + (uintptr_t) ci * 16;
      if (g >= A + run_bytes)
        continue;
      const uint32_t e_lo = g > A ? (uint32_t) (g - A) / ES : 0;
      uint32_t e_hi = (uint32_t) (g + 16 - A) / ES;
      if (e_hi > L)
        e_hi = L;
      /* coordinates of element e_lo inside the run (<= 3 dims), then incremental */
      uint32_t c0 = e_lo, c1 = 0, c2 = 0;
      const uint32_t x0 = p.run_slab == 0 ? k_eff : p.run_size[0];
      const uint32_t x1 = p.n_run > 1 ? (p.run_slab == 1 ? k_eff : p.run_size[1]) : 1;
      if (p.n_run > 1) {
        /* x0 (and x1) are full-size dims here: a run continues past a dim only if it is complete */
        divmod (e_lo, p.run_div[0], c1, c0);
        if (p.n_run > 2) {
          const uint32_t t = c1;
          divmod (t, p.run_div[1], c2, c1);
        }
      }
      uint32_t s = in_off + c0 * p.run_in[0] + c1 * p.run_in[1] + c2 * p.run_in[2];
      const bool full = g >= A && g + 16 <= A + run_bytes;
      /* advance (c0, c1, s) to the next element of the run */
#define NNSB_SLAB_NEXT() \
      do { \
        s += p.run_in[0]; \
        if (++c0 == x0) { \
          c0 = 0; \
          s += p.run_in[1] - x0 * p.run_in[0]; \
          if (++c1 == x1) { \
            c1 = 0; \
            s += p.run_in[2] - x1 * p.run_in[1]; \
          } \
        } \
      } while (0)
      if (full) {
        /* exactly 16 / ES elements: unrolled, the four words stay in registers */
        constexpr int NE = 16 / ES;
        uint32_t wv[4] = { 0, 0, 0, 0 };
#pragma unroll
        for (int i = 0; i < NE; i++) {
          uint32_t hi;
          const uint32_t v = lds_elem (src0 + (size_t) s * ES, ES, &hi);
          if (ES == 1)
            wv[i >> 2] |= v << (8 * (i & 3));
          else if (ES == 2)
            wv[i >> 1] |= v << (16 * (i & 1));
          else if (ES == 4)
            wv[i] = v;
          else {
            wv[2 * i] = v;
            wv[2 * i + 1] = hi;
          }
          NNSB_SLAB_NEXT ();
        }
        asm volatile ("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
            :: "l" (g), "r" (wv[0]), "r" (wv[1]), "r" (wv[2]), "r" (wv[3]) : "memory");
      } else {
        /* a chunk cut by one end of the run: element by element */
        for (uint32_t e = e_lo; e < e_hi; e++) {
          uint32_t hi;
          const uint32_t v = lds_elem (src0 + (size_t) s * ES, ES, &hi);
          unsigned char *d = reinterpret_cast<unsigned char *> (A) + (size_t) e * ES;
          if (ES == 1)
            *d = (unsigned char) v;
          else if (ES == 2)
            *reinterpret_cast<uint16_t *> (d) = (uint16_t) v;
          else if (ES == 4)
            *reinterpret_cast<uint32_t *> (d) = v;
          else
            *reinterpret_cast<uint2 *> (d) = make_uint2 (v, hi);
          NNSB_SLAB_NEXT ();
        }
      }
#undef NNSB_SLAB_NEXT
    }
    __syncthreads ();           /* the buffer is about to be overwritten */
    }                           /* !LANES */
  }
}

template <int ES, bool LANES> int
launch_t (const SlabDev &p, const void *in, void *out, size_t smem, const LaunchCfg &cfg,
    cudaStream_t stream)
{
  static int configured[64];
  int dev = 0;
  cudaGetDevice (&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute (permute_slab_kernel<ES, LANES>,
        cudaFuncAttributeMaxDynamicSharedMemorySize, (int) (SLAB_MAX_BYTES + 64));
    if (e != cudaSuccess)
      return (int) e;
    configured[dev] = 1;
  }
  uint64_t per_sm = (200 * 1024) / (smem + 1024);      /* what the shared memory allows */
  const uint64_t by_regs = LANES ? 6 : 5;
  if (per_sm > by_regs)
    per_sm = by_regs;
  if (per_sm < 2)
    per_sm = 2;
  if (const char *e = getenv ("NNSB_SLAB_CTAS"))
    per_sm = (uint64_t) atoi (e) > 0 ? (uint64_t) atoi (e) : per_sm;
  uint64_t grid = (uint64_t) cfg.sm_count * per_sm;
  if (grid > p.n_slabs)
    grid = p.n_slabs;
  permute_slab_kernel<ES, LANES><<<(unsigned int) grid, SLAB_THREADS, smem, stream>>> (
      (const unsigned char *) in, (unsigned char *) out, p);
  return (int) cudaGetLastError ();
}

}  // namespace

/* -1: this permutation does not suit the slab kernel; else the launch result */
int
launch_permute_slab (const CDim *dims, int n, int es, const void *in, void *out, uint64_t total_elems,
    const LaunchCfg &cfg, cudaStream_t stream, bool dry_run)
{
  if (n < 2 || n > 2 * SLAB_DIMS || (es != 1 && es != 2 && es != 4 && es != 8))
    return -1;
  if (((uintptr_t) in | (uintptr_t) out) % (uintptr_t) es)
    return -1;
  /* how the store phase works: 16-byte chunks per thread (bytes: fewest stores) or one element
   * per lane (everything else: fewest instructions, no alignment cases) */
  int lanes = es >= 2;
  bool chunks_forced = false;
  if (const char *e = getenv ("NNSB_SLAB_STORE")) {
    chunks_forced = strcmp (e, "chunks") == 0;
    lanes = strcmp (e, "lanes") == 0 ? 1 : chunks_forced ? 0 : lanes;
  }
  /* Three channels that are innermost on both sides (a pixel that moves as a whole: transposing
   * an RGB image) cannot be folded into a power-of-two element by the canonicaliser; the lane
   * store does not care, so here the pixel becomes the element (3, 6 or 12 bytes). */
  CDim folded[2 * SLAB_DIMS];
  if (n >= 3 && es <= 4 && dims[0].size == 3 && dims[0].in == 1 && dims[0].out == 1
      && !chunks_forced && !getenv ("NNSB_SLAB_NO_FOLD")) {
    bool ok = total_elems % 3 == 0;
    for (int i = 1; i < n && ok; i++)
      ok = dims[i].in % 3 == 0 && dims[i].out % 3 == 0;
    if (ok) {
      for (int i = 1; i < n; i++) {
        folded[i - 1] = dims[i];
        folded[i - 1].in /= 3;
        folded[i - 1].out /= 3;
      }
      dims = folded;
      n--;
      es *= 3;
      total_elems /= 3;
      lanes = 1;
    }
  }
  /* dims by input stride; the input must be dense */
  int by_in[2 * SLAB_DIMS];
  for (int i = 0; i < n; i++)
    by_in[i] = i;
  for (int i = 1; i < n; i++)
    for (int j = i; j > 0 && dims[by_in[j]].in < dims[by_in[j - 1]].in; j--) {
      const int t = by_in[j];
      by_in[j] = by_in[j - 1];
      by_in[j - 1] = t;
    }
  uint64_t expect = 1;
  for (int i = 0; i < n; i++) {
    if (dims[by_in[i]].in != expect || dims[by_in[i]].size >= (1ull << 31))
      return -1;
    expect *= dims[by_in[i]].size;
  }
  /* the slab dimension: the outermost (on the input side) whose single slice fits */
  int si = -1;
  for (int i = n - 1; i >= 0; i--)
    if (dims[by_in[i]].in * (uint64_t) es <= SLAB_MAX_BYTES) {
      si = i;
      break;
    }
  if (si < 1)
    return -1;                  /* nothing fits, or the slab dimension would be the innermost */
  const int S = by_in[si];
  /* slab height: as tall as fits, but short enough that the grid gets several slabs per CTA slot
   * (a batch of small frames is only a few hundred 72 KB slabs, and with two resident CTAs per
   * SM those quantise into 2-3 rounds of which the last is half empty) */
  uint64_t max_bytes = 24 * 1024;       /* sweep: profiles/r2_slab_sweep.md */
  if (const char *e = getenv ("NNSB_SLAB_KB"))
    if (atoi (e) > 0 && (uint64_t) atoi (e) * 1024 <= SLAB_MAX_BYTES)
      max_bytes = (uint64_t) atoi (e) * 1024;
  if (dims[S].in * (uint64_t) es > max_bytes)
    max_bytes = dims[S].in * (uint64_t) es;
  uint64_t k = max_bytes / (dims[S].in * es);
  if (k > dims[S].size)
    k = dims[S].size;
  bool balance = true;          /* equal slabs, unless a height was picked for the lanes */
  if (lanes && k < dims[S].size) {
    /* When the slab dimension is part of the output runs, a run is per_k * k elements and lanes
     * idle where that is not a multiple of the group width: among the heights down to half the
     * tallest, take the one that fills the lanes best. */
    bool fixed[2 * SLAB_DIMS];
    for (int i = 0; i < n; i++)
      fixed[i] = false;
    for (int i = si + 1; i < n; i++)
      fixed[by_in[i]] = true;
    uint64_t per_k = 0, prod = 1, next_out = 1;
    int cnt = 0;
    for (int d = 0; d < n && cnt < 3; d++, cnt++) {
      if (fixed[d] || dims[d].out != next_out)
        break;
      if (d == S) {
        per_k = prod;
        break;
      }
      prod *= dims[d].size;
      next_out = dims[d].out * dims[d].size;
    }
    if (per_k && per_k * k <= 512) {
      /* (long runs waste at most one group in many; this runs at every launch, so only the
       * handful of heights whose runs are whole groups of 8 lanes are looked at) */
      auto slots = [] (uint64_t L, uint64_t w) { return (L + w - 1) / w * w; };
      double best = -1.0;
      uint64_t best_k = k;
      for (uint64_t kc = k; kc >= 1 && 2 * kc >= k; kc--) {
        if (kc != k && (per_k * kc) % 8)
          continue;
        const uint64_t ns = (dims[S].size + kc - 1) / kc, k_last = dims[S].size - (ns - 1) * kc;
        const uint64_t Lf = per_k * kc, w = Lf <= 8 ? 8 : Lf <= 16 ? 16 : 32;
        if (Lf < 8)
          break;
        const double eff = (double) (per_k * dims[S].size) / (double) ((ns - 1) * slots (Lf, w) + slots (per_k * k_last, w));
        if (eff > best + 0.02) {
          best = eff;
          best_k = kc;
        }
      }
      if (best_k != k) {
        k = best_k;
        balance = false;
      }
    }
  }
  const uint64_t nsl = (dims[S].size + k - 1) / k;
  if (balance)
    k = (dims[S].size + nsl - 1) / nsl;
  if (k * dims[S].in * es >= (1ull << 31))
    return -1;

  SlabDev p;
  memset (&p, 0, sizeof (p));
  p.es = es;
  p.idx_slab = -1;
  p.run_slab = -1;
  bool outer[2 * SLAB_DIMS];
  for (int i = 0; i < n; i++)
    outer[i] = false;
  for (int i = si + 1; i < n; i++) {
    if (p.n_outer >= SLAB_DIMS)
      return -1;
    const int d = by_in[i];
    outer[d] = true;
    p.outer_div[p.n_outer] = make_fastdiv ((uint32_t) dims[d].size);
    p.outer_in[p.n_outer] = dims[d].in;
    p.outer_out[p.n_outer] = dims[d].out;
    p.n_outer++;
  }
  /* local dims in output order (dims[] is in output order already): the run is the prefix that
   * is contiguous in the output */
  bool in_run = true;
  uint64_t next_out = 1, run_elems = 1;
  for (int d = 0; d < n; d++) {
    if (outer[d]) {
      in_run = false;           /* a fixed dim below: whatever follows starts new runs */
      continue;
    }
    if (in_run && dims[d].out == next_out && p.n_run < 3) {
      p.run_size[p.n_run] = (uint32_t) (d == S ? k : dims[d].size);
      p.run_div[p.n_run] = make_fastdiv (p.run_size[p.n_run]);
      p.run_in[p.n_run] = (uint32_t) dims[d].in;
      p.n_run++;
      run_elems *= (d == S ? k : dims[d].size);
      next_out = dims[d].out * dims[d].size;
      if (d == S) {
        p.run_slab = p.n_run - 1;
        if (k < dims[S].size)
          in_run = false;
      }
      continue;
    }
    in_run = false;
    if (p.n_idx >= SLAB_DIMS)
      return -1;
    if (d == S)
      p.idx_slab = p.n_idx;
    p.idx_div[p.n_idx] = make_fastdiv ((uint32_t) (d == S ? k : dims[d].size));
    p.idx_in[p.n_idx] = (uint32_t) dims[d].in;
    p.idx_out[p.n_idx] = dims[d].out;
    p.n_idx++;
  }
  p.lanes = lanes;
  if (p.n_run == 0 || run_elems >= (1ull << 28))
    return -1;
  if (lanes && !getenv ("NNSB_SLAB_ANY_STRIDE")) {
    /* neighbouring lanes read shared memory run_in[0] elements apart: when that stride puts four
     * or more of them into one bank (1080p RGB rows: 5760 bytes = 45 x 128) the gather serialises
     * and the tiled kernels, which pad their tiles, do better */
    const uint64_t stride_b = (uint64_t) p.run_in[0] * es;
    int per_bank[32] = { 0 };
    uint64_t words[32];
    int worst = 0;
    for (int l = 0; l < 32; l++) {
      words[l] = l * stride_b / 4;
      bool seen = false;
      for (int m = 0; m < l; m++)
        seen = seen || words[m] == words[l];
      if (!seen && ++per_bank[words[l] % 32] > worst)
        worst = per_bank[words[l] % 32];
    }
    if (worst >= 4)
      return -1;
  }
  if (p.lanes ? run_elems < 8 : run_elems * es < 128)
    return -1;                  /* short runs: the tiled kernels do better */
  /* slab-relative input strides: dims inside the slab have in < in[S] * k already */
  p.s_size = (uint32_t) dims[S].size;
  p.k = (uint32_t) k;
  p.s_in = dims[S].in;
  p.s_out = dims[S].out;
  p.slabs_per_outer = make_fastdiv ((uint32_t) nsl);
  uint64_t n_slabs = nsl;
  for (int i = 0; i < p.n_outer; i++)
    n_slabs *= p.outer_div[i].d;
  if (n_slabs >= (1ull << 31))
    return -1;
  p.n_slabs = (uint32_t) n_slabs;
  p.total_bytes = total_elems * (uint64_t) es;
  {
    const uint64_t k_last = dims[S].size - (nsl - 1) * k;       /* height of the last slab */
    const uint64_t ks[2] = { k, k_last };
    /* segments: long runs are cut so that the 8 warps of a CTA all find work */
    p.lane_w = (p.lanes && run_elems <= 8) ? 8 : (p.lanes && run_elems <= 16) ? 16 : 32;
    p.seg = (uint32_t) ((run_elems + p.lane_w - 1) / p.lane_w * p.lane_w);
    if (p.seg > 512)
      p.seg = 512;
    for (int v = 0; v < 2; v++) {
      uint64_t run_v = run_elems;
      if (p.run_slab >= 0)
        run_v = run_elems / k * ks[v];
      p.cpr_div[v] = make_fastdiv ((uint32_t) ((run_v * es + 15 + 15) / 16));
      p.keff_div[v] = make_fastdiv ((uint32_t) ks[v]);
      p.nseg_div[v] = make_fastdiv ((uint32_t) ((run_v + p.seg - 1) / p.seg));
    }
    /* a step of lane_w elements along a run, in the mixed radix (x0, x1) of its dims */
    if (p.n_run == 1) {
      p.step_s = p.lane_w * p.run_in[0];
    } else {
      const uint32_t x0 = p.run_size[0], x1 = p.run_size[1];
      p.step_c0 = p.lane_w % x0;
      uint32_t rest = p.lane_w / x0, top = 0;
      if (p.n_run > 2) {
        top = rest / x1;
        rest = rest % x1;
      }
      p.step_c1 = rest;
      p.step_s = p.step_c0 * p.run_in[0] + p.step_c1 * p.run_in[1] + top * p.run_in[2];
      p.wrap0_s = p.run_in[1] - x0 * p.run_in[0];
      p.wrap1_s = p.n_run > 2 ? p.run_in[2] - x1 * p.run_in[1] : 0u;
    }
  }
  const size_t smem = (size_t) (k * dims[S].in * es) + 48;
  if (dry_run)
    return 0;
  if (p.lanes)
    switch (es) {
      case 3: return launch_t<3, true> (p, in, out, smem, cfg, stream);
      case 6: return launch_t<6, true> (p, in, out, smem, cfg, stream);
      case 12: return launch_t<12, true> (p, in, out, smem, cfg, stream);
      case 1: return launch_t<1, true> (p, in, out, smem, cfg, stream);
      case 2: return launch_t<2, true> (p, in, out, smem, cfg, stream);
      case 4: return launch_t<4, true> (p, in, out, smem, cfg, stream);
      default: return launch_t<8, true> (p, in, out, smem, cfg, stream);
    }
  switch (es) {
    case 1: return launch_t<1, false> (p, in, out, smem, cfg, stream);
    case 2: return launch_t<2, false> (p, in, out, smem, cfg, stream);
    case 4: return launch_t<4, false> (p, in, out, smem, cfg, stream);
    default: return launch_t<8, false> (p, in, out, smem, cfg, stream);
  }
}

}  // namespace nnsb
