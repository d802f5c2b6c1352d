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
constexpr uint32_t SLAB_MAX_BYTES = 96 * 1024;        /* two CTAs per SM still overlap their load and store phases */
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
  /* window (lane kernel only): a slab takes only wt slices of the dimension just below the slab
   * dimension on the input side, so that it can be taller -- k row segments instead of one
   * contiguous range.  Row r of the slab sits in shared memory at r * pitch_b (+ the
   * misalignment of row 0): pitch_b = row stride modulo 16, so every row's 16-byte aligned
   * superset lands 16-byte aligned. */
  int win;
  int idx_win;                          /* index of the window dimension among the run-index dims */
  uint32_t w_size, wt;                  /* its extent, slices per window */
  uint64_t w_in, w_out;                 /* its strides, elements */
  FastDiv nwin_div;                     /* windows per slab position */
  FastDiv weff_div[2];                  /* wt, and the width of the last window */
  uint32_t pitch_b, seg_b;              /* bytes between rows in shared memory; bytes of a full row segment */
  /* lane-per-element store (slab_lanes_kernel): groups of lane_w lanes step along a run by
   * lane_w elements, the whole CTA by 256.  A step in the mixed radix of the run dims, and what
   * it adds to the source offset: */
  int lanes;                            /* use slab_lanes_kernel */
  uint32_t piece;                       /* bytes per bulk copy of the slab load (a multiple of 16) */
  uint32_t lane_w, lane_shift;          /* 32, 16 or 8 lanes = 1 << lane_shift */
  uint32_t n_above[2][2];               /* runs per value of the first run-index dim: [last slab][last window] */
  uint32_t step_c0, step_c1, step_s;    /* lane_w = (.. * x1 + step_c1) * x0 + step_c0 */
  uint32_t bstep_c0, bstep_c1, bstep_s; /* the same for 256 */
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

/* Which slab, where it starts on both sides, how tall it is; then its bytes -- the 16-byte
 * aligned superset of them -- go to shared memory (the element at slab byte b sits at
 * buf[mis + b]) and the CTA waits for them. */
template <int ES, bool BULK = false> __device__ __forceinline__ void
slab_begin (const SlabDev &p, uintptr_t in_lo, uintptr_t in_hi, uint32_t slab, unsigned char *buf,
    uint64_t &out_base, uint32_t &k_eff, uint32_t &mis, uint64_t *bar = nullptr, uint32_t *phase = nullptr,
    uint32_t *wt_eff = nullptr)
{
  uint32_t oi, si, w0 = 0;
  if (BULK && p.win) {
    uint32_t rest, wi;
    divmod (slab, p.nwin_div, rest, wi);
    slab = rest;
    w0 = wi * p.wt;
    *wt_eff = (p.w_size - w0 < p.wt) ? p.w_size - w0 : p.wt;
  }
  divmod (slab, p.slabs_per_outer, oi, si);
  uint64_t in_base = (uint64_t) w0 * p.w_in;    /* elements */
  out_base = (uint64_t) w0 * p.w_out;
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
  k_eff = (p.s_size - o0 < p.k) ? p.s_size - o0 : p.k;
  in_base += (uint64_t) o0 * p.s_in;
  out_base += (uint64_t) o0 * p.s_out;
  const uintptr_t abs0 = in_lo + in_base * ES;
  mis = (uint32_t) (abs0 & 15);

  if (BULK) {
    const uint32_t bar_a = (uint32_t) __cvta_generic_to_shared (bar);
    if (threadIdx.x < 32) {
      /* every lane of the first warp arrives once with the bytes of the bulk copies it issues
       * (the copy engine moves them, no thread touches the bytes); what cannot be copied in
       * bulk -- a 16-byte piece that would reach outside the tensor -- the lane moves itself
       * before it arrives */
      const uint32_t rows = p.win ? k_eff : 1u;
      const uint64_t row_b = (uint64_t) p.s_in * ES;
      const uint32_t seg = p.win ? (uint32_t) ((uint64_t) (*wt_eff) * p.w_in * ES) : (uint32_t) (k_eff * p.s_in) * ES;
      uint32_t tx = 0;
      for (int pass = 0; pass < 2; pass++) {
        for (uint32_t r = threadIdx.x; r < rows; r += 32) {
          const uintptr_t a = abs0 + r * row_b;
          uintptr_t g = a & ~(uintptr_t) 15;
          uint32_t len = ((uint32_t) (a - g) + seg + 15) & ~15u;
          /* row r: element 0 at buf[mis + r * pitch_b], so its aligned superset starts here: */
          unsigned char *dst = buf + mis + r * p.pitch_b - (uint32_t) (a - g);
          /* the first / last 16-byte piece may reach outside the tensor (its very first and last
           * bytes only): those pieces are moved byte by byte, never touching what is not ours */
          if (g < in_lo) {
            if (pass == 0)
              for (uint32_t b = 0; b < 16; b++)
                if (g + b >= in_lo && g + b < in_hi)
                  dst[b] = *reinterpret_cast<const unsigned char *> (g + b);
            g += 16;
            dst += 16;
            len -= 16;
          }
          if (len && g + len > in_hi) {
            if (pass == 0)
              for (uint32_t b = len - 16; b < len; b++)
                if (g + b >= in_lo && g + b < in_hi)
                  dst[b] = *reinterpret_cast<const unsigned char *> (g + b);
            len -= 16;
          }
          if (pass == 0) {
            tx += len;
          } else {
            const uint32_t piece = p.piece;
            for (uint32_t off = 0; off < len; off += piece) {
              const uint32_t n = len - off < piece ? len - off : piece;
              asm volatile ("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                  :: "r" ((uint32_t) __cvta_generic_to_shared (dst + off)), "l" (g + off), "r" (n), "r" (bar_a)
                  : "memory");
            }
          }
        }
        if (pass == 0)
          asm volatile ("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
              :: "r" (bar_a), "r" (tx) : "memory");
      }
    }
    asm volatile ("{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" :: "r" (bar_a), "r" (*phase) : "memory");
    *phase ^= 1;
    return;
  }

  /* cp.async by all threads: the 16-byte aligned superset of the slab */
  const uint32_t slab_bytes = (uint32_t) (k_eff * p.s_in) * ES;
  const uintptr_t ga = abs0 & ~(uintptr_t) 15;
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
}

/* One step of `T` elements along a run for a lane that sits at run coordinates (c0, c1, ..):
 * the source offset moves by a constant plus a correction per carry (at most one per dim). */
template <int NRUN> __device__ __forceinline__ void
run_step (uint32_t &sb, uint32_t &c0, uint32_t &c1, uint32_t step_sb, uint32_t step_c0, uint32_t step_c1,
    uint32_t wrap0_sb, uint32_t wrap1_sb, uint32_t x0, uint32_t x1)
{
  sb += step_sb;
  if (NRUN > 1) {
    c0 += step_c0;
    if (c0 >= x0) {
      c0 -= x0;
      sb += wrap0_sb;
      if (NRUN > 2)
        c1 += 1;
    }
    if (NRUN > 2) {
      c1 += step_c1;
      if (c1 >= x1) {
        c1 -= x1;
        sb += wrap1_sb;
      }
    }
  }
}

/* The slab kernel with a store phase of one element per lane: consecutive lanes write
 * consecutive elements of a run (a group of W lanes stores W * ES contiguous bytes, whatever
 * the alignment of the run) and read their sources from shared memory at the stride the
 * permutation gives them.  Nothing depends on alignment.  Three shapes of loop:
 *   runs of at most W elements   a group of W = 8 / 16 / 32 lanes owns a run, takes one element per
 *                                lane and moves on to the group's next run: bases advance by
 *                                constants, ~8 instructions per run and lane;
 *   runs of at most 512          the same with an inner loop along the run (steps of W);
 *   longer runs                  the whole CTA walks one run after the other in steps of 256.
 * Coordinates inside a run advance with at most one carry per dim (run_step).  NRUN = dims
 * inside a run.  ES may be 3, 6 or 12 (a folded RGB pixel, see launch_permute_slab). */
template <int ES, int NRUN>
__global__ void __launch_bounds__ (SLAB_THREADS, 6)
slab_lanes_kernel (const unsigned char *__restrict__ in, unsigned char *__restrict__ out,
    const __grid_constant__ SlabDev p)
{
  extern __shared__ __align__ (16) unsigned char buf[];
  const uintptr_t in_lo = (uintptr_t) in, in_hi = in_lo + p.total_bytes;
  __shared__ __align__ (8) uint64_t bar;
  const uint32_t tid = threadIdx.x;
  if (tid == 0) {
    asm volatile ("mbarrier.init.shared::cta.b64 [%0], 32;"      /* the lanes of the first warp, slab_begin */
        :: "r" ((uint32_t) __cvta_generic_to_shared (&bar)) : "memory");
    asm volatile ("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads ();
  uint32_t phase = 0;
  const uint32_t W = p.lane_w;
  const uint32_t lane = tid & (W - 1), group = tid >> p.lane_shift, n_groups = SLAB_THREADS >> p.lane_shift;
  const uint32_t x0 = p.run_size[0], x1 = p.run_size[1];        /* radices: always full dims */
  /* where lane `lane` (group loops) and thread `tid` (CTA-wide loop) start inside a run */
  uint32_t lc0 = lane, lc1 = 0, lc2 = 0, tc0 = tid, tc1 = 0, tc2 = 0;
  if (NRUN > 1) {
    divmod (lane, p.run_div[0], lc1, lc0);
    divmod (tid, p.run_div[0], tc1, tc0);
    if (NRUN > 2) {
      uint32_t t = lc1;
      divmod (t, p.run_div[1], lc2, lc1);
      t = tc1;
      divmod (t, p.run_div[1], tc2, tc1);
    }
  }
  uint32_t ls = lc0 * p.run_in[0], ts = tc0 * p.run_in[0];
  if (NRUN > 1) {
    ls += lc1 * p.run_in[1];
    ts += tc1 * p.run_in[1];
  }
  if (NRUN > 2) {
    ls += lc2 * p.run_in[2];
    ts += tc2 * p.run_in[2];
  }
  const uint32_t in0_b = p.n_idx > 0 ? p.idx_in[0] * ES : 0;
  const uint64_t out0_b = p.n_idx > 0 ? p.idx_out[0] * ES : 0;
  const uint32_t g_in_b = n_groups * in0_b;
  const uint64_t g_out_b = n_groups * out0_b;
  const uint32_t wstep_sb = p.step_s * ES, wwrap0 = p.wrap0_s * ES, wwrap1 = p.wrap1_s * ES;
  const uint32_t bstep_sb = p.bstep_s * ES;

  for (uint32_t slab = blockIdx.x; slab < p.n_slabs; slab += gridDim.x) {
    uint64_t out_base;
    uint32_t k_eff, mis, wt_eff = 0;
    slab_begin<ES, true> (p, in_lo, in_hi, slab, buf, out_base, k_eff, mis, &bar, &phase, &wt_eff);
    const unsigned char *src0 = buf + mis;

    uint32_t L = 1;
#pragma unroll
    for (int d = 0; d < NRUN; d++)
      L *= (d == p.run_slab) ? k_eff : p.run_size[d];
    uint32_t n_runs = 1;
#pragma unroll
    for (int d = 0; d < SLAB_DIMS; d++)
      if (d < p.n_idx)
        n_runs *= (d == p.idx_slab) ? k_eff : (p.win && d == p.idx_win) ? wt_eff : p.idx_div[d].d;
    const FastDiv keff_div = p.keff_div[k_eff != p.k];
    const int short_w = p.win && wt_eff != p.wt;
    const FastDiv weff_div = p.weff_div[short_w];

    if (L <= 512) {
      /* groups of W lanes, a run each; from run to run along the first run-index dim only the
       * bases move */
      const uint32_t n0 = p.n_idx > 0 ? (p.idx_slab == 0 ? k_eff : (p.win && p.idx_win == 0) ? wt_eff : p.idx_div[0].d) : 1;
      const uint32_t n_above = p.n_above[k_eff != p.k][short_w];
      const uint32_t mine = n0 > group ? (n0 - group + n_groups - 1) >> (8 - p.lane_shift) : 0;
      if (L <= W && n_above <= 4) {
        /* one element per lane and run, and so few values of the upper run-index dims (the three
         * channel planes of a CHW output) that their offsets stay in registers: the loop over
         * the first run-index dim carries all of them */
        uint32_t in_ob[4];
        uint64_t out_ob[4];
#pragma unroll
        for (int o = 0; o < 4; o++) {
          uint64_t out_o = 0;
          uint32_t in_o = 0, r = o;
#pragma unroll
          for (int d = 1; d < SLAB_DIMS; d++)
            if (d < p.n_idx) {
              uint32_t q, c;
              if (d == p.idx_slab)
                divmod (r, keff_div, q, c);
              else if (p.win && d == p.idx_win)
                divmod (r, weff_div, q, c);
              else
                divmod (r, p.idx_div[d], q, c);
              r = q;
              out_o += (uint64_t) c * p.idx_out[d];
              in_o += c * p.idx_in[d];
            }
          in_ob[o] = in_o * ES;
          out_ob[o] = out_o * ES;
        }
        if (lane < L) {
          uint32_t sb = ls * ES + group * in0_b;
          unsigned char *dp = out + (out_base + lane) * ES + group * out0_b;
#pragma unroll 2
          for (uint32_t i = 0; i < mine; i++) {
#pragma unroll
            for (int o = 0; o < 4; o++)
              if ((uint32_t) o < n_above)
                copy_elem<ES> (dp + out_ob[o], src0 + sb + in_ob[o]);
            sb += g_in_b;
            dp += g_out_b;
          }
        }
      } else
      for (uint32_t o = 0; o < n_above; o++) {
        uint64_t out_o = out_base;
        uint32_t in_o = 0, r = o;
#pragma unroll
        for (int d = 1; d < SLAB_DIMS; d++)
          if (d < p.n_idx) {
            uint32_t q, c;
            if (d == p.idx_slab)
              divmod (r, keff_div, q, c);
            else if (p.win && d == p.idx_win)
              divmod (r, weff_div, q, c);
            else
              divmod (r, p.idx_div[d], q, c);
            r = q;
            out_o += (uint64_t) c * p.idx_out[d];
            in_o += c * p.idx_in[d];
          }
        uint32_t sb = (in_o + ls) * ES + group * in0_b;
        unsigned char *dp = out + (out_o + lane) * ES + group * out0_b;
        if (L <= W) {
          if (lane < L) {
#pragma unroll 4
            for (uint32_t i = 0; i < mine; i++) {
              copy_elem<ES> (dp, src0 + sb);
              sb += g_in_b;
              dp += g_out_b;
            }
          }
        } else {
          /* the lane's position inside a run on the outside, the group's runs on the inside: for
           * one position only the bases move from run to run (the same tight loop as above), and
           * the carry arithmetic along the run is done once per position, not once per element */
          uint32_t s = 0, c0 = lc0, c1 = lc1;
          for (uint32_t j = lane; j < L; j += W) {
            uint32_t sbt = sb + s;
            unsigned char *d = dp + (j - lane) * ES;
#pragma unroll 4
            for (uint32_t i = 0; i < mine; i++) {
              copy_elem<ES> (d, src0 + sbt);
              sbt += g_in_b;
              d += g_out_b;
            }
            run_step<NRUN> (s, c0, c1, wstep_sb, p.step_c0, p.step_c1, wwrap0, wwrap1, x0, x1);
          }
        }
      }
    } else {
      /* long runs: the CTA walks each of them in steps of 256 elements */
      for (uint32_t run = 0; run < n_runs; run++) {
        uint64_t out_r = out_base;
        uint32_t in_r = 0, r = run;
        if (p.n_idx == 1) {
          out_r += (uint64_t) run * p.idx_out[0];
          in_r = run * p.idx_in[0];
        } else
#pragma unroll
        for (int d = 0; d < SLAB_DIMS; d++)
          if (d < p.n_idx) {
            uint32_t q, c;
            if (d == p.idx_slab)
              divmod (r, keff_div, q, c);
            else if (p.win && d == p.idx_win)
              divmod (r, weff_div, q, c);
            else
              divmod (r, p.idx_div[d], q, c);
            r = q;
            out_r += (uint64_t) c * p.idx_out[d];
            in_r += c * p.idx_in[d];
          }
        uint32_t s = (in_r + ts) * ES, c0 = tc0, c1 = tc1;
        unsigned char *d = out + (out_r + tid) * ES;
#pragma unroll 4
        for (uint32_t j = tid; j < L; j += SLAB_THREADS) {
          copy_elem<ES> (d, src0 + s);
          d += SLAB_THREADS * ES;
          run_step<NRUN> (s, c0, c1, bstep_sb, p.bstep_c0, p.bstep_c1, wwrap0, wwrap1, x0, x1);
        }
      }
    }
    __syncthreads ();           /* the buffer is about to be overwritten */
  }
}

template <int ES>
__global__ void __launch_bounds__ (SLAB_THREADS, 5)
permute_slab_kernel (const unsigned char *__restrict__ in, unsigned char *__restrict__ out,
    const __grid_constant__ SlabDev p)
{
  extern __shared__ __align__ (16) unsigned char buf[];
  const uintptr_t in_lo = (uintptr_t) in, in_hi = in_lo + p.total_bytes;

  for (uint32_t slab = blockIdx.x; slab < p.n_slabs; slab += gridDim.x) {
    uint64_t out_base;
    uint32_t k_eff, mis;
    slab_begin<ES> (p, in_lo, in_hi, slab, buf, out_base, k_eff, mis);

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
    {
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
    }
  }
}

static uint64_t
slab_grid (size_t smem, uint64_t by_regs, const SlabDev &p, const LaunchCfg &cfg)
{
  uint64_t per_sm = (200 * 1024) / (smem + 1024);      /* what the shared memory allows */
  if (per_sm > by_regs)
    per_sm = by_regs;
  if (per_sm < 2)
    per_sm = 2;
  if (const char *e = getenv ("NNSB_SLAB_CTAS"))
    per_sm = (uint64_t) atoi (e) > 0 ? (uint64_t) atoi (e) : per_sm;
  const uint64_t grid = (uint64_t) cfg.sm_count * per_sm;
  return grid > p.n_slabs ? p.n_slabs : grid;
}

template <typename K> static int
slab_configure (K kernel, int *configured)
{
  int dev = 0;
  cudaGetDevice (&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute (kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
        (int) (SLAB_MAX_BYTES + 64));
    if (e != cudaSuccess)
      return (int) e;
    configured[dev] = 1;
  }
  return 0;
}

template <int ES> int
launch_chunks (const SlabDev &p, const void *in, void *out, size_t smem, const LaunchCfg &cfg,
    cudaStream_t stream)
{
  static int configured[64];
  if (int r = slab_configure (permute_slab_kernel<ES>, configured))
    return r;
  permute_slab_kernel<ES><<<(unsigned int) slab_grid (smem, 5, p, cfg), SLAB_THREADS, smem, stream>>> (
      (const unsigned char *) in, (unsigned char *) out, p);
  return (int) cudaGetLastError ();
}

template <int ES, int NRUN> int
launch_lanes_n (const SlabDev &p, const void *in, void *out, size_t smem, const LaunchCfg &cfg,
    cudaStream_t stream)
{
  static int configured[64];
  if (int r = slab_configure (slab_lanes_kernel<ES, NRUN>, configured))
    return r;
  slab_lanes_kernel<ES, NRUN><<<(unsigned int) slab_grid (smem, 6, p, cfg), SLAB_THREADS, smem, stream>>> (
      (const unsigned char *) in, (unsigned char *) out, p);
  return (int) cudaGetLastError ();
}

template <int ES> int
launch_lanes (const SlabDev &p, const void *in, void *out, size_t smem, const LaunchCfg &cfg,
    cudaStream_t stream)
{
  switch (p.n_run) {
    case 1: return launch_lanes_n<ES, 1> (p, in, out, smem, cfg, stream);
    case 2: return launch_lanes_n<ES, 2> (p, in, out, smem, cfg, stream);
    default: return launch_lanes_n<ES, 3> (p, in, out, smem, cfg, stream);
  }
}

}  // namespace

/* lanes that share a run of L elements: a warp, or a half / quarter of one when that fills them
 * better (short runs, or runs that are whole multiples of 16 / 8 only) */
static uint32_t
lane_width (uint64_t L)
{
  if (L <= 8)
    return 8;
  if (L <= 16)
    return 16;
  if (L % 32 == 0 || L > 128)
    return 32;
  if (L % 16 == 0)
    return 16;
  if (L <= 64 && L % 8 == 0)
    return 8;
  return 32;
}

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
   * store does not care, so here the pixel can become the element (3, 6 or 12 bytes).  Measured
   * on u16 225x113x3 0:2:1:3 this is slower than leaving the channels as the innermost run dim
   * (three strided 2-byte stores per lane: 1.9 against 2.3 TB/s), so it is opt-in: NNSB_SLAB_FOLD. */
  CDim folded[2 * SLAB_DIMS];
  if (n >= 3 && es <= 4 && dims[0].size == 3 && dims[0].in == 1 && dims[0].out == 1
      && !chunks_forced && getenv ("NNSB_SLAB_FOLD")) {
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
  /* elements of an output run per slice of the slab dimension (0: the slab dimension is not part
   * of the runs) */
  uint64_t per_k = 0;
  bool pre_run[2 * SLAB_DIMS];          /* dims the output runs cover below the slab dimension */
  for (int i = 0; i < n; i++)
    pre_run[i] = false;
  {
    bool fixed[2 * SLAB_DIMS];
    for (int i = 0; i < n; i++)
      fixed[i] = false;
    for (int i = si + 1; i < n; i++)
      fixed[by_in[i]] = true;
    uint64_t prod = 1, next_out = 1;
    int cnt = 0;
    for (int d = 0; d < n && cnt < 3; d++, cnt++) {
      if (fixed[d] || dims[d].out != next_out)
        break;
      if (d == S) {
        per_k = prod;
        break;
      }
      pre_run[d] = true;
      prod *= dims[d].size;
      next_out = dims[d].out * dims[d].size;
    }
  }
  const uint64_t slice_bytes = dims[S].in * (uint64_t) es;
  uint64_t max_bytes = 24 * 1024;       /* sweep: profiles/r2_slab_sweep.md */
  /* ... unless the runs would then be shorter than a 128-byte line: every 32-byte sector of the
   * output is then written in pieces by different slabs, which is harmless while the output
   * stays in the L2 and halves the rate once it does not (f32 227x227x3 2:1:0:3, 8-row slabs:
   * 3.4 TB/s on 64 frames, 1.9 on 512; 16-row slabs: 3.5 and 2.8) */
  if (lanes && per_k && per_k * (max_bytes / slice_bytes) * es < 128) {
    const uint64_t k_line = (128 + per_k * es - 1) / (per_k * es);
    max_bytes = k_line * slice_bytes;
    if (max_bytes > 48 * 1024)          /* 16 rows of 227 floats: 22.4 us / 226 us (64 / 512 frames); 32 rows: 25.3 / 237 */
      max_bytes = 48 * 1024;
  }
  if (const char *e = getenv ("NNSB_SLAB_KB"))
    if (atoi (e) > 0 && (uint64_t) atoi (e) * 1024 <= SLAB_MAX_BYTES)
      max_bytes = (uint64_t) atoi (e) * 1024;
  if (slice_bytes > max_bytes)
    max_bytes = slice_bytes;
  uint64_t k = max_bytes / slice_bytes;
  if (k > dims[S].size)
    k = dims[S].size;
  bool balance = true;          /* equal slabs, unless a height was picked for the lanes */
  /* Window: when the slab dimension is what the output runs run along and a slab of whole slices
   * is only a few of them tall (f32 227x227x3 2:1:0:3: 16 rows = 64-byte runs), take only a part
   * of the dimension below it instead and make the slab taller: runs of >= 512 bytes (or the
   * whole dimension), row segments of >= 256 bytes, about 32 KB per slab. */
  int Wd = -1;
  uint64_t wt = 0, nwin = 1, pitch_b = 0;
  {
    /* measured (profiles/r2_slab_sweep.md): f32 227x227x3 2:1:0:3 on 512 frames 221 -> 167-173 us,
     * u16 225x113x3 0:2:1:3 on 2048 frames 263 -> 207 us; on 64 / 256 frames, whose output stays in
     * the L2, the many small row copies cost more than the longer runs give (24 -> 30 us), hence
     * only for outputs that do not fit the L2 */
    const char *e = getenv ("NNSB_SLAB_WINDOW");
    const bool allowed = e ? atoi (e) != 0 : total_elems * (uint64_t) es >= 100ull * 1000 * 1000;
    const int cand = by_in[si - 1];
    if (allowed && lanes && per_k && es <= 8 && k < dims[S].size && per_k * k * es < 512
        && !pre_run[cand] && dims[cand].size > 1 && !getenv ("NNSB_SLAB_KB")) {
      const uint64_t unit_b = dims[cand].in * (uint64_t) es;    /* one slice of the window dim inside a row */
      const uint64_t k_need = (512 + per_k * es - 1) / (per_k * es), budget = 32 * 1024;
      uint64_t kk = dims[S].size <= 2 * k_need ? dims[S].size : k_need;
      if (const char *r = getenv ("NNSB_SLAB_WIN_ROWS"))
        if (atoi (r) > 0 && (uint64_t) atoi (r) <= dims[S].size)
          kk = (uint64_t) atoi (r);
      uint64_t w = budget / (kk * unit_b);
      const uint64_t w_min = (256 + unit_b - 1) / unit_b;       /* (144-byte segments: 234 us where 264-byte ones take 173) */
      if (w < w_min) {
        w = w_min;
        kk = budget / (w * unit_b + 46);
      }
      if (kk > dims[S].size)
        kk = dims[S].size;
      if (w >= 1 && w < dims[cand].size && kk >= 2 * k) {
        nwin = (dims[cand].size + w - 1) / w;
        w = (dims[cand].size + nwin - 1) / nwin;
        nwin = (dims[cand].size + w - 1) / w;
        const uint64_t ns = (dims[S].size + kk - 1) / kk;
        kk = (dims[S].size + ns - 1) / ns;
        const uint64_t p0 = w * unit_b + 30;
        pitch_b = p0 + ((slice_bytes % 16) + 16 - (p0 % 16)) % 16;      /* = row stride modulo 16 */
        if (kk * pitch_b + 64 <= SLAB_MAX_BYTES) {
          Wd = cand;
          wt = w;
          k = kk;
          balance = false;
        }
      }
    }
  }
  if (lanes && k < dims[S].size && Wd < 0) {
    /* When the slab dimension is part of the output runs, a run is per_k * k elements and lanes
     * idle where that is not a multiple of the group width: among the heights down to half the
     * tallest, take the one that fills the lanes best. */
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
        const uint64_t Lf = per_k * kc, w = lane_width (Lf);
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
  p.idx_win = -1;
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
      p.run_in[p.n_run] = (uint32_t) ((d == S && Wd >= 0) ? pitch_b / es : dims[d].in);
      p.n_run++;
      run_elems *= (d == S ? k : dims[d].size);
      next_out = dims[d].out * dims[d].size;
      if (d == S) {
        p.run_slab = p.n_run - 1;
        if (k < dims[S].size || Wd >= 0)
          in_run = false;       /* (with a window the dim above may not join the run: its extent varies) */
      }
      continue;
    }
    in_run = false;
    if (p.n_idx >= SLAB_DIMS)
      return -1;
    if (d == S)
      p.idx_slab = p.n_idx;
    if (d == Wd)
      p.idx_win = p.n_idx;
    p.idx_div[p.n_idx] = make_fastdiv ((uint32_t) (d == S ? k : d == Wd ? wt : dims[d].size));
    p.idx_in[p.n_idx] = (uint32_t) ((d == S && Wd >= 0) ? pitch_b / es : dims[d].in);
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
  uint64_t n_slabs = nsl * nwin;
  for (int i = 0; i < p.n_outer; i++)
    n_slabs *= p.outer_div[i].d;
  if (Wd >= 0) {
    if (p.idx_win < 0 || p.idx_slab >= 0 || p.run_slab < 0)
      return -1;                /* (the window dimension must enumerate runs, the slab dimension be inside them) */
    p.win = 1;
    p.w_size = (uint32_t) dims[Wd].size;
    p.wt = (uint32_t) wt;
    p.w_in = dims[Wd].in;
    p.w_out = dims[Wd].out;
    p.nwin_div = make_fastdiv ((uint32_t) nwin);
    p.weff_div[0] = make_fastdiv ((uint32_t) wt);
    p.weff_div[1] = make_fastdiv ((uint32_t) (dims[Wd].size - (nwin - 1) * wt));
    p.pitch_b = (uint32_t) pitch_b;
    p.seg_b = (uint32_t) (wt * dims[Wd].in * es);
  }
  if (n_slabs >= (1ull << 31))
    return -1;
  p.n_slabs = (uint32_t) n_slabs;
  p.total_bytes = total_elems * (uint64_t) es;
  {
    const uint64_t k_last = dims[S].size - (nsl - 1) * k;       /* height of the last slab */
    const uint64_t ks[2] = { k, k_last };
    p.piece = 1u << 20;         /* one copy: pieces of 1-16 KB were measured 0-5 % slower (profiles/r2_slab_sweep.md) */
    if (const char *e = getenv ("NNSB_SLAB_PIECE"))
      if (atoi (e) >= 16)
        p.piece = (uint32_t) atoi (e) / 16 * 16;
    p.lane_w = p.lanes ? lane_width (run_elems) : 32;
    p.lane_shift = p.lane_w == 8 ? 3 : p.lane_w == 16 ? 4 : 5;
    for (int v = 0; v < 2; v++)
      for (int u = 0; u < 2; u++) {
        uint64_t above = 1;
        for (int d = 1; d < p.n_idx; d++)
          above *= (d == p.idx_slab) ? ks[v] : (p.win && d == p.idx_win) ? p.weff_div[u].d : p.idx_div[d].d;
        p.n_above[v][u] = (uint32_t) above;
      }
    for (int v = 0; v < 2; v++) {
      uint64_t run_v = run_elems;
      if (p.run_slab >= 0)
        run_v = run_elems / k * ks[v];
      p.cpr_div[v] = make_fastdiv ((uint32_t) ((run_v * es + 15 + 15) / 16));
      p.keff_div[v] = make_fastdiv ((uint32_t) ks[v]);
    }
    /* a step of T elements along a run, in the mixed radix (x0, x1) of its dims */
    auto steps = [&p] (uint32_t T, uint32_t &c0, uint32_t &c1, uint32_t &sv) {
      if (p.n_run == 1) {
        c0 = c1 = 0;
        sv = T * p.run_in[0];
        return;
      }
      const uint32_t x0 = p.run_size[0], x1 = p.run_size[1];
      c0 = T % x0;
      uint32_t rest = T / x0, top = 0;
      if (p.n_run > 2) {
        top = rest / x1;
        rest = rest % x1;
      }
      c1 = rest;
      sv = c0 * p.run_in[0] + c1 * p.run_in[1] + top * p.run_in[2];
    };
    steps (p.lane_w, p.step_c0, p.step_c1, p.step_s);
    steps (SLAB_THREADS, p.bstep_c0, p.bstep_c1, p.bstep_s);
    if (p.n_run > 1) {
      p.wrap0_s = p.run_in[1] - p.run_size[0] * p.run_in[0];
      p.wrap1_s = p.n_run > 2 ? p.run_in[2] - p.run_size[1] * p.run_in[1] : 0u;
    }
  }
  const size_t smem = (Wd >= 0 ? (size_t) (k * pitch_b) : (size_t) (k * dims[S].in * es)) + 48;
  if (dry_run)
    return 0;
  if (p.lanes)
    switch (es) {
      case 3: return launch_lanes<3> (p, in, out, smem, cfg, stream);
      case 6: return launch_lanes<6> (p, in, out, smem, cfg, stream);
      case 12: return launch_lanes<12> (p, in, out, smem, cfg, stream);
      case 1: return launch_lanes<1> (p, in, out, smem, cfg, stream);
      case 2: return launch_lanes<2> (p, in, out, smem, cfg, stream);
      case 4: return launch_lanes<4> (p, in, out, smem, cfg, stream);
      default: return launch_lanes<8> (p, in, out, smem, cfg, stream);
    }
  switch (es) {
    case 1: return launch_chunks<1> (p, in, out, smem, cfg, stream);
    case 2: return launch_chunks<2> (p, in, out, smem, cfg, stream);
    case 4: return launch_chunks<4> (p, in, out, smem, cfg, stream);
    default: return launch_chunks<8> (p, in, out, smem, cfg, stream);
  }
}

}  // namespace nnsb
