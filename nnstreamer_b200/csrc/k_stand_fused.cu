/*
 * k_stand_fused.cu -- stand (standardisation) of a BATCH of float32 tensors with ONE read of
 * the input from HBM: a persistent, warp-specialised kernel that keeps each tensor on the SMs
 * between its statistics pass and its normalisation pass -- in shared memory AND in TENSOR
 * MEMORY (TMEM), so that two and a half tensors are resident at any time.
 *
 * Reference semantics: gst_tensor_transform_stand (gsttensor_transform.c:1512-1619) +
 * gst_tensor_data_raw_average / _std [_per_channel] (tensor_data.c:314-493), see k_stand.cuh.
 *
 * Why: the three-launch scheme (k_stand.cuh) reads the input twice -- 12 B of DRAM traffic per
 * element for 8 algorithmic bytes -- and is therefore capped at 2/3 of the HBM roof.  A 1080p
 * RGB float32 frame is 24.9 MB; split over the 148 SMs that is 168 KB per SM, which fits in the
 * 227 KB of shared memory of one CTA -- but only once, and the statistics of a tensor need a
 * grid-wide reduction (~3 us of latency) before it can be normalised; with one resident tensor
 * the SMs sit idle during it (measured: 13.6 us per tensor, 3.6 TB/s).  The 256 KB of tensor
 * memory next to the tensor cores are idle in this library (nothing here is a contraction), and
 * tcgen05.st / tcgen05.ld move registers to and from it.  So:
 *
 *   grid    one CTA per SM, launched cooperatively (all CTAs must be co-resident: they wait for
 *           one another); CTA b owns the same slice of every tensor of the batch: tiles
 *           [first_b, first_b + n_b) of TILE = 3072 elements (12 KB; a multiple of 1..4 channels
 *           and of the thread strides below, so a thread's channel phase never changes).
 *   places  a tile of a tensor is PARKED, from the moment it is loaded until it has been
 *           normalised, in one of 35 places that are used round robin: 21 in TMEM (24 columns x
 *           128 lanes each) and 14 in shared memory.  35 places = 2.5 tensors of 13-14 tiles: the
 *           loads run up to 2.5 tensors ahead of the normalisation, which takes the grid-wide
 *           reduction off the critical path.  Four more shared-memory slots are the landing
 *           FIFO of the tiles bound for TMEM.  Every place / slot has its mbarrier(s).
 *   warp 0  (producer) one elected thread issues cp.async.bulk global -> shared for the tiles of
 *           tensor 0, 1, 2, ... in order as their place (or a landing slot) becomes free (+ an L2
 *           prefetch a few tiles ahead).
 *   8 warps (statistics) read every tile once as it lands: fp64 sum (x - K) and sum (x - K)^2
 *           per channel, K = first element of the channel (the shifted one-pass formula of
 *           k_stand.cuh).  Tiles bound for TMEM go there with tcgen05.st straight from the
 *           registers they were just loaded into, and the landing slot is freed.  After the last
 *           tile of a tensor each warp leaves its sums in shared memory and arrives on
 *           `stats_done[f & 1]`.
 *   warp 1  (control) adds the warp sums, writes the CTA's partial to global scratch and bumps one
 *           counter per tensor (fence + atomicAdd).  The CTA that arrives LAST reads the 148
 *           partials, folds them in a fixed order, computes {mean, 1/std} and publishes them
 *           behind a flag; everybody polls that flag and copies the one result line (148 CTAs
 *           each reading all partials made the reduction a 4-10 us hot spot in L2).
 *   8 warps (normalisation) wait for `ready`, fetch the tensor's tiles again -- tcgen05.ld from
 *           TMEM or ld.shared from a parking slot, freeing the place -- fp64
 *           (x - mean) * (1 / std), |.|, convert, 128-bit stores to global.
 *
 * DRAM sees one read + one write = the algorithmic 8 B/element.  A thread of a statistics warp
 * and the thread of a normalisation warp with the same (warp % 4, lane) own the same vectors of
 * every tile (a warp reaches only its own quarter of the TMEM lanes), which also fixes the TMEM
 * layout: place p = columns [24 p, 24 p + 24), 128 lanes x 24 words = 3072 elements.
 *
 * Restrictions (else the three-launch path is used, nnsb_runtime.cu): float32 in and out,
 * 1..4 channels, frame_elems % 4 == 0, 16-byte aligned bases, 4 <= tiles per CTA <= 16.
 * Summation order differs from k_stand.cuh (tiles instead of a grid-stride), so results agree
 * with it within stand's tolerance, not bit for bit; the normalisation arithmetic is identical.
 */
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "nnsb_device.cuh"
#include "nnsb_kernels.h"

namespace nnsb {

namespace {

constexpr int TILE = 3072;                      /* elements per tile */
constexpr int TILE_BYTES = TILE * 4;
constexpr int LAND = 4;                         /* landing FIFO slots (shared memory) */
constexpr int PARK = 14;                        /* parking slots in shared memory */
constexpr int NSLOT = LAND + PARK;              /* 18 x 12 KB */
constexpr int TPLACES = 21;                     /* parking places in TMEM: 21 x 24 of the 512 columns */
constexpr int PLACES = TPLACES + PARK;          /* 35 */
constexpr int MAX_TILES = 16;                   /* per CTA and tensor; must stay well below PLACES */
constexpr int STATS_WARPS = 8, APPLY_WARPS = 8; /* two of each per TMEM lane quarter */
constexpr int STATS_THREADS = STATS_WARPS * 32;
constexpr int APPLY_THREADS = APPLY_WARPS * 32;
constexpr int THREADS = 64 + STATS_THREADS + APPLY_THREADS;     /* producer + control + ... = 576 */
constexpr int VPT = 3;                          /* vectors per tile and thread: 768 / 256 */
constexpr int TILE_COLS = 24;                   /* TMEM columns per tile: 2 halves x VPT x 4 words */
constexpr int TMEM_COLS = 512;
constexpr int MAX_CTAS = 160;
constexpr int PUB_SLOTS = 4;                    /* scratch slots cycled by tensor index */
constexpr int LOOKAHEAD_DEFAULT = 8;            /* tiles the L2 prefetch runs ahead of the ring */

__device__ __forceinline__ uint32_t smem_u32 (const void *p)
{
  return (uint32_t) __cvta_generic_to_shared (p);
}
__device__ __forceinline__ void mbar_init (uint64_t *bar, uint32_t count)
{
  asm volatile ("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r" (smem_u32 (bar)), "r" (count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx (uint64_t *bar, uint32_t bytes)
{
  asm volatile ("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
      :: "r" (smem_u32 (bar)), "r" (bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive (uint64_t *bar)
{
  asm volatile ("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r" (smem_u32 (bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait (uint64_t *bar, uint32_t parity)
{
  const uint32_t addr = smem_u32 (bar);
  uint32_t done;
  do {
    asm volatile (
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r" (done) : "r" (addr), "r" (parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void bulk_g2s (void *smem_dst, const void *gmem_src, uint32_t bytes,
    uint64_t *bar)
{
  asm volatile ("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      :: "r" (smem_u32 (smem_dst)), "l" (gmem_src), "r" (bytes), "r" (smem_u32 (bar)) : "memory");
}
__device__ __forceinline__ void bulk_prefetch_l2 (const void *gmem_src, uint32_t bytes)
{
  asm volatile ("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l" (gmem_src), "r" (bytes) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_u32 (const uint32_t *p)
{
  uint32_t v;
  asm volatile ("ld.acquire.gpu.global.u32 %0, [%1];" : "=r" (v) : "l" (p) : "memory");
  return v;
}
/* tensor memory: a warp reaches lanes [32 (warp % 4), +32); thread i of the warp = lane 32 q + i,
 * four consecutive 32-bit columns per instruction */
__device__ __forceinline__ void tmem_st4 (uint32_t taddr, const float4 &v)
{
  asm volatile ("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};"
      :: "r" (taddr), "r" (__float_as_uint (v.x)), "r" (__float_as_uint (v.y)),
         "r" (__float_as_uint (v.z)), "r" (__float_as_uint (v.w)) : "memory");
}
__device__ __forceinline__ float4 tmem_ld4 (uint32_t taddr)
{
  uint32_t a, b, c, d;
  asm volatile ("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
      : "=r" (a), "=r" (b), "=r" (c), "=r" (d) : "r" (taddr) : "memory");
  return make_float4 (__uint_as_float (a), __uint_as_float (b), __uint_as_float (c), __uint_as_float (d));
}

struct FusedArgs
{
  const float *in;
  float *out;
  uint64_t frame_elems;
  uint32_t batch;
  uint32_t n_tiles;             /* tiles per tensor (the last one may be short) */
  uint32_t tiles_lo, n_hi;      /* CTAs [0, n_hi) own tiles_lo + 1 tiles, the others tiles_lo */
  uint32_t *counter;            /* [PUB_SLOTS] arrivals, then [PUB_SLOTS] flags (one per 128-byte line), zero on entry */
  double *partial;              /* [PUB_SLOTS][4][2][gridDim.x] */
  double *final_ms;             /* [PUB_SLOTS][4][4]: mean, 1/std, std, - */
  uint32_t lookahead;
  double *stats;                /* [batch][nch][2] = {mean, std}, for nnsb_plan_stand_stats */
  int default_mode;
  unsigned long long *dbg;      /* NNSB_STAND_TRACE: [cta][tensor][TRACE_POINTS] globaltimer ns, else NULL */
};

constexpr int TRACE_POINTS = 10;
__device__ __forceinline__ void
trace (const FusedArgs &a, uint32_t cta, uint32_t f, int point)
{
  if (a.dbg) {
    unsigned long long t;
    asm volatile ("mov.u64 %0, %%globaltimer;" : "=l" (t));
    a.dbg[((size_t) cta * a.batch + f) * TRACE_POINTS + point] = t;
  }
}

struct __align__ (16) Shared
{
  uint64_t land_full[LAND], land_empty[LAND];
  uint64_t park_full[PARK], park_empty[PARK];
  uint64_t tmem_empty[TPLACES]; /* the normalisation warps have read the tile out of that place */
  uint64_t ready[2];            /* {mean, 1/std} of tensor f is in ms[f & 1] */
  uint64_t stats_done[2];       /* the warp sums of tensor f are in red[f & 1] */
  double ms[2][4][2];           /* mean, 1/std per channel */
  double red[2][2][4][STATS_WARPS];
  uint32_t tmem_base;
};

/* one element: the arithmetic of stand_one (k_stand.cuh) */
template <bool DEFAULT_MODE> __device__ __forceinline__ float
norm_one (float x, double mean, double rsd)
{
  const double d = __dsub_rn ((double) x, mean);
  return DEFAULT_MODE ? (float) fabs (__dmul_rn (d, rsd)) : (float) d;
}

template <int NCH>
__global__ void __launch_bounds__ (THREADS, 1)
stand_fused_kernel (FusedArgs a)
{
  extern __shared__ __align__ (128) unsigned char smem_raw[];
  float *ring = reinterpret_cast<float *> (smem_raw);
  Shared &sh = *reinterpret_cast<Shared *> (smem_raw + (size_t) NSLOT * TILE_BYTES);

  const uint32_t cta = blockIdx.x, n_cta = gridDim.x;
  const uint32_t my_tiles = a.tiles_lo + (cta < a.n_hi ? 1u : 0u);
  const uint32_t first_tile = cta * a.tiles_lo + (cta < a.n_hi ? cta : a.n_hi);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < LAND; i++) {
      mbar_init (&sh.land_full[i], 1);
      mbar_init (&sh.land_empty[i], STATS_WARPS);
    }
    for (int i = 0; i < PARK; i++) {
      mbar_init (&sh.park_full[i], 1);
      mbar_init (&sh.park_empty[i], APPLY_WARPS);
    }
    for (int i = 0; i < TPLACES; i++)
      mbar_init (&sh.tmem_empty[i], APPLY_WARPS);
    mbar_init (&sh.ready[0], 1);
    mbar_init (&sh.ready[1], 1);
    mbar_init (&sh.stats_done[0], STATS_WARPS);
    mbar_init (&sh.stats_done[1], STATS_WARPS);
    asm volatile ("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    /* all 512 columns: this CTA has the SM to itself */
    asm volatile ("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
        :: "r" (smem_u32 (&sh.tmem_base)), "n" (TMEM_COLS) : "memory");
    asm volatile ("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile ("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads ();
  asm volatile ("tcgen05.fence::after_thread_sync;" ::: "memory");

  /* elements of tile t of a tensor: all tiles are full except possibly the last one.  (No
   * division anywhere below: slot and phase of the ring advance incrementally -- a 64-bit
   * modulo per tile and thread costs more than the tile's arithmetic.) */
  const uint32_t last_tile = a.n_tiles - 1;
  const uint32_t last_elems = (uint32_t) (a.frame_elems - (uint64_t) last_tile * TILE);
  auto tile_elems = [&] (uint32_t t) -> uint32_t { return t == last_tile ? last_elems : (uint32_t) TILE; };

  /* Places in use: two and a half tensors' worth of tiles, TMEM first.  Not more: the control
   * warp reuses ms[f & 1] / ready[f & 1] for tensor f + 2, which is safe because tensor f + 2
   * cannot be loaded completely before every normalisation warp has freed the first tiles of
   * tensor f -- i.e. has long read ms[f & 1]. */
  const uint32_t n_places = (5 * my_tiles / 2 < (uint32_t) PLACES) ? 5 * my_tiles / 2 : (uint32_t) PLACES;
  const uint32_t n_tplaces = n_places < (uint32_t) TPLACES ? n_places : (uint32_t) TPLACES;

  /* statistics warp 2 + i and normalisation warp 10 + i: TMEM quarter q = warp % 4, half h = i / 4;
   * thread (q, h, lane) owns the vectors v = 192 q + 96 h + 32 k + lane, k = 0..2, of every tile,
   * kept in TMEM columns 24 j + 12 h + 4 k .. + 3 of lanes 32 q + lane */
  const uint32_t role_idx = warp >= 2 + STATS_WARPS ? warp - 2 - STATS_WARPS : warp - 2;
  const uint32_t quarter = warp & 3, half = (role_idx >> 2) & 1;
  const uint32_t v0 = quarter * 192 + half * 96 + lane;
  const uint32_t tmem_thread = sh.tmem_base + ((quarter * 32) << 16) + half * 12;
  /* lane k' of vector v0 + 32 k belongs to channel (4 v0 + 128 k + k') % NCH = (c0 + 2 k + k') % NCH
   * for 3 channels (128 = 2 mod 3) and (c0 + k') % NCH for 1, 2, 4 */
  const uint32_t c0 = (v0 * 4) % NCH;
  constexpr int KROT = (NCH == 3) ? 2 : 0;

  if (warp == 0) {
    /* ------------------------------ producer ------------------------------ */
    if (lane == 0 && my_tiles) {
      const uint32_t total = a.batch * my_tiles;        /* tiles this CTA moves (fits 32 bits) */
      /* the L2 prefetch cursor (tensor pf_f, tile pf_j) runs `lookahead` tiles ahead of q */
      uint32_t pf = 0, pf_f = 0, pf_j = 0;
      /* the tile sequence walks the 35 places round robin: pl = place, cyc = how often it has
       * been used; ls / lround the same for the landing slots of the TMEM-bound tiles */
      uint32_t f = 0, j = 0, pl = 0, cyc = 0, ls = 0, lround = 0;
      for (uint32_t q = 0; q < total; q++) {
        const uint32_t pf_end = (total - q > a.lookahead) ? q + a.lookahead : total;
        for (; pf < pf_end; pf++) {
          if (pf >= q) {
            const uint32_t t = first_tile + pf_j;
            bulk_prefetch_l2 (a.in + (uint64_t) pf_f * a.frame_elems + (uint64_t) t * TILE, tile_elems (t) * 4);
          }
          if (++pf_j == my_tiles) {
            pf_j = 0;
            pf_f++;
          }
        }
        uint64_t *full;
        uint32_t slot;
        if (pl < n_tplaces) {
          if (lround)
            mbar_wait (&sh.land_empty[ls], (lround - 1) & 1);
          full = &sh.land_full[ls];
          slot = ls;
          if (++ls == LAND) {
            ls = 0;
            lround++;
          }
        } else {
          const uint32_t ps = pl - n_tplaces;
          if (cyc)
            mbar_wait (&sh.park_empty[ps], (cyc - 1) & 1);
          full = &sh.park_full[ps];
          slot = LAND + ps;
        }
        if (++pl == n_places) {
          pl = 0;
          cyc++;
        }
        const uint32_t t = first_tile + j;
        const uint32_t bytes = tile_elems (t) * 4;
        mbar_expect_tx (full, bytes);
        bulk_g2s (ring + (size_t) slot * TILE, a.in + (uint64_t) f * a.frame_elems + (uint64_t) t * TILE,
            bytes, full);
        if (j == 0)
          trace (a, cta, f, 8);
        if (++j == my_tiles) {
          trace (a, cta, f, 9);
          j = 0;
          f++;
        }
      }
    }
  } else if (warp == 1) {
    /* ------------------------------ control ------------------------------ */
    for (uint32_t f = 0; f < a.batch; f++) {
      const uint32_t pub = f % PUB_SLOTS;
      const float *base = a.in + (uint64_t) f * a.frame_elems;
      const double k0 = lane < NCH ? (double) __ldg (base + lane) : 0.0;       /* issued early */
      mbar_wait (&sh.stats_done[f & 1], (f >> 1) & 1);
      if (lane == 0)
        trace (a, cta, f, 4);
      /* lane (c, k): channel c, k = 0 sum / 1 sum of squares */
      if (lane < NCH * 2) {
        const uint32_t c = lane >> 1, k = lane & 1;
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < STATS_WARPS; w++)
          t += sh.red[f & 1][k][c][w];
        a.partial[(((size_t) pub * 4 + c) * 2 + k) * n_cta + cta] = t;
      }
      __syncwarp ();
      uint32_t last = 0;
      if (lane == 0) {
        __threadfence ();
        /* this slot's (f / PUB_SLOTS)-th use: everybody has arrived at n_cta * (uses so far) */
        const uint32_t target = n_cta * (f / PUB_SLOTS + 1);
        last = atomicAdd (a.counter + pub * 32, 1u) == target - 1;
        trace (a, cta, f, 5);
      }
      last = __shfl_sync (0xffffffffu, last, 0);
      if (last) {
        /* the last CTA to arrive folds the n_cta partials: lane l takes CTAs l, l + 32, ... in
         * order, then a fixed shuffle tree -- the result does not depend on who is last */
        __threadfence ();
        double tot[NCH][2];
#pragma unroll
        for (int c = 0; c < NCH; c++)
#pragma unroll
          for (int k = 0; k < 2; k++) {
            const double *p = a.partial + (((size_t) pub * 4 + c) * 2 + k) * n_cta;
            double v[(MAX_CTAS + 31) / 32];
#pragma unroll
            for (int i = 0; i < (MAX_CTAS + 31) / 32; i++)
              v[i] = (lane + i * 32 < n_cta) ? __ldcg (p + lane + i * 32) : 0.0;
            double t = 0.0;
#pragma unroll
            for (int i = 0; i < (MAX_CTAS + 31) / 32; i++)
              t += v[i];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
              t += __shfl_xor_sync (0xffffffffu, t, o);
            tot[c][k] = t;
          }
        if (lane < NCH) {
          double s1 = 0.0, s2 = 0.0;
#pragma unroll
          for (int c = 0; c < NCH; c++)
            if (lane == (uint32_t) c) {
              s1 = tot[c][0];
              s2 = tot[c][1];
            }
          const double n = (double) (a.frame_elems / NCH);
          const double mean = k0 + s1 / n;
          double sd = 0.0;
          if (a.default_mode) {
            const double var = (s2 - s1 * (s1 / n)) / n;
            /* tensor_data.c:438,488: std = (sum != 0) ? sqrt (sum) : 1e-10 */
            sd = (var > 0.0) ? sqrt (var) : 1e-10;
          }
          double *fm = a.final_ms + ((size_t) pub * 4 + lane) * 4;
          fm[0] = mean;
          fm[1] = a.default_mode ? __drcp_rn (sd) : 0.0;
          fm[2] = sd;
          a.stats[((size_t) f * NCH + lane) * 2 + 0] = mean;
          a.stats[((size_t) f * NCH + lane) * 2 + 1] = sd;
        }
        __syncwarp ();
        if (lane == 0) {
          __threadfence ();
          asm volatile ("st.release.gpu.global.u32 [%0], %1;"
              :: "l" (a.counter + (PUB_SLOTS + pub) * 32), "r" (f + 1) : "memory");
        }
      }
      if (lane == 0) {
        while (ld_acquire_u32 (a.counter + (PUB_SLOTS + pub) * 32) != f + 1)
          __nanosleep (32);
        trace (a, cta, f, 6);
      }
      __syncwarp ();
      if (lane < NCH) {
        /* ms[f & 1] was last read by the normalisation of tensor f - 2, which is over: the
         * places its tiles lived in had to be free before tensor f could be loaded completely */
        const double *fm = a.final_ms + ((size_t) pub * 4 + lane) * 4;
        sh.ms[f & 1][lane][0] = __ldcg (fm);
        sh.ms[f & 1][lane][1] = __ldcg (fm + 1);
      }
      __syncwarp ();
      if (lane == 0) {
        trace (a, cta, f, 7);
        mbar_arrive (&sh.ready[f & 1]);
      }
    }
  } else if (warp < 2 + STATS_WARPS) {
    /* ------------------------------ statistics ------------------------------ */
    uint32_t pl = 0, cyc = 0, ls = 0, lround = 0;       /* as in the producer */
    for (uint32_t f = 0; f < a.batch; f++) {
      const float *base = a.in + (uint64_t) f * a.frame_elems;
      double kk[NCH], a1[NCH], a2[NCH];
#pragma unroll
      for (int r = 0; r < NCH; r++) {
        kk[r] = (double) __ldg (base + (c0 + r) % NCH);
        a1[r] = 0.0;
        a2[r] = 0.0;
      }
      for (uint32_t j = 0; j < my_tiles; j++) {
        const bool to_tmem = pl < n_tplaces;
        const uint32_t place = pl, use = cyc;
        uint64_t *full, *land_empty = nullptr;
        uint32_t slot, parity;
        if (to_tmem) {
          full = &sh.land_full[ls];
          land_empty = &sh.land_empty[ls];
          slot = ls;
          parity = lround & 1;
          if (++ls == LAND) {
            ls = 0;
            lround++;
          }
        } else {
          full = &sh.park_full[pl - n_tplaces];
          slot = LAND + pl - n_tplaces;
          parity = cyc & 1;
        }
        if (++pl == n_places) {
          pl = 0;
          cyc++;
        }
        mbar_wait (full, parity);
        if (role_idx == 0 && lane == 0 && j + 1 == my_tiles)
          trace (a, cta, f, 2);
        const float4 *tile = reinterpret_cast<const float4 *> (ring + (size_t) slot * TILE);
        const uint32_t nv = tile_elems (first_tile + j) / 4;
        float4 x[VPT];
#pragma unroll
        for (int k = 0; k < VPT; k++) {
          const uint32_t v = v0 + 32 * k;
          x[k] = v < nv ? tile[v] : make_float4 (0.f, 0.f, 0.f, 0.f);
        }
        if (to_tmem) {
          /* the tile that was parked in this place 35 tiles ago must have been normalised */
          if (use)
            mbar_wait (&sh.tmem_empty[place], (use - 1) & 1);
          asm volatile ("tcgen05.fence::after_thread_sync;" ::: "memory");
          __syncwarp ();                        /* the spin loops above may have split the warp */
#pragma unroll
          for (int k = 0; k < VPT; k++)
            tmem_st4 (tmem_thread + place * TILE_COLS + 4 * k, x[k]);
        }
#pragma unroll
        for (int k = 0; k < VPT; k++) {
          const uint32_t v = v0 + 32 * k;
          if (v < nv) {
            const float e[4] = { x[k].x, x[k].y, x[k].z, x[k].w };
#pragma unroll
            for (int i = 0; i < 4; i++) {
              const int r = (KROT * k + i) % NCH;
              const double d = __dsub_rn ((double) e[i], kk[r]);
              a1[r] += d;
              a2[r] = fma (d, d, a2[r]);
            }
          }
        }
        if (to_tmem) {
          /* the landing slot is only a FIFO: free it as soon as the copy is out */
          __syncwarp ();
          asm volatile ("tcgen05.wait::st.sync.aligned;" ::: "memory");
          __syncwarp ();
          if (lane == 0)
            mbar_arrive (land_empty);
        }
      }
      /* actual-channel order, a fixed shuffle tree per warp; the control warp adds the warps.
       * red[f & 1] was consumed (tensor f - 2) long before these tiles could even be loaded */
#pragma unroll
      for (int c = 0; c < NCH; c++) {
        double s1 = 0.0, s2 = 0.0;
#pragma unroll
        for (int r = 0; r < NCH; r++)
          if ((c0 + r) % NCH == (uint32_t) c) {
            s1 = a1[r];
            s2 = a2[r];
          }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          s1 += __shfl_xor_sync (0xffffffffu, s1, o);
          s2 += __shfl_xor_sync (0xffffffffu, s2, o);
        }
        if (lane == 0) {
          sh.red[f & 1][0][c][role_idx] = s1;
          sh.red[f & 1][1][c][role_idx] = s2;
        }
      }
      asm volatile ("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp ();
      if (role_idx == 0 && lane == 0)
        trace (a, cta, f, 3);
      if (lane == 0)
        mbar_arrive (&sh.stats_done[f & 1]);
    }
  } else {
    /* ------------------------------ normalisation ------------------------------ */
    uint32_t pl = 0;
    for (uint32_t f = 0; f < a.batch; f++) {
      mbar_wait (&sh.ready[f & 1], (f >> 1) & 1);
      if (role_idx == 0 && lane == 0)
        trace (a, cta, f, 0);
      asm volatile ("tcgen05.fence::after_thread_sync;" ::: "memory");
      double mean[NCH], rsd[NCH];
#pragma unroll
      for (int r = 0; r < NCH; r++) {
        mean[r] = sh.ms[f & 1][(c0 + r) % NCH][0];
        rsd[r] = sh.ms[f & 1][(c0 + r) % NCH][1];
      }
      float *obase = a.out + (uint64_t) f * a.frame_elems;
      for (uint32_t j = 0; j < my_tiles; j++) {
        const bool from_tmem = pl < n_tplaces;
        const uint32_t place = pl;
        if (++pl == n_places)
          pl = 0;
        const uint32_t t = first_tile + j;
        const uint32_t nv = tile_elems (t) / 4;
        float4 *dst = reinterpret_cast<float4 *> (obase + (uint64_t) t * TILE);
        float4 x[VPT];
        if (from_tmem) {
          __syncwarp ();
#pragma unroll
          for (int k = 0; k < VPT; k++)
            x[k] = tmem_ld4 (tmem_thread + place * TILE_COLS + 4 * k);
          asm volatile ("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          /* the values are in registers: the place can take another tile */
          asm volatile ("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp ();
          if (lane == 0)
            mbar_arrive (&sh.tmem_empty[place]);
        } else {
          const float4 *tile = reinterpret_cast<const float4 *> (ring + (size_t) (LAND + place - n_tplaces) * TILE);
#pragma unroll
          for (int k = 0; k < VPT; k++) {
            const uint32_t v = v0 + 32 * k;
            x[k] = v < nv ? tile[v] : make_float4 (0.f, 0.f, 0.f, 0.f);
          }
          __syncwarp ();
          if (lane == 0)
            mbar_arrive (&sh.park_empty[place - n_tplaces]);
        }
#pragma unroll
        for (int k = 0; k < VPT; k++) {
          const uint32_t v = v0 + 32 * k;
          if (v < nv) {
            const float e[4] = { x[k].x, x[k].y, x[k].z, x[k].w };
            float y[4];
#pragma unroll
            for (int i = 0; i < 4; i++) {
              const int r = (KROT * k + i) % NCH;
              y[i] = a.default_mode ? norm_one<true> (e[i], mean[r], rsd[r])
                  : norm_one<false> (e[i], mean[r], rsd[r]);
            }
            asm volatile ("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                :: "l" (dst + v), "r" (__float_as_uint (y[0])), "r" (__float_as_uint (y[1])),
                   "r" (__float_as_uint (y[2])), "r" (__float_as_uint (y[3])) : "memory");
          }
        }
      }
      if (role_idx == 0 && lane == 0)
        trace (a, cta, f, 1);
    }
  }

  /* everybody is done with TMEM: give it back */
  __syncwarp ();
  asm volatile ("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads ();
  if (warp == 1) {
    asm volatile ("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile ("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;"
        :: "r" (sh.tmem_base), "n" (TMEM_COLS) : "memory");
  }
}

size_t
shared_bytes ()
{
  return (size_t) NSLOT * TILE_BYTES + sizeof (Shared);
}

struct Geometry
{
  uint32_t n_tiles, grid, tiles_lo, n_hi;
};

bool
geometry (size_t frame_elems, const LaunchCfg &cfg, Geometry *g)
{
  if (cfg.sm_count < 1 || cfg.sm_count > MAX_CTAS)
    return false;
  const uint64_t n_tiles = (frame_elems + TILE - 1) / TILE;
  uint64_t grid = (uint64_t) cfg.sm_count;
  if (n_tiles < grid * 4)
    return false;               /* too little work per CTA: the hand-shake would dominate */
  const uint64_t lo = n_tiles / grid, hi = n_tiles % grid;
  if (lo + (hi ? 1 : 0) > (uint64_t) MAX_TILES)
    return false;               /* the slice must fit in the ring (with two slots to spare) and in TMEM */
  g->n_tiles = (uint32_t) n_tiles;
  g->grid = (uint32_t) grid;
  g->tiles_lo = (uint32_t) lo;
  g->n_hi = (uint32_t) hi;
  return true;
}

template <int NCH> int
launch_n (const FusedArgs &args, uint32_t grid, cudaStream_t stream)
{
  static int configured[64];
  int dev = 0;
  cudaGetDevice (&dev);
  const size_t smem = shared_bytes ();
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute (stand_fused_kernel<NCH>,
        cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e != cudaSuccess)
      return (int) e;
    configured[dev] = 1;
  }
  FusedArgs a = args;
  void *params[] = { &a };
  return (int) cudaLaunchCooperativeKernel ((const void *) stand_fused_kernel<NCH>, dim3 (grid),
      dim3 (THREADS), params, smem, stream);
}

}  // namespace

bool
stand_fused_supported (int in_type, int out_type, const void *in, const void *out, size_t frame_elems,
    uint32_t nch, uint32_t batch, const LaunchCfg &cfg)
{
  Geometry g;
  if (in_type != NNSB_FLOAT32 || out_type != NNSB_FLOAT32)
    return false;
  if (nch < 1 || nch > 4 || batch < 2 || frame_elems % 4 != 0 || frame_elems % nch != 0)
    return false;
  if (((uintptr_t) in | (uintptr_t) out) & 15)
    return false;
  return geometry (frame_elems, cfg, &g);
}

constexpr size_t COUNTER_BYTES = (size_t) 2 * PUB_SLOTS * 128;  /* arrival counters + flags, one 128-byte line each */
constexpr size_t FINAL_BYTES = (size_t) PUB_SLOTS * 4 * 4 * sizeof (double);

size_t
stand_fused_scratch_bytes (const LaunchCfg &cfg)
{
  /* counters and flags, the published statistics, then the partials */
  return COUNTER_BYTES + FINAL_BYTES + (size_t) PUB_SLOTS * cfg.sm_count * 4 * 2 * sizeof (double);
}

int
launch_stand_fused (int stand_mode, const void *in, void *out, size_t frame_elems, uint32_t nch,
    uint32_t batch, void *scratch, double *stats, const LaunchCfg &cfg, cudaStream_t stream,
    int *n_launches)
{
  Geometry g;
  if (!geometry (frame_elems, cfg, &g))
    return (int) cudaErrorInvalidValue;
  cudaError_t e = cudaMemsetAsync (scratch, 0, COUNTER_BYTES, stream);
  if (e != cudaSuccess)
    return (int) e;
  FusedArgs a;
  a.in = (const float *) in;
  a.out = (float *) out;
  a.frame_elems = frame_elems;
  a.batch = batch;
  a.n_tiles = g.n_tiles;
  a.tiles_lo = g.tiles_lo;
  a.n_hi = g.n_hi;
  a.counter = (uint32_t *) scratch;
  a.final_ms = (double *) ((char *) scratch + COUNTER_BYTES);
  a.partial = (double *) ((char *) scratch + COUNTER_BYTES + FINAL_BYTES);
  a.lookahead = LOOKAHEAD_DEFAULT;
  if (const char *env = getenv ("NNSB_STAND_LOOKAHEAD"))
    a.lookahead = (uint32_t) atoi (env);
  a.stats = stats;
  a.default_mode = stand_mode == NNSB_STAND_DEFAULT;
  a.dbg = nullptr;
  const char *trace_path = getenv ("NNSB_STAND_TRACE");
  const size_t trace_n = (size_t) g.grid * batch * TRACE_POINTS;
  if (trace_path && cudaMalloc ((void **) &a.dbg, trace_n * 8) == cudaSuccess)
    cudaMemsetAsync (a.dbg, 0, trace_n * 8, stream);
  int err;
  switch (nch) {
    case 1: err = launch_n<1> (a, g.grid, stream); break;
    case 2: err = launch_n<2> (a, g.grid, stream); break;
    case 3: err = launch_n<3> (a, g.grid, stream); break;
    case 4: err = launch_n<4> (a, g.grid, stream); break;
    default: err = (int) cudaErrorInvalidValue; break;
  }
  if (a.dbg) {
    /* developer aid (profiles/): per CTA and tensor, when each role reached each point */
    std::vector<unsigned long long> h (trace_n);
    cudaStreamSynchronize (stream);
    cudaMemcpy (h.data (), a.dbg, trace_n * 8, cudaMemcpyDeviceToHost);
    cudaFree (a.dbg);
    if (FILE *fp = fopen (trace_path, "wb")) {
      const unsigned long long hdr[3] = { g.grid, batch, (unsigned long long) TRACE_POINTS };
      fwrite (hdr, 8, 3, fp);
      fwrite (h.data (), 8, trace_n, fp);
      fclose (fp);
    }
  }
  if (n_launches)
    *n_launches = 1;
  return err;
}

}  // namespace nnsb
