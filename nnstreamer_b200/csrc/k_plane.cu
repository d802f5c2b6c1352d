/*
 * k_plane.cu -- K3 fast path: interleaved <-> planar (HWC <-> CHW) through the TMA engine.
 *
 * The video-shaped transposes of tensor_transform -- `transpose 1:2:0:3` / `dimchg 0:2`
 * (C:W:H -> W:H:C, i.e. HWC -> CHW) and the inverse `transpose 2:0:1:3` -- canonicalise to a
 * [P][C] <-> [C][P] transpose with a tiny C (2..4 channels).  The reference does one memcpy
 * of `typesize` bytes per element (gsttensor_transform.c:1417-1432) or of one element-sized
 * block per channel per pixel (:1165-1178).
 *
 * Here no thread ever touches global memory:
 *   - an elected thread streams tiles global -> shared with `cp.async.bulk` (1-D TMA bulk
 *     copy, completion on an mbarrier) through a STAGES-deep ring (3-5, by element size), and shared -> global with
 *     `cp.async.bulk ... bulk_group` (one copy per channel plane);
 *   - the 256 threads only re-order inside shared memory: each reads C x 16 bytes
 *     (16/ES pixels, interleaved) with LDS.128, permutes them in registers and writes C x 16
 *     bytes (one vector per plane) with STS.128.
 * So the instruction stream per element is ~0.5 shared-memory instructions plus the register
 * permute, and the copy engine keeps STAGES x 12 KiB per CTA in flight.
 *
 * Requirements (checked by the launcher, else the generic tiled kernel is used): 16-byte
 * aligned bases, plane rows and outer strides that are multiples of 16 bytes.
 * Algorithmic bytes: 2 * ES per element.
 */
#include "nnsb_device.cuh"
#include "nnsb_kernels.h"
#include "k_perm_common.cuh"

namespace nnsb {

constexpr int PLANE_THREADS = 256;              /* threads that re-order tiles */
constexpr int PLANE_BLOCK = PLANE_THREADS + 32; /* + the copy-engine warp */
/* ring depth per element size, from a sweep on B200 (4K frames, 3 channels): 4-byte elements
 * only move registers, so one CTA per SM with a deep ring is best (5 stages: 6.3 TB/s vs 5.7 with
 * 4 stages and two CTAs); 1-byte elements spend ~50 ALU instructions per thread and tile on the
 * byte shuffles and want more CTAs per SM to hide them (3 stages, three CTAs: 5.6 vs 5.2). */
#ifndef NNSB_PLANE_STAGES_1
#define NNSB_PLANE_STAGES_1 3
#endif
#ifndef NNSB_PLANE_STAGES_2
#define NNSB_PLANE_STAGES_2 3
#endif
#ifndef NNSB_PLANE_STAGES_4
#define NNSB_PLANE_STAGES_4 5
#endif
template <int ES> struct PlaneStages
{
  static constexpr int N = ES == 4 ? NNSB_PLANE_STAGES_4 : ES == 2 ? NNSB_PLANE_STAGES_2
      : NNSB_PLANE_STAGES_1;
};
constexpr int PLANE_MAX_OUTER = NNSB_PERM_MAX_RANK;

struct PlaneDev
{
  uint64_t P;                   /* pixels per plane */
  FastDiv tiles_per_plane;      /* ceil (P / TP) */
  uint32_t n_tiles;             /* tiles_per_plane * prod (outer), < 2^31 */
  FastDiv o_div[PLANE_MAX_OUTER];
  uint64_t plane_stride;        /* elements between channel planes on the planar side */
  int n_outer;
  uint32_t o_size[PLANE_MAX_OUTER];
  uint64_t o_il[PLANE_MAX_OUTER];     /* outer strides (elements) on the interleaved side */
  uint64_t o_pl[PLANE_MAX_OUTER];     /* ... on the planar side */
};

/* ---- PTX wrappers ------------------------------------------------------------------- */
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
__device__ __forceinline__ void bulk_s2g (void *gmem_dst, const void *smem_src, uint32_t bytes)
{
  asm volatile ("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
      :: "l" (gmem_dst), "r" (smem_u32 (smem_src)), "r" (bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit ()
{
  asm volatile ("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N> __device__ __forceinline__ void bulk_wait_read ()
{
  asm volatile ("cp.async.bulk.wait_group.read %0;" :: "n" (N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all ()
{
  asm volatile ("cp.async.bulk.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void fence_async_smem ()
{
  asm volatile ("fence.proxy.async.shared::cta;" ::: "memory");
}

/* ---- the kernel ------------------------------------------------------------------------ */
template <int ES> struct ElemOf;
template <> struct ElemOf<1> { typedef uint8_t type; };
template <> struct ElemOf<2> { typedef uint16_t type; };
template <> struct ElemOf<4> { typedef uint32_t type; };

template <int C, int ES> struct PlaneGeom
{
  static constexpr int G = 16 / ES;                       /* pixels per thread */
  static constexpr int TP = PLANE_THREADS * G;            /* pixels per tile */
  static constexpr int TILE_BYTES = TP * C * ES;          /* 256 * 16 * C */
  static constexpr int PLANE_BYTES = TP * ES;             /* 4096 */
  static constexpr int FIT = (216 * 1024) / (2 * TILE_BYTES);        /* what one SM can hold */
  static constexpr int STAGES = PlaneStages<ES>::N < FIT ? PlaneStages<ES>::N : FIT;
  static constexpr size_t SMEM = (size_t) STAGES * 2 * TILE_BYTES + 3 * STAGES * 8 + 32;
};

/* element i (ES bytes wide) of a little-endian array of 32-bit words; i is a compile-time
 * constant after unrolling, so these become PRMT / plain register moves */
template <int ES> __device__ __forceinline__ uint32_t get_elem (const uint32_t *w, int i)
{
  if constexpr (ES == 4) return w[i];
  else if constexpr (ES == 2) return (w[i >> 1] >> (16 * (i & 1))) & 0xffffu;
  else return (w[i >> 2] >> (8 * (i & 3))) & 0xffu;
}
template <int ES> __device__ __forceinline__ void put_elem (uint32_t *w, int i, uint32_t v)
{
  if constexpr (ES == 4) w[i] = v;
  else if constexpr (ES == 2) w[i >> 1] |= v << (16 * (i & 1));
  else w[i >> 2] |= v << (8 * (i & 3));
}

/* TO_PLANAR: in = interleaved [P][C], out = planar [C][P].  Otherwise the inverse. */
template <int C, int ES, bool TO_PLANAR>
__global__ void __launch_bounds__ (PLANE_BLOCK)
plane_kernel (const unsigned char *__restrict__ in, unsigned char *__restrict__ out,
    const __grid_constant__ PlaneDev p)
{
  typedef PlaneGeom<C, ES> Gm;
  constexpr int G = Gm::G;
  constexpr int PLANE_STAGES = Gm::STAGES;
  extern __shared__ __align__ (128) unsigned char smem[];
  unsigned char *in_buf = smem;                                          /* [STAGES][TILE_BYTES] */
  unsigned char *out_buf = smem + (size_t) PLANE_STAGES * Gm::TILE_BYTES;   /* [STAGES][TILE_BYTES] */
  /* per stage: full  = the input tile has landed (1 arrival + transaction bytes)
   *            empty = the previous bulk store out of out_buf[s] has drained (thread 0 arrives)
   *            done  = all 256 threads have re-ordered the tile (each arrives once)
   * No CTA-wide barrier sits in the steady-state loop: only thread 0 ever waits for the other
   * threads, so the remaining warps run ahead to the next tile as soon as its data is there. */
  uint64_t *full = reinterpret_cast<uint64_t *> (smem + (size_t) PLANE_STAGES * 2 * Gm::TILE_BYTES);
  uint64_t *empty = full + PLANE_STAGES;
  uint64_t *done = empty + PLANE_STAGES;

  const int t = threadIdx.x;
  if (t == 0) {
    for (int s = 0; s < PLANE_STAGES; s++) {
      mbar_init (&full[s], 1);
      mbar_init (&empty[s], 1);
      mbar_init (&done[s], PLANE_THREADS);
    }
    asm volatile ("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads ();
  if (t == 0)
    for (int s = 0; s < PLANE_STAGES; s++)
      mbar_arrive (&empty[s]);          /* every output buffer starts out free (phase 0) */

  /* tile k of this CTA is global tile blockIdx.x + k * gridDim.x */
  auto tile_of = [&] (uint32_t k) { return blockIdx.x + k * gridDim.x; };

  /* global addresses of one tile (only the elected thread needs them) */
  auto locate = [&] (uint32_t tile, uint64_t &il_off, uint64_t &pl_off, uint32_t &npix) {
    uint32_t q, tp;
    divmod (tile, p.tiles_per_plane, q, tp);
    const uint64_t p0 = (uint64_t) tp * Gm::TP;
    uint64_t il = 0, pl = 0;
    for (int i = 0; i < p.n_outer; i++) {
      uint32_t idx;
      divmod (q, p.o_div[i], q, idx);
      il += idx * p.o_il[i];
      pl += idx * p.o_pl[i];
    }
    il_off = (il + p0 * C) * ES;
    pl_off = (pl + p0) * ES;
    const uint64_t left = p.P - p0;
    npix = (uint32_t) (left < (uint64_t) Gm::TP ? left : (uint64_t) Gm::TP);
  };

  auto issue_load = [&] (uint32_t tile, int s) {
    uint64_t il_off, pl_off;
    uint32_t npix;
    locate (tile, il_off, pl_off, npix);
    mbar_expect_tx (&full[s], npix * C * ES);
    if (TO_PLANAR) {
      bulk_g2s (in_buf + (size_t) s * Gm::TILE_BYTES, in + il_off, npix * C * ES, &full[s]);
    } else {
#pragma unroll
      for (int c = 0; c < C; c++)
        bulk_g2s (in_buf + (size_t) s * Gm::TILE_BYTES + (size_t) c * Gm::PLANE_BYTES,
            in + pl_off + (uint64_t) c * p.plane_stride * ES, npix * ES, &full[s]);
    }
  };

  auto issue_store = [&] (uint32_t tile, int s) {
    uint64_t il_off, pl_off;
    uint32_t npix;
    locate (tile, il_off, pl_off, npix);
    if (TO_PLANAR) {
#pragma unroll
      for (int c = 0; c < C; c++)
        bulk_s2g (out + pl_off + (uint64_t) c * p.plane_stride * ES,
            out_buf + (size_t) s * Gm::TILE_BYTES + (size_t) c * Gm::PLANE_BYTES, npix * ES);
    } else {
      bulk_s2g (out + il_off, out_buf + (size_t) s * Gm::TILE_BYTES, npix * C * ES);
    }
    bulk_commit ();
  };

  uint32_t my_tiles = 0;
  if (blockIdx.x < p.n_tiles)
    my_tiles = (p.n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;

  if (t >= PLANE_THREADS) {
    /* ---- copy-engine warp: one thread keeps loads and stores flowing; it never touches the
     * data, so the 256 re-ordering threads never wait for its address arithmetic ---- */
    if (t == PLANE_THREADS) {
      for (uint32_t k = 0; k < (uint32_t) PLANE_STAGES && k < my_tiles; k++)
        issue_load (tile_of (k), (int) k);
      for (uint32_t k = 0; k < my_tiles; k++) {
        const int s = (int) (k % PLANE_STAGES);
        const uint32_t parity = (k / PLANE_STAGES) & 1u;
        mbar_wait (&done[s], parity);   /* everyone has read in_buf[s] and written out_buf[s] */
        issue_store (tile_of (k), s);
        if (k + PLANE_STAGES < my_tiles)
          issue_load (tile_of (k + PLANE_STAGES), s);
        /* stores complete in order: once all but the newest STAGES-1 have been read out of
         * shared memory, the buffer of tile k-STAGES+1 -- the one tile k+1 writes -- is free */
        bulk_wait_read<PLANE_STAGES - 1> ();
        if (k + 1 >= (uint32_t) PLANE_STAGES)
          mbar_arrive (&empty[(k + 1) % PLANE_STAGES]);
      }
      bulk_wait_all ();         /* stores must complete before the CTA's shared memory goes away */
    }
    return;
  }

  for (uint32_t k = 0; k < my_tiles; k++) {
    const int s = (int) (k % PLANE_STAGES);
    const uint32_t parity = (k / PLANE_STAGES) & 1u;
    mbar_wait (&full[s], parity);
    mbar_wait (&empty[s], parity);      /* the store that last read out_buf[s] has drained */

    const unsigned char *src = in_buf + (size_t) s * Gm::TILE_BYTES;
    unsigned char *dst = out_buf + (size_t) s * Gm::TILE_BYTES;
    /* registers, as 32-bit words: il = G pixels x C channels interleaved, pl = C planes x G */
    uint32_t il[4 * C], pl[4 * C];
    if (TO_PLANAR) {
#pragma unroll
      for (int j = 0; j < C; j++) {
        const uint4 v = *reinterpret_cast<const uint4 *> (src + (size_t) t * 16 * C + 16 * j);
        il[4 * j + 0] = v.x; il[4 * j + 1] = v.y; il[4 * j + 2] = v.z; il[4 * j + 3] = v.w;
      }
#pragma unroll
      for (int i = 0; i < 4 * C; i++)
        pl[i] = 0;
#pragma unroll
      for (int c = 0; c < C; c++)
#pragma unroll
        for (int g = 0; g < G; g++)
          put_elem<ES> (pl, c * G + g, get_elem<ES> (il, g * C + c));
#pragma unroll
      for (int c = 0; c < C; c++)
        *reinterpret_cast<uint4 *> (dst + (size_t) c * Gm::PLANE_BYTES + (size_t) t * 16) =
            make_uint4 (pl[4 * c + 0], pl[4 * c + 1], pl[4 * c + 2], pl[4 * c + 3]);
    } else {
#pragma unroll
      for (int c = 0; c < C; c++) {
        const uint4 v = *reinterpret_cast<const uint4 *> (src + (size_t) c * Gm::PLANE_BYTES + (size_t) t * 16);
        pl[4 * c + 0] = v.x; pl[4 * c + 1] = v.y; pl[4 * c + 2] = v.z; pl[4 * c + 3] = v.w;
      }
#pragma unroll
      for (int i = 0; i < 4 * C; i++)
        il[i] = 0;
#pragma unroll
      for (int c = 0; c < C; c++)
#pragma unroll
        for (int g = 0; g < G; g++)
          put_elem<ES> (il, g * C + c, get_elem<ES> (pl, c * G + g));
#pragma unroll
      for (int j = 0; j < C; j++)
        *reinterpret_cast<uint4 *> (dst + (size_t) t * 16 * C + 16 * j) =
            make_uint4 (il[4 * j + 0], il[4 * j + 1], il[4 * j + 2], il[4 * j + 3]);
    }
    fence_async_smem ();        /* generic-proxy writes -> visible to the bulk store */
    mbar_arrive (&done[s]);
  }
}

template <int C, int ES, bool TO_PLANAR> static int
launch_plane_t (const PlaneDev &p, const void *in, void *out, const LaunchCfg &cfg,
    cudaStream_t stream)
{
  typedef PlaneGeom<C, ES> Gm;
  /* per device: the attribute and the occupancy answer belong to a device's context */
  static int resident_of[64];
  int dev_id = 0;
  cudaGetDevice (&dev_id);
  int &resident = resident_of[(dev_id >= 0 && dev_id < 64) ? dev_id : 0];
  if (resident == 0 || dev_id >= 64) {
    cudaError_t e = cudaFuncSetAttribute (plane_kernel<C, ES, TO_PLANAR>,
        cudaFuncAttributeMaxDynamicSharedMemorySize, (int) Gm::SMEM);
    if (e != cudaSuccess)
      return (int) e;
    resident = resident_blocks_per_sm (plane_kernel<C, ES, TO_PLANAR>, PLANE_BLOCK, Gm::SMEM);
  }
  uint64_t grid = (uint64_t) cfg.sm_count * (cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : resident);
  if (grid > p.n_tiles)
    grid = p.n_tiles;
  if (grid == 0)
    return 0;
  plane_kernel<C, ES, TO_PLANAR><<<(unsigned int) grid, PLANE_BLOCK, Gm::SMEM, stream>>> (
      (const unsigned char *) in, (unsigned char *) out, p);
  return (int) cudaGetLastError ();
}

template <int C, int ES> static int
launch_plane_dir (bool to_planar, const PlaneDev &p, const void *in, void *out,
    const LaunchCfg &cfg, cudaStream_t stream)
{
  return to_planar ? launch_plane_t<C, ES, true> (p, in, out, cfg, stream)
      : launch_plane_t<C, ES, false> (p, in, out, cfg, stream);
}

template <int C> static int
launch_plane_es (int es, bool to_planar, const PlaneDev &p, const void *in, void *out,
    const LaunchCfg &cfg, cudaStream_t stream)
{
  switch (es) {
    case 1: return launch_plane_dir<C, 1> (to_planar, p, in, out, cfg, stream);
    case 2: return launch_plane_dir<C, 2> (to_planar, p, in, out, cfg, stream);
    case 4: return launch_plane_dir<C, 4> (to_planar, p, in, out, cfg, stream);
    default: return -1;
  }
}

/* Returns -1 when the problem does not qualify (caller falls back), else the launch status.
 *   channels, es      : 2..4 channels of 1/2/4-byte elements
 *   pixels            : extent of the plane dimension
 *   plane_stride      : elements between planes on the planar side
 *   outer dims        : size, stride on the interleaved side, stride on the planar side */
int launch_plane (bool to_planar, int channels, int es, uint64_t pixels, uint64_t plane_stride,
    int n_outer, const uint32_t *o_size, const uint64_t *o_il, const uint64_t *o_pl,
    const void *in, void *out, const LaunchCfg &cfg, cudaStream_t stream)
{
  if (channels < 2 || channels > 4 || (es != 1 && es != 2 && es != 4) || n_outer > PLANE_MAX_OUTER)
    return -1;
  if (((uintptr_t) in % 16) || ((uintptr_t) out % 16))
    return -1;
  if ((pixels * es) % 16 || (plane_stride * es) % 16)
    return -1;
  PlaneDev p;
  memset (&p, 0, sizeof (p));
  p.P = pixels;
  p.plane_stride = plane_stride;
  const uint64_t tp = (uint64_t) PLANE_THREADS * (16 / es);
  const uint64_t tiles_per_plane = (pixels + tp - 1) / tp;
  uint64_t n_tiles = tiles_per_plane;
  p.n_outer = n_outer;
  for (int i = 0; i < n_outer; i++) {
    if ((o_il[i] * es) % 16 || (o_pl[i] * es) % 16)
      return -1;
    p.o_size[i] = o_size[i];
    p.o_div[i] = make_fastdiv (o_size[i]);
    p.o_il[i] = o_il[i];
    p.o_pl[i] = o_pl[i];
    n_tiles *= o_size[i];
  }
  if (n_tiles >= (1ull << 31) || tiles_per_plane >= (1ull << 31))
    return -1;
  p.tiles_per_plane = make_fastdiv ((uint32_t) tiles_per_plane);
  p.n_tiles = (uint32_t) n_tiles;
  /* tiny problems are not worth a pipeline */
  if (p.n_tiles * (uint64_t) PLANE_THREADS * 16 * channels < 64 * 1024)
    return -1;
  switch (channels) {
    case 2: return launch_plane_es<2> (es, to_planar, p, in, out, cfg, stream);
    case 3: return launch_plane_es<3> (es, to_planar, p, in, out, cfg, stream);
    case 4: return launch_plane_es<4> (es, to_planar, p, in, out, cfg, stream);
    default: return -1;
  }
}

}  // namespace nnsb
