/*
 * k_perm_vec.cu -- K3/K4 vector path: rank-N permutation copy with 16-byte global accesses on
 * both sides, for every element size (an innermost block that is common to both sides --
 * RGBA bytes, a float4 -- is folded into the element first; 1- and 2-byte elements gather
 * 16 / 8 of them per output vector).
 *
 * Same job as permute_kernel (k_perm.cu): the per-element nns_memcpy nest of
 * gst_tensor_transform_transpose (gsttensor_transform.c:1417-1500) and the block memcpy nest
 * of gst_tensor_transform_dimchg (:1155-1178), one N-dimensional box per tile.  What differs:
 *
 *   - The box is chosen so that 16-byte vectors never straddle a run on either side, so a
 *     thread loads with one cp.async.cg of 16 bytes (LDGSTS.128, L1 bypassed) and stores
 *     with one 16-byte st.global; between the two it gathers the VE = 16/es elements of an
 *     output vector from shared memory with VE scalar LDS.
 *   - Everything that depends only on the box -- per-vector global byte offsets, the swizzled
 *     shared-memory slot of every element, packed in-box coordinates for the edge test -- is
 *     worked out on the HOST when the plan is created and kept in a small device table
 *     (<= 36 KiB).  A block reads its 4 entries per thread once; the tile loop is nothing
 *     but pointer + offset.  The host also verifies the table by brute force (every vector
 *     contiguous and 16-byte aligned on both sides, never cut by an edge) -- a box that fails
 *     is simply not eligible and the scalar kernel is used.
 *   - The shared-memory layout is a 16-byte-chunk XOR swizzle chunk ^ ((chunk >> s) & 7); the
 *     host simulates the bank conflicts of the gather for each candidate s and keeps the best.
 *   - The tile loop is unrolled over the PV_STAGES ring slots, so every shared-memory access
 *     is [register + immediate].
 *
 * Algorithmic bytes: 2 * es per element.
 */
#include <cstdlib>
#include <cstring>
#include <vector>

#include "nnsb_device.cuh"
#include "nnsb_kernels.h"
#include "k_perm_common.cuh"

namespace nnsb {

#ifndef NNSB_PV_THREADS
#define NNSB_PV_THREADS 256
#define NNSB_PV_MIN_BLOCKS 4
#endif
constexpr int PV_THREADS = NNSB_PV_THREADS;
constexpr int PV_VPT = 1024 / PV_THREADS;                 /* vectors per thread per tile */
constexpr int PV_NVT = PV_THREADS * PV_VPT;               /* vector slots per stage: 1024 */
constexpr int PV_STAGE_BYTES = PV_NVT * 16;               /* 16 KiB */
#ifndef NNSB_PV_STAGES
#define NNSB_PV_STAGES 3
#endif
constexpr int PV_STAGES = NNSB_PV_STAGES;
constexpr int PV_DIMS = NNSB_PERM_MAX_RANK + 1;           /* + batch */
constexpr int PV_BOXED = 4;                               /* dims with a box extent > 1 */
constexpr uint32_t PV_INVALID = 0xffffffffu;

/* table layout (bytes) */
constexpr size_t PV_T_LD_OFF = 0;                         /* u32 [NVT] */
constexpr size_t PV_T_LD_SLOT = PV_T_LD_OFF + 4 * PV_NVT; /* u32 [NVT] */
constexpr size_t PV_T_ST_OFF = PV_T_LD_SLOT + 4 * PV_NVT; /* u32 [NVT] */
constexpr size_t PV_T_LD_COORD = PV_T_ST_OFF + 4 * PV_NVT;    /* u64 [NVT] */
constexpr size_t PV_T_ST_COORD = PV_T_LD_COORD + 8 * PV_NVT;  /* u64 [NVT] */
constexpr size_t PV_T_ST_SLOT = PV_T_ST_COORD + 8 * PV_NVT;   /* u16 [NVT][VE] */

struct PermVecDev
{
  int n;                                /* dims incl. batch, output order */
  FastDiv ntile[PV_DIMS];               /* boxes per dim */
  uint32_t tile[PV_DIMS];
  uint32_t size[PV_DIMS];
  uint64_t in_stride_b[PV_DIMS];        /* bytes */
  uint64_t out_stride_b[PV_DIMS];
  int box_slot[PV_DIMS];
  uint64_t limits_init;
  uint32_t n_vec;
  uint32_t n_tiles;
};

/* one vector global -> shared: 16 bytes bypass L1 (.cg), 8 and 4 bytes exist only as .ca */
template <int VB> __device__ __forceinline__ void
cp_async_vec (uint32_t smem_dst, const void *gmem_src)
{
  if constexpr (VB == 16)
    asm volatile ("cp.async.cg.shared.global [%0], [%1], 16;" :: "r" (smem_dst), "l" (gmem_src) : "memory");
  else if constexpr (VB == 8)
    asm volatile ("cp.async.ca.shared.global [%0], [%1], 8;" :: "r" (smem_dst), "l" (gmem_src) : "memory");
  else
    asm volatile ("cp.async.ca.shared.global [%0], [%1], 4;" :: "r" (smem_dst), "l" (gmem_src) : "memory");
}

template <int VB> __device__ __forceinline__ void
cp_async_vec_if (bool on, uint32_t smem_dst, const void *gmem_src)
{
  if constexpr (VB == 16) {
    asm volatile ("{\n"
        ".reg .pred p;\n"
        "setp.ne.u32 p, %2, 0;\n"
        "@p cp.async.cg.shared.global [%0], [%1], 16;\n"
        "}\n" :: "r" (smem_dst), "l" (gmem_src), "r" ((uint32_t) on) : "memory");
  } else {
    if (on)
      cp_async_vec<VB> (smem_dst, gmem_src);
  }
}

template <int VB> __device__ __forceinline__ void
st_stream_vec_if (bool on, void *dst, const uint32_t *w)
{
  if constexpr (VB == 16) {
    asm volatile ("{\n"
        ".reg .pred p;\n"
        "setp.ne.u32 p, %5, 0;\n"
        "@p st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};\n"
        "}\n" :: "l" (dst), "r" (w[0]), "r" (w[1]), "r" (w[2]), "r" (w[3]), "r" ((uint32_t) on)
        : "memory");
  } else {
    if (on)
      st_stream<VB> (dst, w);
  }
}

/* all four 16-bit fields of c below the matching field of lim */
__device__ __forceinline__ bool
inside4 (uint64_t c, uint64_t lim)
{
  bool ok = true;
#pragma unroll
  for (int b = 0; b < PV_BOXED; b++)
    ok = ok && (((uint32_t) (c >> (16 * b)) & 0xffffu) < ((uint32_t) (lim >> (16 * b)) & 0xffffu));
  return ok;
}

/* slots per store vector kept in registers: one per element for 4/8/16-byte elements, two
 * 16-bit slots per register for 1- and 2-byte elements (16 / 8 elements per vector) */
template <int ES, int VB> struct PvSlots
{
  static constexpr int VE = VB / ES;
  static constexpr int N = ES < 4 ? (VE + 1) / 2 : VE;
};

/* the VB / ES elements of one output vector out of shared memory */
template <int ES, int VB> __device__ __forceinline__ void
gather (uint32_t stage_addr, const uint32_t *slot, uint32_t *w)
{
  constexpr int VE = VB / ES;
  if constexpr (ES == 1) {
#pragma unroll
    for (int q = 0; q < VB / 4; q++)
      w[q] = 0;
#pragma unroll
    for (int e = 0; e < VE; e++) {
      const uint32_t sl = (slot[e >> 1] >> (16 * (e & 1))) & 0xffffu;
      uint32_t v;
      asm volatile ("ld.shared.u8 %0, [%1];" : "=r" (v) : "r" (stage_addr + sl));
      w[e >> 2] |= v << (8 * (e & 3));
    }
  } else if constexpr (ES == 2) {
#pragma unroll
    for (int q = 0; q < VB / 4; q++)
      w[q] = 0;
#pragma unroll
    for (int e = 0; e < VE; e++) {
      const uint32_t sl = (slot[e >> 1] >> (16 * (e & 1))) & 0xffffu;
      uint32_t v;
      asm volatile ("ld.shared.u16 %0, [%1];" : "=r" (v) : "r" (stage_addr + sl));
      w[e >> 1] |= v << (16 * (e & 1));
    }
  } else if constexpr (ES == 4) {
#pragma unroll
    for (int e = 0; e < VE; e++)
      asm volatile ("ld.shared.u32 %0, [%1];" : "=r" (w[e]) : "r" (stage_addr + slot[e]));
  } else if constexpr (ES == 8) {
#pragma unroll
    for (int e = 0; e < VE; e++)
      asm volatile ("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r" (w[2 * e]), "=r" (w[2 * e + 1])
          : "r" (stage_addr + slot[e]));
  } else {
    asm volatile ("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r" (w[0]), "=r" (w[1]),
        "=r" (w[2]), "=r" (w[3]) : "r" (stage_addr + slot[0]));
  }
}

/* MICRO: the VE store vectors of a thread form VE x VE register micro-transposes: VE LDS.128
 * (whole shared-memory chunks; entry [j] of st_slot holds chunk j % VE of block j / VE) feed VE
 * output vectors, instead of VE scalar LDS per vector. */
template <int ES, int VB, bool MICRO>
__global__ void __launch_bounds__ (PV_THREADS, ES < 4 ? 2 : NNSB_PV_MIN_BLOCKS)
permute_vec_kernel (const unsigned char *__restrict__ in, unsigned char *__restrict__ out,
    const unsigned char *__restrict__ table, const __grid_constant__ PermVecDev p)
{
  constexpr int VE = VB / ES;
  constexpr uint32_t STAGE_BYTES = PV_NVT * VB;
  extern __shared__ __align__ (128) unsigned char ring[];       /* [PV_STAGES][PV_STAGE_BYTES] */
  const uint32_t ring_addr = (uint32_t) __cvta_generic_to_shared (ring);

  /* this thread's share of the box, from the host-built table */
  uint32_t ld_off[PV_VPT], ld_slot[PV_VPT], st_off[PV_VPT], st_slot[PV_VPT][PvSlots<ES, VB>::N];
  {
    const uint32_t *t_ld_off = reinterpret_cast<const uint32_t *> (table + PV_T_LD_OFF);
    const uint32_t *t_ld_slot = reinterpret_cast<const uint32_t *> (table + PV_T_LD_SLOT);
    const uint32_t *t_st_off = reinterpret_cast<const uint32_t *> (table + PV_T_ST_OFF);
    const uint16_t *t_st_slot = reinterpret_cast<const uint16_t *> (table + PV_T_ST_SLOT);
#pragma unroll
    for (int j = 0; j < PV_VPT; j++) {
      const uint32_t i = threadIdx.x + j * PV_THREADS;
      ld_off[j] = t_ld_off[i];
      ld_slot[j] = ring_addr + t_ld_slot[i];
      st_off[j] = t_st_off[i];
      if constexpr (ES < 4) {
        /* packed pairs, straight from the table */
        const uint32_t *packed = reinterpret_cast<const uint32_t *> (t_st_slot + (size_t) i * VE);
#pragma unroll
        for (int q = 0; q < PvSlots<ES, VB>::N; q++)
          st_slot[j][q] = packed[q];
      } else if constexpr (VE == 4) {
        const uint2 s = *reinterpret_cast<const uint2 *> (t_st_slot + (size_t) i * 4);
        st_slot[j][0] = s.x & 0xffffu;
        st_slot[j][1] = s.x >> 16;
        st_slot[j][2] = s.y & 0xffffu;
        st_slot[j][3] = s.y >> 16;
      } else if constexpr (VE == 2) {
        const uint32_t s = *reinterpret_cast<const uint32_t *> (t_st_slot + (size_t) i * 2);
        st_slot[j][0] = s & 0xffffu;
        st_slot[j][1] = s >> 16;
      } else {
        st_slot[j][0] = t_st_slot[i];
      }
    }
  }
  const uint64_t *t_ld_coord = reinterpret_cast<const uint64_t *> (table + PV_T_LD_COORD);
  const uint64_t *t_st_coord = reinterpret_cast<const uint64_t *> (table + PV_T_ST_COORD);
  const bool whole = (p.n_vec == (uint32_t) PV_NVT);

  uint32_t my_tiles = 0;
  if (blockIdx.x < p.n_tiles)
    my_tiles = (p.n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;

  /* one thread per tile works out where the box sits */
  struct TileInfo
  {
    const unsigned char *src;
    unsigned char *dst;
    uint64_t limits;            /* packed min (box extent, size - origin) of the boxed dims */
    int full;
  };
  constexpr int INFO = PV_STAGES + 1;
  __shared__ TileInfo info[INFO];
  auto locate = [&] (uint32_t k) {
    TileInfo &ti = info[k % INFO];
    uint32_t q = blockIdx.x + k * gridDim.x, td;
    bool full = true;
    uint64_t ib = 0, ob = 0, lim = p.limits_init;
    for (int d = 0; d < p.n; d++) {
      divmod (q, p.ntile[d], q, td);
      const uint32_t o = td * p.tile[d];
      ib += (uint64_t) o * p.in_stride_b[d];
      ob += (uint64_t) o * p.out_stride_b[d];
      const uint32_t left = p.size[d] - o;
      const uint32_t l = left < p.tile[d] ? left : p.tile[d];
      full = full && (left >= p.tile[d]);
      if (p.box_slot[d] >= 0)
        lim |= (uint64_t) l << (16 * p.box_slot[d]);
    }
    ti.src = in + ib;
    ti.dst = out + ob;
    ti.limits = lim;
    ti.full = full ? 1 : 0;
  };

  /* bit j: vector j of this thread exists in the box (always, when the box fills the stage) */
  uint32_t present = 0, present_st = 0;     /* (the store side differs under MICRO) */
#pragma unroll
  for (int j = 0; j < PV_VPT; j++) {
    present |= (ld_off[j] != PV_INVALID ? 1u : 0u) << j;
    present_st |= (st_off[j] != PV_INVALID ? 1u : 0u) << j;
  }

  /* stage_off is a compile-time ring slot: shared-memory addresses are register + immediate.
   * Vectors that are absent, or outside the tensor in an edge box, are predicated off (their
   * offsets / slots are still valid numbers, so the address arithmetic needs no branch). */
  auto issue = [&] (uint32_t k, uint32_t stage_off) {
    if (k < my_tiles) {
      const TileInfo &ti = info[k % INFO];
      const unsigned char *src = ti.src;
      if (ti.full && whole) {
#pragma unroll
        for (int j = 0; j < PV_VPT; j++)
          cp_async_vec<VB> (ld_slot[j] + stage_off, src + ld_off[j]);
      } else if (ti.full) {
#pragma unroll
        for (int j = 0; j < PV_VPT; j++)
          cp_async_vec_if<VB> ((present >> j) & 1u, ld_slot[j] + stage_off, src + ld_off[j]);
      } else {
        const uint64_t lim = ti.limits;
#pragma unroll
        for (int j = 0; j < PV_VPT; j++)
          cp_async_vec_if<VB> (((present >> j) & 1u)
              && inside4 (t_ld_coord[threadIdx.x + j * PV_THREADS], lim),
              ld_slot[j] + stage_off, src + ld_off[j]);
      }
    }
    asm volatile ("cp.async.commit_group;" ::: "memory");     /* possibly empty: keeps the count */
  };
  auto drain = [&] (uint32_t k, uint32_t stage_off) {
    const TileInfo &ti = info[k % INFO];
    unsigned char *dst = ti.dst;
    const uint32_t stage_addr = ring_addr + stage_off;
    uint32_t w[PV_VPT][4];
    if constexpr (MICRO) {
#pragma unroll
      for (int b = 0; b < PV_VPT / VE; b++) {
        uint32_t r[VE][4];
#pragma unroll
        for (int e = 0; e < VE; e++)
          asm volatile ("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r" (r[e][0]), "=r" (r[e][1]),
              "=r" (r[e][2]), "=r" (r[e][3]) : "r" (stage_addr + st_slot[b * VE + e][0]));
#pragma unroll
        for (int jj = 0; jj < VE; jj++)
#pragma unroll
          for (int e = 0; e < VE; e++) {
            if constexpr (ES == 4) {
              w[b * VE + jj][e] = r[e][jj];
            } else {
              w[b * VE + jj][2 * e] = r[e][2 * jj];
              w[b * VE + jj][2 * e + 1] = r[e][2 * jj + 1];
            }
          }
      }
    } else {
#pragma unroll
      for (int j = 0; j < PV_VPT; j++)
        gather<ES, VB> (stage_addr, st_slot[j], w[j]);
    }
    if (ti.full && whole) {
#pragma unroll
      for (int j = 0; j < PV_VPT; j++)
        st_stream<VB> (dst + st_off[j], w[j]);
    } else if (ti.full) {
#pragma unroll
      for (int j = 0; j < PV_VPT; j++)
        st_stream_vec_if<VB> ((present_st >> j) & 1u, dst + st_off[j], w[j]);
    } else {
      const uint64_t lim = ti.limits;
#pragma unroll
      for (int j = 0; j < PV_VPT; j++)
        st_stream_vec_if<VB> (((present_st >> j) & 1u)
            && inside4 (t_st_coord[threadIdx.x + j * PV_THREADS], lim), dst + st_off[j], w[j]);
    }
  };
  /* One barrier per tile: past it, every thread's part of tile k has landed AND every thread
   * has finished draining tile k-1, whose ring slot the next issue overwrites. */
  auto step = [&] (uint32_t k, uint32_t stage_off, uint32_t refill_off) {
    asm volatile ("cp.async.wait_group %0;" :: "n" (PV_STAGES - 2) : "memory");
    __syncthreads ();
    if (threadIdx.x == 0 && k + PV_STAGES < my_tiles)
      locate (k + PV_STAGES);   /* slot (k - 1) % (S + 1): last read while draining tile k-1 */
    issue (k + PV_STAGES - 1, refill_off);
    drain (k, stage_off);
  };

  if (threadIdx.x == 0)
    for (uint32_t k = 0; k < (uint32_t) PV_STAGES && k < my_tiles; k++)
      locate (k);
  __syncthreads ();
  /* prologue: PV_STAGES - 1 tiles in flight; the loop is unrolled over the ring slots so that
   * every shared-memory offset is a compile-time constant */
#pragma unroll
  for (int s = 0; s < PV_STAGES - 1; s++)
    issue ((uint32_t) s, (uint32_t) s * STAGE_BYTES);
  for (uint32_t k = 0; k < my_tiles; k += PV_STAGES) {
#pragma unroll
    for (int s = 0; s < PV_STAGES; s++)
      if (k + s < my_tiles)
        step (k + s, (uint32_t) s * STAGE_BYTES,
            (uint32_t) ((s + PV_STAGES - 1) % PV_STAGES) * STAGE_BYTES);
  }
}

/* ------------------------------------------------------------------------------------- */
/* host: box selection, table, verification                                               */
/* ------------------------------------------------------------------------------------- */
namespace {

uint64_t
gcd_u64 (uint64_t a, uint64_t b)
{
  while (b) {
    const uint64_t t = a % b;
    a = b;
    b = t;
  }
  return a;
}

uint64_t
box_volume (const uint64_t *T, int n)
{
  uint64_t v = 1;
  for (int i = 0; i < n; i++)
    v *= T[i];
  return v;
}

/* walk the dims in `order` until the run holds `run` elements; extents are rounded up so the
 * running product stays a multiple of the vector length where the dim allows it */
void
ensure_run (const CDim *dims, int n, const int *order, uint64_t run, uint64_t ve, uint64_t *T)
{
  uint64_t prod = 1;
  for (int k = 0; k < n && prod < run; k++) {
    const int d = order[k];
    uint64_t want = (run + prod - 1) / prod;
    if (want >= dims[d].size) {
      want = dims[d].size;
    } else {
      const uint64_t step = ve / gcd_u64 (ve, prod);
      want = (want + step - 1) / step * step;
      if (want > dims[d].size)
        want = dims[d].size;
    }
    if (T[d] < want)
      T[d] = want;
    prod *= T[d];
  }
}

/* shared-memory swizzle of 16-byte chunk q: XOR of its low three bits with bits picked higher
 * up (s1, s2 < 0 = unused).  With s < 3 the source overlaps the target bits; the builder only
 * keeps candidates that are permutations of the stage's chunks. */
struct Swizzle
{
  int s1, s2;
  uint32_t operator () (uint32_t q) const
  {
    uint32_t x = 0;
    if (s1 >= 0)
      x ^= q >> s1;
    if (s2 >= 0)
      x ^= q >> s2;
    return q ^ (x & 7u);
  }
  bool is_permutation () const
  {
    bool seen[PV_NVT];
    memset (seen, 0, sizeof (seen));
    for (uint32_t q = 0; q < (uint32_t) PV_NVT; q++) {
      const uint32_t r = (*this) (q);
      if (r >= (uint32_t) PV_NVT || seen[r])
        return false;
      seen[r] = true;
    }
    return true;
  }
};

/* wavefronts of the gather (one LDS per element of every output vector) under a swizzle:
 * 32-bit accesses are served a warp at a time over 32 banks, 64/128-bit ones a half/quarter
 * warp at a time over groups of 2/4 banks */
uint64_t
gather_cost (const uint32_t *lin, uint32_t n_vec, int ve, int es, const Swizzle &sw)
{
  uint64_t cost = 0;
  /* 1/2-byte accesses: a warp at a time, 32 banks of 4 bytes, distinct words per bank */
  const int lanes = es < 4 ? 32 : 128 / es, banks = lanes;
  const int per_word = es < 4 ? 4 / es : 1;
  for (uint32_t base = 0; base < n_vec; base += lanes) {
    for (int e = 0; e < ve; e++) {
      uint32_t words[32][32];
      int count[32];
      memset (count, 0, sizeof (count));
      int worst = 0;
      for (int l = 0; l < lanes && base + l < n_vec; l++) {
        const uint32_t li = lin[(size_t) (base + l) * ve + e];
        const uint32_t w = (sw (li / ve) * ve + li % ve) / per_word;  /* bank-unit word index */
        const int b = (int) (w % banks);
        bool dup = false;
        for (int k = 0; k < count[b]; k++)
          dup = dup || words[b][k] == w;
        if (!dup)
          words[b][count[b]++] = w;
        if (count[b] > worst)
          worst = count[b];
      }
      cost += worst;
    }
  }
  return cost;
}

/* One permutation problem in canonical form, and the checks / table for a candidate box */
struct Builder
{
  CDim dims[PV_DIMS];
  int n, es, vb;                /* vb: vector bytes (16, 8 or 4) */
  uint64_t VE, TILE_MAX;
  int in_order[PV_DIMS], out_order[PV_DIMS];

  /* Static conditions + brute-force verification of box T.  With tab != null also fills the
   * offsets / coordinates of the table and st_lin (load-order index of every output element). */
  bool check (const uint64_t *T, int *box_slot, unsigned char *tab, uint32_t *st_lin) const
  {
    const uint64_t volume = box_volume (T, n);
    if (volume % VE || volume > TILE_MAX)
      return false;
    const uint32_t n_vec = (uint32_t) (volume / VE);
    uint64_t span_in = 0, span_out = 0, tiles = 1;
    int n_boxed = 0, slot[PV_DIMS];
    for (int i = 0; i < n; i++) {
      /* box origins vector aligned on both sides */
      if (T[i] < dims[i].size && ((T[i] * dims[i].in * es) % vb || (T[i] * dims[i].out * es) % vb))
        return false;
      span_in += (T[i] - 1) * dims[i].in * es;
      span_out += (T[i] - 1) * dims[i].out * es;
      tiles *= (dims[i].size + T[i] - 1) / T[i];
      slot[i] = -1;
      if (T[i] > 1) {
        if (n_boxed == PV_BOXED || T[i] > 0xffff)
          return false;
        slot[i] = n_boxed++;
      }
    }
    if (span_in >= (1ull << 32) - 16 || span_out >= (1ull << 32) - 16 || tiles >= (1ull << 31))
      return false;
    uint32_t lstride[PV_DIMS];
    uint32_t ls = 1;
    for (int k = 0; k < n; k++) {
      lstride[in_order[k]] = ls;
      ls *= (uint32_t) T[in_order[k]];
    }
    uint64_t rem[PV_DIMS];              /* extent of the last, partial box (0 = none) */
    for (int i = 0; i < n; i++)
      rem[i] = dims[i].size % T[i];

    uint32_t *t_ld_off = tab ? (uint32_t *) (tab + PV_T_LD_OFF) : nullptr;
    uint32_t *t_st_off = tab ? (uint32_t *) (tab + PV_T_ST_OFF) : nullptr;
    uint64_t *t_ld_coord = tab ? (uint64_t *) (tab + PV_T_LD_COORD) : nullptr;
    uint64_t *t_st_coord = tab ? (uint64_t *) (tab + PV_T_ST_COORD) : nullptr;
    for (int side = 0; side < 2; side++) {
      const int *order = side == 0 ? in_order : out_order;
      for (uint32_t v = 0; v < n_vec; v++) {
        uint64_t off0 = 0, coord0 = 0;
        bool cut0[PV_DIMS];
        for (uint64_t e = 0; e < VE; e++) {
          uint64_t q = (uint64_t) v * VE + e, off = 0, coord = 0, lin = 0;
          bool cut[PV_DIMS];
          for (int k = 0; k < n; k++) {
            const int d = order[k];
            const uint64_t c = q % T[d];
            q /= T[d];
            off += c * (side == 0 ? dims[d].in : dims[d].out);
            lin += c * lstride[d];
            if (slot[d] >= 0)
              coord |= c << (16 * slot[d]);
            cut[d] = rem[d] != 0 && c >= rem[d];
          }
          if (e == 0) {
            off0 = off;
            coord0 = coord;
            for (int d = 0; d < n; d++)
              cut0[d] = cut[d];
          } else {
            if (off != off0 + e)
              return false;             /* not contiguous in global memory */
            for (int d = 0; d < n; d++)
              if (cut[d] != cut0[d])
                return false;           /* an edge would cut through the vector */
          }
          if (side == 1 && st_lin)
            st_lin[(size_t) v * VE + e] = (uint32_t) lin;
        }
        if ((off0 * es) % vb)
          return false;
        if (tab) {
          if (side == 0) {
            t_ld_off[v] = (uint32_t) (off0 * es);
            t_ld_coord[v] = coord0;
          } else {
            t_st_off[v] = (uint32_t) (off0 * es);
            t_st_coord[v] = coord0;
          }
        }
      }
    }
    if (box_slot)
      for (int i = 0; i < n; i++)
        box_slot[i] = slot[i];
    return true;
  }
  bool ok (const uint64_t *T) const
  {
    return check (T, nullptr, nullptr, nullptr);
  }
};

}  // namespace

static int perm_vec_build_vb (const PermDesc &d0, size_t frame_elems, int vb, PermVecPlan *pl,
    void **host_table);

/* widest vector that the shape allows: 16 bytes, else 8, else 4 (rows of 3240 bytes -- 1080 RGB
 * pixels -- are 8-byte but not 16-byte multiples) */
int
perm_vec_build (const PermDesc &d0, size_t frame_elems, PermVecPlan *pl, void **host_table)
{
  memset (pl, 0, sizeof (*pl));
  *host_table = nullptr;
  if (getenv ("NNSB_PERM_NO_VEC"))
    return -1;
  const char *env_vb = getenv ("NNSB_PERM_VEC_BYTES");
  for (int vb = 16; vb >= 4; vb /= 2) {
    if (env_vb && atoi (env_vb) != vb)
      continue;
    if (perm_vec_build_vb (d0, frame_elems, vb, pl, host_table) == 0)
      return 0;
  }
  memset (pl, 0, sizeof (*pl));
  return -1;
}

static int
perm_vec_build_vb (const PermDesc &d0, size_t frame_elems, int vb, PermVecPlan *pl,
    void **host_table)
{
  memset (pl, 0, sizeof (*pl));
  *host_table = nullptr;
  if (((uint64_t) frame_elems * d0.es) % vb)
    return -1;
  /* launches require vector-aligned bases, so an innermost block may fold up to one vector */
  Builder B;
  B.es = d0.es;
  B.vb = vb;
  B.n = canonicalise_perm (d0.size, d0.in_stride, d0.rank, frame_elems, 0, vb, vb, B.dims, &B.es);
  const int n = B.n, es = B.es;
  if (n < 2)
    return -1;                  /* identity: the flat copy */
  if ((es != 1 && es != 2 && es != 4 && es != 8 && es != 16) || es > vb)
    return -1;
  const uint64_t VE = B.VE = vb / es;
  const uint64_t TILE_MAX = B.TILE_MAX = (uint64_t) PV_NVT * vb / es;
  const CDim *dims = B.dims;
  for (int i = 0; i < n; i++)
    if (dims[i].size > 0xffffffffull)
      return -1;
  int *in_order = B.in_order, *out_order = B.out_order;
  for (int i = 0; i < n; i++)
    in_order[i] = out_order[i] = i;
  for (int i = 1; i < n; i++)
    for (int j = i; j > 0 && dims[in_order[j]].in < dims[in_order[j - 1]].in; j--) {
      const int t = in_order[j];
      in_order[j] = in_order[j - 1];
      in_order[j - 1] = t;
    }

  /* 1. the box.  Start from runs of >= RUN bytes on both sides, made of whole 32-byte sectors
   *    where the sizes allow (16-byte multiples otherwise) ... */
  const char *env_run = getenv ("NNSB_PERM_VEC_RUN");
  const uint64_t RUN = (env_run ? (uint64_t) atoi (env_run) : 128) / es;
  uint64_t T[PV_DIMS];
  bool have = false;
  for (uint64_t gran = 2 * VE; gran >= VE && !have; gran /= 2) {
    for (int i = 0; i < n; i++)
      T[i] = 1;
    ensure_run (dims, n, in_order, RUN, gran, T);
    ensure_run (dims, n, out_order, RUN, gran, T);
    for (int guard = 0; box_volume (T, n) > TILE_MAX && guard < 64; guard++) {
      int big = -1;
      for (int i = 0; i < n; i++)
        if (T[i] > gran && (big < 0 || T[i] > T[big]))
          big = i;
      if (big < 0)
        break;
      T[big] = ((T[big] + 1) / 2 + gran - 1) / gran * gran;
    }
    have = B.ok (T);
  }
  if (!have)
    return -1;
  /* ... double the two runs in turn while the box fits 16 KiB (keeps them balanced) ... */
  for (int round = 0; round < 64; round++) {
    bool grew = false;
    for (int side = 0; side < 2; side++) {
      const int *order = side == 0 ? out_order : in_order;
      for (int k = 0; k < n; k++) {
        const int d = order[k];
        if (T[d] < dims[d].size) {
          const uint64_t old = T[d];
          T[d] = old * 2 > dims[d].size ? dims[d].size : old * 2;
          if (box_volume (T, n) <= TILE_MAX && B.ok (T))
            grew = true;
          else
            T[d] = old;
          break;                /* only the first non-full dim of each side */
        }
      }
    }
    if (!grew)
      break;
  }
  /* ... then fill what is left of the stage in smaller power-of-two quanta */
  for (int side = 0; side < 2; side++) {
    const int *order = side == 0 ? out_order : in_order;
    for (int k = 0; k < n; k++) {
      const int d = order[k];
      if (T[d] == dims[d].size)
        continue;
      const uint64_t old = T[d];
      uint64_t most = TILE_MAX / (box_volume (T, n) / old);
      if (most > dims[d].size)
        most = dims[d].size;
      uint64_t quantum = old & (~old + 1);      /* largest power of two dividing the extent */
      bool grown = false;
      for (; quantum >= 1 && !grown; quantum /= 2) {
        const uint64_t nt = most / quantum * quantum;
        if (nt <= old)
          continue;
        T[d] = nt;
        if (B.ok (T))
          grown = true;
        else
          T[d] = old;
      }
      break;
    }
  }
  const uint64_t volume = box_volume (T, n);
  const uint32_t n_vec = (uint32_t) (volume / VE);

  /* 2. the table, verified element by element */
  const size_t table_bytes = PV_T_ST_SLOT + (size_t) PV_NVT * VE * 2;
  unsigned char *tab = (unsigned char *) calloc (1, table_bytes);
  uint32_t *st_lin = (uint32_t *) malloc (sizeof (uint32_t) * volume);
  if (!tab || !st_lin) {
    free (tab);
    free (st_lin);
    return -1;
  }
  uint32_t *t_ld_off = (uint32_t *) (tab + PV_T_LD_OFF);
  uint32_t *t_ld_slot = (uint32_t *) (tab + PV_T_LD_SLOT);
  uint32_t *t_st_off = (uint32_t *) (tab + PV_T_ST_OFF);
  uint16_t *t_st_slot = (uint16_t *) (tab + PV_T_ST_SLOT);
  for (int i = 0; i < PV_NVT; i++)
    t_ld_off[i] = t_st_off[i] = PV_INVALID;
  if (!B.check (T, pl->box_slot, tab, st_lin)) {
    free (tab);
    free (st_lin);
    return -1;
  }
  int n_boxed = 0;
  for (int i = 0; i < n; i++)
    if (pl->box_slot[i] >= 0)
      n_boxed++;
  uint64_t limits_init = 0;
  for (int b = n_boxed; b < PV_BOXED; b++)
    limits_init |= 1ull << (16 * b);
  uint64_t tiles = 1;
  for (int i = 0; i < n; i++)
    tiles *= (dims[i].size + T[i] - 1) / T[i];

  /* 2b. register micro-transposes: possible when the fastest output dim O is not the fastest
   *     input dim, every shared-memory chunk lies at one O coordinate and O-vectors are whole.
   *     A block = one chunk column x VE consecutive O: VE chunks in, VE store vectors out. */
  bool micro = false;
  uint32_t lstride_o = 0;
  std::vector<uint32_t> blocks;         /* representative store vector (element 0 at lin % VE == 0) */
  std::vector<int32_t> vec_of_lin;
  if (vb == 16 && (es == 4 || es == 8) && out_order[0] != in_order[0]
      && !getenv ("NNSB_PERM_NO_MICRO")) {
    uint32_t ls = 1;
    for (int k = 0; k < n && in_order[k] != out_order[0]; k++)
      ls *= (uint32_t) T[in_order[k]];
    lstride_o = ls;
    if (lstride_o % VE == 0 && T[out_order[0]] % VE == 0) {
      vec_of_lin.assign (volume, -1);
      for (uint32_t v = 0; v < n_vec; v++)
        vec_of_lin[st_lin[(size_t) v * VE]] = (int32_t) v;
      for (uint32_t v = 0; v < n_vec; v++)
        if (st_lin[(size_t) v * VE] % VE == 0)
          blocks.push_back (v);
      micro = blocks.size () * VE == n_vec;
      for (size_t mb = 0; mb < blocks.size () && micro; mb++) {
        const uint32_t lin0 = st_lin[(size_t) blocks[mb] * VE];
        for (uint32_t jj = 0; jj < VE && micro; jj++) {
          const int32_t u = lin0 + jj < volume ? vec_of_lin[lin0 + jj] : -1;
          micro = u >= 0;
          for (uint32_t e = 0; e < VE && micro; e++)
            micro = st_lin[(size_t) u * VE + e] == lin0 + jj + e * lstride_o;
        }
      }
    }
  }
  /* chunk read by LDS.128 number e of block mb (micro mode) */
  auto micro_cost = [&] (const Swizzle &sw) {
    uint64_t cost = 0;
    for (size_t base = 0; base < blocks.size (); base += 8)
      for (uint32_t e = 0; e < VE; e++) {
        int count[8] = { 0, 0, 0, 0, 0, 0, 0, 0 }, worst = 0;
        for (size_t l = 0; l < 8 && base + l < blocks.size (); l++) {
          const uint32_t lin0 = st_lin[(size_t) blocks[base + l] * VE];
          const int b = (int) (sw ((lin0 + e * lstride_o) / (uint32_t) VE) % 8);
          if (++count[b] > worst)
            worst = count[b];
        }
        cost += worst;
      }
    return cost;
  };
  auto cost_of = [&] (const Swizzle &sw) {
    return micro ? micro_cost (sw) : gather_cost (st_lin, n_vec, (int) VE, es, sw);
  };

  /* 3. shared-memory swizzle: the candidate with the fewest gather wavefronts */
  Swizzle best_sw = { -1, -1 };
  {
    const char *env_sw = getenv ("NNSB_PERM_VEC_SWIZZLE");
    if (env_sw) {
      best_sw.s1 = atoi (env_sw);          /* "s1" or "s1,s2" */
      if (strchr (env_sw, ','))
        best_sw.s2 = atoi (strchr (env_sw, ',') + 1);
      if (!best_sw.is_permutation ())
        best_sw.s1 = best_sw.s2 = -1;
      pl->gather_wavefronts = cost_of (best_sw);
    } else {
      uint64_t best = cost_of (best_sw);
      /* s1 = 1 is left out: same simulated gather cost, but measured 8 % slower (the
       * cp.async writes seem to dislike chunk pairs being swapped) */
      for (int s1 = 2; s1 <= 9; s1++)
        for (int s2 = -1; s2 <= 9; s2++) {
          if (s2 >= 0 && s2 <= s1)
            continue;
          const Swizzle sw = { s1, s2 };
          if (!sw.is_permutation ())
            continue;
          const uint64_t c = cost_of (sw);
          if (c < best) {
            best = c;
            best_sw = sw;
          }
        }
      pl->gather_wavefronts = best;
    }
  }
  for (uint32_t v = 0; v < n_vec; v++)
    t_ld_slot[v] = best_sw (v) * vb;
  if (micro) {
    /* entry t + j * 256 of the store-side arrays = vector j % VE of block t + (j / VE) * 256 */
    uint64_t *t_st_coord = (uint64_t *) (tab + PV_T_ST_COORD);
    std::vector<uint32_t> off (t_st_off, t_st_off + PV_NVT);
    std::vector<uint64_t> coord (t_st_coord, t_st_coord + PV_NVT);
    for (int i = 0; i < PV_NVT; i++) {
      t_st_off[i] = PV_INVALID;
      t_st_coord[i] = 0;
    }
    for (size_t mb = 0; mb < blocks.size (); mb++) {
      const uint32_t lin0 = st_lin[(size_t) blocks[mb] * VE];
      for (uint32_t jj = 0; jj < VE; jj++) {
        const size_t entry = (mb % PV_THREADS) + ((mb / PV_THREADS) * VE + jj) * PV_THREADS;
        const uint32_t u = (uint32_t) vec_of_lin[lin0 + jj];
        t_st_off[entry] = off[u];
        t_st_coord[entry] = coord[u];
        t_st_slot[entry * VE] = (uint16_t) (best_sw ((lin0 + jj * lstride_o) / (uint32_t) VE) * vb);
      }
    }
  } else {
    for (uint32_t v = 0; v < n_vec; v++)
      for (uint64_t e = 0; e < VE; e++) {
        const uint32_t li = st_lin[(size_t) v * VE + e];
        t_st_slot[(size_t) v * VE + e] = (uint16_t) (best_sw (li / (uint32_t) VE) * vb + (li % VE) * es);
      }
  }
  free (st_lin);

  pl->valid = 1;
  pl->es = es;
  pl->n = n;
  for (int i = 0; i < n; i++) {
    pl->size[i] = (uint32_t) dims[i].size;
    pl->tile[i] = (uint32_t) T[i];
    pl->in_stride[i] = dims[i].in;
    pl->out_stride[i] = dims[i].out;
  }
  pl->limits_init = limits_init;
  pl->n_vec = n_vec;
  pl->tiles_per_frame = tiles;
  pl->table_bytes = table_bytes;
  pl->frame_bytes = (uint64_t) frame_elems * d0.es;
  pl->swizzle_shift = best_sw.s1;
  pl->swizzle_shift2 = best_sw.s2;
  pl->micro = micro ? 1 : 0;
  pl->vb = vb;
  *host_table = tab;
  return 0;
}

template <int ES, int VB, bool MICRO> static int
launch_permute_vec_t (const PermVecDev &p, const void *d_table, const void *in, void *out,
    const LaunchCfg &cfg, cudaStream_t stream)
{
  const size_t smem = (size_t) PV_STAGES * PV_NVT * VB;
  /* per device: the attribute and the occupancy answer belong to a device's context */
  static int resident_of[64];
  int dev_id = 0;
  cudaGetDevice (&dev_id);
  int &resident = resident_of[(dev_id >= 0 && dev_id < 64) ? dev_id : 0];
  if (resident == 0 || dev_id >= 64) {
    cudaError_t e = cudaFuncSetAttribute (permute_vec_kernel<ES, VB, MICRO>,
        cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e != cudaSuccess)
      return (int) e;
    resident = resident_blocks_per_sm (permute_vec_kernel<ES, VB, MICRO>, PV_THREADS, smem);
  }
  uint64_t grid = p.n_tiles;
  /* two CTAs per SM, although four fit: fewer tiles in flight keep the DRAM pages of a tile
   * row open (measured on the 4K float32 transposes: 6.0 / 5.5 / 5.2 TB/s with two, 5.4 / 4.7 /
   * 5.0 with four, 4.7 / 4.7 / 4.3 with one) */
  const int per_sm = cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : (resident < 2 ? resident : 2);
  const uint64_t cap = (uint64_t) cfg.sm_count * per_sm;
  if (grid > cap)
    grid = cap;
  permute_vec_kernel<ES, VB, MICRO><<<(unsigned int) grid, PV_THREADS, smem, stream>>> (
      (const unsigned char *) in, (unsigned char *) out, (const unsigned char *) d_table, p);
  return (int) cudaGetLastError ();
}

int
launch_permute_vec (const PermVecPlan &pl, const void *d_table, const void *in, void *out,
    size_t frame_elems, uint32_t batch, const LaunchCfg &cfg, cudaStream_t stream)
{
  if (!pl.valid || !d_table || batch == 0)
    return -1;
  if ((uintptr_t) in % pl.vb || (uintptr_t) out % pl.vb)
    return -1;
  const uint64_t tiles = pl.tiles_per_frame * batch;
  if (tiles >= (1ull << 31))
    return -1;
  PermVecDev p;
  memset (&p, 0, sizeof (p));
  int n = pl.n;
  for (int i = 0; i < n; i++) {
    p.tile[i] = pl.tile[i];
    p.size[i] = pl.size[i];
    p.ntile[i] = make_fastdiv ((pl.size[i] + pl.tile[i] - 1) / pl.tile[i]);
    p.in_stride_b[i] = pl.in_stride[i] * pl.es;
    p.out_stride_b[i] = pl.out_stride[i] * pl.es;
    p.box_slot[i] = pl.box_slot[i];
  }
  if (batch > 1) {
    p.tile[n] = 1;
    p.size[n] = batch;
    p.ntile[n] = make_fastdiv (batch);
    p.in_stride_b[n] = pl.frame_bytes;
    p.out_stride_b[n] = pl.frame_bytes;
    p.box_slot[n] = -1;
    n++;
  }
  p.n = n;
  p.limits_init = pl.limits_init;
  p.n_vec = pl.n_vec;
  p.n_tiles = (uint32_t) tiles;
#define NNSB_PV(ES, VB, M) launch_permute_vec_t<ES, VB, M> (p, d_table, in, out, cfg, stream)
  if (pl.vb == 16) {
    switch (pl.es) {
      case 1: return NNSB_PV (1, 16, false);
      case 2: return NNSB_PV (2, 16, false);
      case 4: return pl.micro ? NNSB_PV (4, 16, true) : NNSB_PV (4, 16, false);
      case 8: return pl.micro ? NNSB_PV (8, 16, true) : NNSB_PV (8, 16, false);
      case 16: return NNSB_PV (16, 16, false);
      default: return -1;
    }
  } else if (pl.vb == 8) {
    switch (pl.es) {
      case 1: return NNSB_PV (1, 8, false);
      case 2: return NNSB_PV (2, 8, false);
      case 4: return NNSB_PV (4, 8, false);
      case 8: return NNSB_PV (8, 8, false);
      default: return -1;
    }
  } else if (pl.vb == 4) {
    switch (pl.es) {
      case 1: return NNSB_PV (1, 4, false);
      case 2: return NNSB_PV (2, 4, false);
      case 4: return NNSB_PV (4, 4, false);
      default: return -1;
    }
  }
  return -1;
#undef NNSB_PV
}

}  // namespace nnsb
