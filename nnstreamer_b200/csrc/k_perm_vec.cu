/*
 * k_perm_vec.cu -- K3/K4 vector path: rank-N permutation copy with 16-byte global accesses on
 * both sides, for 4- and 8-byte elements.
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

#include "nnsb_device.cuh"
#include "nnsb_kernels.h"
#include "k_perm_common.cuh"

namespace nnsb {

constexpr int PV_THREADS = 256;
constexpr int PV_VPT = 4;                                 /* vectors per thread per tile */
constexpr int PV_NVT = PV_THREADS * PV_VPT;               /* vector slots per stage: 1024 */
constexpr int PV_STAGE_BYTES = PV_NVT * 16;               /* 16 KiB */
constexpr int PV_STAGES = 3;
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

__device__ __forceinline__ void
cp_async_16 (uint32_t smem_dst, const void *gmem_src)
{
  asm volatile ("cp.async.cg.shared.global [%0], [%1], 16;" :: "r" (smem_dst), "l" (gmem_src) : "memory");
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

template <int ES> __device__ __forceinline__ void
gather (uint32_t stage_addr, const uint32_t *slot, uint32_t *w)
{
  if constexpr (ES == 4) {
#pragma unroll
    for (int e = 0; e < 4; e++)
      asm volatile ("ld.shared.u32 %0, [%1];" : "=r" (w[e]) : "r" (stage_addr + slot[e]));
  } else {
#pragma unroll
    for (int e = 0; e < 2; e++)
      asm volatile ("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r" (w[2 * e]), "=r" (w[2 * e + 1])
          : "r" (stage_addr + slot[e]));
  }
}

template <int ES>
__global__ void __launch_bounds__ (PV_THREADS, 4)
permute_vec_kernel (const unsigned char *__restrict__ in, unsigned char *__restrict__ out,
    const unsigned char *__restrict__ table, const __grid_constant__ PermVecDev p)
{
  constexpr int VE = 16 / ES;
  extern __shared__ __align__ (128) unsigned char ring[];       /* [PV_STAGES][PV_STAGE_BYTES] */
  const uint32_t ring_addr = (uint32_t) __cvta_generic_to_shared (ring);

  /* this thread's share of the box, from the host-built table */
  uint32_t ld_off[PV_VPT], ld_slot[PV_VPT], st_off[PV_VPT], st_slot[PV_VPT][VE];
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
      if constexpr (VE == 4) {
        const uint2 s = *reinterpret_cast<const uint2 *> (t_st_slot + (size_t) i * 4);
        st_slot[j][0] = s.x & 0xffffu;
        st_slot[j][1] = s.x >> 16;
        st_slot[j][2] = s.y & 0xffffu;
        st_slot[j][3] = s.y >> 16;
      } else {
        const uint32_t s = *reinterpret_cast<const uint32_t *> (t_st_slot + (size_t) i * 2);
        st_slot[j][0] = s & 0xffffu;
        st_slot[j][1] = s >> 16;
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

  /* S / NEXT are compile-time ring slots: shared-memory addresses become register + immediate */
  auto issue = [&] (uint32_t k, uint32_t stage_off) {
    if (k < my_tiles) {
      const TileInfo &ti = info[k % INFO];
      const unsigned char *src = ti.src;
      if (ti.full && whole) {
#pragma unroll
        for (int j = 0; j < PV_VPT; j++)
          cp_async_16 (ld_slot[j] + stage_off, src + ld_off[j]);
      } else if (ti.full) {
#pragma unroll
        for (int j = 0; j < PV_VPT; j++)
          if (ld_off[j] != PV_INVALID)
            cp_async_16 (ld_slot[j] + stage_off, src + ld_off[j]);
      } else {
        const uint64_t lim = ti.limits;
#pragma unroll
        for (int j = 0; j < PV_VPT; j++)
          if (ld_off[j] != PV_INVALID && inside4 (t_ld_coord[threadIdx.x + j * PV_THREADS], lim))
            cp_async_16 (ld_slot[j] + stage_off, src + ld_off[j]);
      }
    }
    asm volatile ("cp.async.commit_group;" ::: "memory");     /* possibly empty: keeps the count */
  };
  auto drain = [&] (uint32_t k, uint32_t stage_off) {
    const TileInfo &ti = info[k % INFO];
    unsigned char *dst = ti.dst;
    const uint32_t stage_addr = ring_addr + stage_off;
    if (ti.full && whole) {
      uint32_t w[PV_VPT][4];
#pragma unroll
      for (int j = 0; j < PV_VPT; j++)
        gather<ES> (stage_addr, st_slot[j], w[j]);
#pragma unroll
      for (int j = 0; j < PV_VPT; j++)
        st_stream<16> (dst + st_off[j], w[j]);
    } else {
      const bool full = ti.full != 0;
      const uint64_t lim = ti.limits;
#pragma unroll
      for (int j = 0; j < PV_VPT; j++)
        if (st_off[j] != PV_INVALID
            && (full || inside4 (t_st_coord[threadIdx.x + j * PV_THREADS], lim))) {
          uint32_t w[4];
          gather<ES> (stage_addr, st_slot[j], w);
          st_stream<16> (dst + st_off[j], w);
        }
    }
  };
  auto step = [&] (uint32_t k, uint32_t stage_off, uint32_t next_off) {
    issue (k + PV_STAGES - 1, next_off);
    asm volatile ("cp.async.wait_group %0;" :: "n" (PV_STAGES - 1) : "memory");
    __syncthreads ();
    if (threadIdx.x == 0 && k + PV_STAGES < my_tiles)
      locate (k + PV_STAGES);   /* slot (k + S) % (S + 1): not read before the next barrier */
    drain (k, stage_off);
    __syncthreads ();           /* the stage is overwritten by the issue of the next step */
  };

  if (threadIdx.x == 0)
    for (uint32_t k = 0; k < (uint32_t) PV_STAGES && k < my_tiles; k++)
      locate (k);
  __syncthreads ();
  static_assert (PV_STAGES == 3, "the tile loop below is unrolled for three ring slots");
  issue (0, 0);
  issue (1, PV_STAGE_BYTES);
  for (uint32_t k = 0; k < my_tiles; k += 3) {
    step (k, 0, 2 * PV_STAGE_BYTES);
    if (k + 1 < my_tiles)
      step (k + 1, PV_STAGE_BYTES, 0);
    if (k + 2 < my_tiles)
      step (k + 2, 2 * PV_STAGE_BYTES, PV_STAGE_BYTES);
  }
}

/* ------------------------------------------------------------------------------------- */
/* host: box selection, table, verification                                               */
/* ------------------------------------------------------------------------------------- */
namespace {

struct CDim
{
  uint64_t size, in, out;       /* extent, input stride, output stride (elements) */
};

/* drop extent-1 dims, merge dims adjacent on both sides; returns the dim count */
int
canonicalise (const PermDesc &d0, CDim *dims)
{
  int n = 0;
  uint64_t ostride = 1;
  for (int i = 0; i < d0.rank; i++) {
    if (d0.size[i] != 1) {
      dims[n].size = d0.size[i];
      dims[n].in = d0.in_stride[i];
      dims[n].out = ostride;
      n++;
    }
    ostride *= d0.size[i];
  }
  int m = 0;
  for (int i = 0; i < n; i++) {
    if (m > 0 && dims[m - 1].in * dims[m - 1].size == dims[i].in
        && dims[m - 1].out * dims[m - 1].size == dims[i].out)
      dims[m - 1].size *= dims[i].size;
    else
      dims[m++] = dims[i];
  }
  return m;
}

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

/* wavefronts of the gather (one LDS per element of every output vector) for a swizzle */
uint64_t
gather_cost (const uint32_t *lin, uint32_t n_vec, int ve, int es, int shift)
{
  uint64_t cost = 0;
  const int lanes = es == 8 ? 16 : 32;          /* 64-bit accesses go half a warp at a time */
  const int banks = es == 8 ? 16 : 32;          /* ... over 16 bank pairs */
  for (uint32_t base = 0; base < (uint32_t) PV_NVT; base += lanes) {
    for (int e = 0; e < ve; e++) {
      uint32_t words[32];
      int nw = 0;
      for (int l = 0; l < lanes; l++) {
        const uint32_t i = base + l;
        if (i >= n_vec)
          continue;
        const uint32_t li = lin[(size_t) i * ve + e];
        uint32_t q = li / ve;
        if (shift >= 0)
          q ^= (q >> shift) & 7u;
        words[nw++] = q * ve + li % ve;         /* element-sized word index */
      }
      int worst = 0;
      for (int b = 0; b < banks; b++) {
        uint32_t seen[32];
        int ns = 0;
        for (int k = 0; k < nw; k++)
          if ((int) (words[k] % banks) == b) {
            bool dup = false;
            for (int s = 0; s < ns; s++)
              dup = dup || seen[s] == words[k];
            if (!dup)
              seen[ns++] = words[k];
          }
        if (ns > worst)
          worst = ns;
      }
      cost += worst;
    }
  }
  return cost;
}

}  // namespace

int
perm_vec_build (const PermDesc &d0, size_t frame_elems, PermVecPlan *pl, void **host_table)
{
  memset (pl, 0, sizeof (*pl));
  *host_table = nullptr;
  if (getenv ("NNSB_PERM_NO_VEC"))
    return -1;
  const int es = d0.es;
  if (es != 4 && es != 8)
    return -1;
  const uint64_t VE = 16 / es;
  const uint64_t TILE_MAX = PV_STAGE_BYTES / es;
  if (((uint64_t) frame_elems * es) % 16)
    return -1;
  CDim dims[PV_DIMS];
  const int n = canonicalise (d0, dims);
  if (n < 2)
    return -1;                  /* identity: the flat copy */
  for (int i = 0; i < n; i++)
    if (dims[i].size > 0xffffffffull)
      return -1;

  int in_order[PV_DIMS], out_order[PV_DIMS];
  for (int i = 0; i < n; i++)
    in_order[i] = out_order[i] = i;
  for (int i = 1; i < n; i++)
    for (int j = i; j > 0 && dims[in_order[j]].in < dims[in_order[j - 1]].in; j--) {
      const int t = in_order[j];
      in_order[j] = in_order[j - 1];
      in_order[j - 1] = t;
    }

  /* 1. the box: runs of >= RUN bytes on both sides, then grown to <= 16 KiB */
  const char *env_run = getenv ("NNSB_PERM_VEC_RUN");
  const uint64_t RUN = (env_run ? (uint64_t) atoi (env_run) : 128) / es;
  uint64_t T[PV_DIMS];
  for (int i = 0; i < n; i++)
    T[i] = 1;
  ensure_run (dims, n, in_order, RUN, VE, T);
  ensure_run (dims, n, out_order, RUN, VE, T);
  for (int guard = 0; box_volume (T, n) > TILE_MAX && guard < 64; guard++) {
    int big = -1;
    for (int i = 0; i < n; i++)
      if (T[i] > VE && (big < 0 || T[i] > T[big]))
        big = i;
    if (big < 0)
      return -1;
    T[big] = ((T[big] + 1) / 2 + VE - 1) / VE * VE;
  }
  if (box_volume (T, n) > TILE_MAX)
    return -1;
  for (int round = 0; round < 64; round++) {
    bool grew = false;
    for (int side = 0; side < 2; side++) {
      const int *order = side == 0 ? out_order : in_order;
      for (int k = 0; k < n; k++) {
        const int d = order[k];
        if (T[d] < dims[d].size) {
          uint64_t nt = T[d] * 2;
          if (nt > dims[d].size)
            nt = dims[d].size;
          if (box_volume (T, n) / T[d] * nt <= TILE_MAX) {
            T[d] = nt;
            grew = true;
          }
          break;                /* only the first non-full dim of each side */
        }
      }
    }
    if (!grew)
      break;
  }
  const uint64_t volume = box_volume (T, n);
  if (volume % VE || volume > TILE_MAX)
    return -1;
  const uint32_t n_vec = (uint32_t) (volume / VE);

  /* 2. static conditions: box origins 16-byte aligned on both sides, offsets fit 32 bits,
   *    at most PV_BOXED dims with an extent > 1 */
  uint64_t span_in = 0, span_out = 0, tiles = 1;
  int n_boxed = 0;
  for (int i = 0; i < n; i++) {
    if ((T[i] * dims[i].in * es) % 16 || (T[i] * dims[i].out * es) % 16)
      if (T[i] < dims[i].size)
        return -1;
    span_in += (T[i] - 1) * dims[i].in * es;
    span_out += (T[i] - 1) * dims[i].out * es;
    tiles *= (dims[i].size + T[i] - 1) / T[i];
    pl->box_slot[i] = -1;
    if (T[i] > 1) {
      if (n_boxed == PV_BOXED || T[i] > 0xffff)
        return -1;
      pl->box_slot[i] = n_boxed++;
    }
  }
  if (span_in >= (1ull << 32) - 16 || span_out >= (1ull << 32) - 16 || tiles >= (1ull << 31))
    return -1;
  uint64_t limits_init = 0;
  for (int b = n_boxed; b < PV_BOXED; b++)
    limits_init |= 1ull << (16 * b);

  uint32_t lstride[PV_DIMS];
  {
    uint32_t ls = 1;
    for (int k = 0; k < n; k++) {
      lstride[in_order[k]] = ls;
      ls *= (uint32_t) T[in_order[k]];
    }
  }
  uint64_t rem[PV_DIMS];                /* extent of the last, partial box (0 = none) */
  for (int i = 0; i < n; i++)
    rem[i] = dims[i].size % T[i];

  /* 3. the table, verified element by element */
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
  uint64_t *t_ld_coord = (uint64_t *) (tab + PV_T_LD_COORD);
  uint64_t *t_st_coord = (uint64_t *) (tab + PV_T_ST_COORD);
  uint16_t *t_st_slot = (uint16_t *) (tab + PV_T_ST_SLOT);
  for (int i = 0; i < PV_NVT; i++)
    t_ld_off[i] = t_st_off[i] = PV_INVALID;

  bool ok = true;
  for (int side = 0; side < 2 && ok; side++) {
    const int *order = side == 0 ? in_order : out_order;
    for (uint32_t v = 0; v < n_vec && ok; v++) {
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
          if (pl->box_slot[d] >= 0)
            coord |= c << (16 * pl->box_slot[d]);
          cut[d] = rem[d] != 0 && c >= rem[d];
        }
        if (e == 0) {
          off0 = off;
          coord0 = coord;
          memcpy (cut0, cut, sizeof (cut));
        } else {
          ok = ok && off == off0 + e;   /* contiguous in global memory */
          for (int d = 0; d < n; d++)
            ok = ok && cut[d] == cut0[d];       /* an edge never cuts through a vector */
        }
        if (side == 1)
          st_lin[(size_t) v * VE + e] = (uint32_t) lin;
      }
      ok = ok && (off0 * es) % 16 == 0;
      if (side == 0) {
        t_ld_off[v] = (uint32_t) (off0 * es);
        t_ld_coord[v] = coord0;
      } else {
        t_st_off[v] = (uint32_t) (off0 * es);
        t_st_coord[v] = coord0;
      }
    }
  }
  if (!ok) {
    free (tab);
    free (st_lin);
    return -1;
  }

  /* 4. shared-memory swizzle: the candidate with the fewest gather wavefronts */
  int best_shift = -1;
  {
    const char *env_sw = getenv ("NNSB_PERM_VEC_SWIZZLE");
    if (env_sw) {
      best_shift = atoi (env_sw);
    } else {
      uint64_t best = gather_cost (st_lin, n_vec, (int) VE, es, -1);
      for (int s = 3; s <= 9; s++) {
        const uint64_t c = gather_cost (st_lin, n_vec, (int) VE, es, s);
        if (c < best) {
          best = c;
          best_shift = s;
        }
      }
    }
  }
  auto swz = [&] (uint32_t q) { return best_shift >= 0 ? q ^ ((q >> best_shift) & 7u) : q; };
  for (uint32_t v = 0; v < n_vec; v++) {
    t_ld_slot[v] = swz (v) * 16;
    for (uint64_t e = 0; e < VE; e++) {
      const uint32_t li = st_lin[(size_t) v * VE + e];
      t_st_slot[(size_t) v * VE + e] = (uint16_t) (swz (li / (uint32_t) VE) * 16 + (li % VE) * es);
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
  pl->swizzle_shift = best_shift;
  *host_table = tab;
  return 0;
}

template <int ES> static int
launch_permute_vec_t (const PermVecDev &p, const void *d_table, const void *in, void *out,
    const LaunchCfg &cfg, cudaStream_t stream)
{
  const size_t smem = (size_t) PV_STAGES * PV_STAGE_BYTES;
  static int resident = 0;
  if (resident == 0) {
    cudaError_t e = cudaFuncSetAttribute (permute_vec_kernel<ES>,
        cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e != cudaSuccess)
      return (int) e;
    resident = resident_blocks_per_sm (permute_vec_kernel<ES>, PV_THREADS, smem);
  }
  uint64_t grid = p.n_tiles;
  const uint64_t cap = (uint64_t) cfg.sm_count * (cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : resident);
  if (grid > cap)
    grid = cap;
  permute_vec_kernel<ES><<<(unsigned int) grid, PV_THREADS, smem, stream>>> (
      (const unsigned char *) in, (unsigned char *) out, (const unsigned char *) d_table, p);
  return (int) cudaGetLastError ();
}

int
launch_permute_vec (const PermVecPlan &pl, const void *d_table, const void *in, void *out,
    size_t frame_elems, uint32_t batch, const LaunchCfg &cfg, cudaStream_t stream)
{
  if (!pl.valid || !d_table || batch == 0)
    return -1;
  if ((uintptr_t) in % 16 || (uintptr_t) out % 16)
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
    p.in_stride_b[n] = (uint64_t) frame_elems * pl.es;
    p.out_stride_b[n] = (uint64_t) frame_elems * pl.es;
    p.box_slot[n] = -1;
    n++;
  }
  p.n = n;
  p.limits_init = pl.limits_init;
  p.n_vec = pl.n_vec;
  p.n_tiles = (uint32_t) tiles;
  return pl.es == 4 ? launch_permute_vec_t<4> (p, d_table, in, out, cfg, stream)
      : launch_permute_vec_t<8> (p, d_table, in, out, cfg, stream);
}

}  // namespace nnsb
