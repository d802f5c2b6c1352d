/*
 * k_perm.cu -- K3/K4/K8: data-movement kernels.
 *
 *  permute : rank-N permutation copy through a shared-memory tile.  Replaces the per-element
 *            nns_memcpy nest of gst_tensor_transform_transpose (gsttensor_transform.c:1417-1500)
 *            and the block memcpy nest of gst_tensor_transform_dimchg (:1155-1178).  Both lower
 *            to the same problem after the dimensions are canonicalised (drop extent-1 dims,
 *            merge dims that are adjacent on both sides).
 *            The tile is an N-dimensional BOX chosen so that it holds a run of >= 32
 *            contiguous elements on the input side (walking the dims by input stride) AND on
 *            the output side (walking them by output stride), whichever dims that takes --
 *            e.g. C:W:H -> H:W:C needs a box over all three dims.  A block reads its box in
 *            input order (dense warp loads), parks it in shared memory, and writes it in
 *            output order (dense warp stores).  The in-box coordinates of the <= 8 elements a
 *            thread moves depend only on the box shape, so their global offsets and
 *            shared-memory slots are computed once per kernel; per tile and element only an
 *            add, a load and a store remain.  Shared memory is padded by one word every 32.
 *            Video-shaped interleaved <-> planar cases go to the TMA pipeline in k_plane.cu.
 *  copy    : flat vector copy for permutations that canonicalise to the identity
 *            (dimchg from == to, :1133-1139; transpose 0:1:2:3, :1464-1469).
 *  strip   : tensor_converter's video row-padding removal (gsttensor_converter.c:1087-1100):
 *            rows of row_bytes at pitch RU4 (row_bytes) -> dense.  Rows are 1..3 bytes off
 *            word alignment, so each thread assembles aligned 16-byte output chunks from
 *            aligned 4-byte source words with funnel shifts.
 *  pad     : the inverse, tensor_decoder direct_video (tensordec-directvideo.c:326-339):
 *            dense rows -> pitch RU4 (row_bytes), padding bytes written as zero.
 *
 * Algorithmic bytes: 2 * es per element (permute, copy), 2 * row_bytes per row (strip).
 */
#include <cstdlib>

#include "nnsb_device.cuh"
#include "nnsb_kernels.h"
#include "k_perm_common.cuh"

namespace nnsb {

constexpr int PERM_THREADS = 256;
constexpr int PERM_EPT = 8;             /* elements per thread per tile */
constexpr int PERM_TILE_MAX = PERM_THREADS * PERM_EPT;   /* 2048 elements */
constexpr int PERM_DIMS = NNSB_PERM_MAX_RANK + 1;        /* rank 8 + batch */

struct PermDev
{
  int n;                                /* dims, in output order (out stride ascending) */
  uint32_t size[PERM_DIMS];
  FastDiv tile[PERM_DIMS];              /* box extent per dim */
  FastDiv ntile[PERM_DIMS];             /* boxes per dim */
  uint64_t in_stride[PERM_DIMS];
  uint64_t out_stride[PERM_DIMS];
  int in_order[PERM_DIMS];              /* dims sorted by input stride (fastest first) */
  uint32_t lstride[PERM_DIMS];          /* stride of dim d in the load-order linear index */
  uint32_t volume;                      /* elements per box (<= 2048) */
  uint32_t n_tiles;                     /* < 2^31 */
};

__device__ __forceinline__ uint32_t smem_slot (uint32_t linear)
{
  return linear + (linear >> 5);        /* one pad word every 32: breaks power-of-two strides */
}

constexpr int PERM_STAGES = 4;

template <typename E> __device__ __forceinline__ void
cp_async_elem (E *smem_dst, const E *gmem_src)
{
  const uint32_t d = (uint32_t) __cvta_generic_to_shared (smem_dst);
  if constexpr (sizeof (E) == 8)
    asm volatile ("cp.async.ca.shared.global [%0], [%1], 8;" :: "r" (d), "l" (gmem_src) : "memory");
  else if constexpr (sizeof (E) == 4)
    asm volatile ("cp.async.ca.shared.global [%0], [%1], 4;" :: "r" (d), "l" (gmem_src) : "memory");
}

/* ASYNC: loads are cp.async (LDGSTS) into a ring of PERM_STAGES tiles, PERM_STAGES-1 tiles of
 * reads in flight per block (4/8-byte elements: cp.async moves 4, 8 or 16 bytes).
 * !ASYNC: loads go through registers, one tile at a time (1/2-byte elements). */
template <typename E, bool ASYNC>
__global__ void __launch_bounds__ (PERM_THREADS)
permute_kernel (const E *__restrict__ in, E *__restrict__ out, const __grid_constant__ PermDev p)
{
  extern __shared__ __align__ (16) unsigned char smem_raw[];
  constexpr uint32_t STAGE_ELEMS = PERM_TILE_MAX + PERM_TILE_MAX / 32 + 1;
  E *ring = reinterpret_cast<E *> (smem_raw);

  /* in-box coordinates of this thread's elements: computed once per kernel */
  /* (in-box offsets fit 32 bits: the launcher checks the box span) */
  uint32_t l_sm[PERM_EPT], s_sm[PERM_EPT];
  uint32_t l_off[PERM_EPT], s_off[PERM_EPT];
#pragma unroll
  for (int j = 0; j < PERM_EPT; j++) {
    const uint32_t i = threadIdx.x + j * PERM_THREADS;
    uint32_t q = i, c;
    uint32_t off = 0;
    for (int k = 0; k < p.n; k++) {
      const int d = p.in_order[k];
      divmod (q, p.tile[d], q, c);
      off += c * (uint32_t) p.in_stride[d];
    }
    l_off[j] = off;
    l_sm[j] = smem_slot (i);
    q = i;
    off = 0;
    uint32_t lin = 0;
    for (int d = 0; d < p.n; d++) {
      divmod (q, p.tile[d], q, c);
      off += c * (uint32_t) p.out_stride[d];
      lin += c * p.lstride[d];
    }
    s_off[j] = off;
    s_sm[j] = smem_slot (lin);
  }
  const bool whole = (p.volume == (uint32_t) PERM_TILE_MAX);   /* every thread owns 8 elements */

  uint32_t my_tiles = 0;
  if (blockIdx.x < p.n_tiles)
    my_tiles = (p.n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;

  /* Where a tile sits (its base offsets, whether the box lies fully inside the tensor, its
   * origin for the edge test) is worked out by ONE thread per tile and shared through a small
   * ring: the decode walks all dims and is far too long to be repeated by 256 threads. */
  struct TileInfo
  {
    uint64_t in_base, out_base;
    uint32_t origin[PERM_DIMS];
    int full;
  };
  constexpr int INFO = PERM_STAGES + 1;
  __shared__ TileInfo info[INFO];
  auto locate = [&] (uint32_t k) {
    TileInfo &ti = info[k % INFO];
    uint32_t q = blockIdx.x + k * gridDim.x, td;
    bool full = true;
    uint64_t ib = 0, ob = 0;
    for (int d = 0; d < p.n; d++) {
      divmod (q, p.ntile[d], q, td);
      const uint32_t o = td * p.tile[d].d;
      ti.origin[d] = o;
      ib += (uint64_t) o * p.in_stride[d];
      ob += (uint64_t) o * p.out_stride[d];
      full = full && (o + p.tile[d].d <= p.size[d]);
    }
    ti.in_base = ib;
    ti.out_base = ob;
    ti.full = full ? 1 : 0;
  };
  auto inside = [&] (uint32_t i, const uint32_t *origin, bool load_order) {
    uint32_t q = i, c;
    bool ok = true;
    for (int k = 0; k < p.n; k++) {
      const int d = load_order ? p.in_order[k] : k;
      divmod (q, p.tile[d], q, c);
      ok = ok && (origin[d] + c < p.size[d]);
    }
    return ok;
  };

  if constexpr (ASYNC) {
    auto issue = [&] (uint32_t k) {
      if (k < my_tiles) {
        const TileInfo &ti = info[k % INFO];
        E *stage = ring + (size_t) (k % PERM_STAGES) * STAGE_ELEMS;
        const E *src = in + ti.in_base;
        if (ti.full && whole) {
#pragma unroll
          for (int j = 0; j < PERM_EPT; j++)
            cp_async_elem (stage + l_sm[j], src + l_off[j]);
        } else if (ti.full) {
#pragma unroll
          for (int j = 0; j < PERM_EPT; j++)
            if (threadIdx.x + j * PERM_THREADS < p.volume)
              cp_async_elem (stage + l_sm[j], src + l_off[j]);
        } else {
          for (int j = 0; j < PERM_EPT; j++) {
            const uint32_t i = threadIdx.x + j * PERM_THREADS;
            if (i < p.volume && inside (i, ti.origin, true))
              cp_async_elem (stage + l_sm[j], src + l_off[j]);
          }
        }
      }
      asm volatile ("cp.async.commit_group;" ::: "memory");   /* possibly empty: keeps the count */
    };
    if (threadIdx.x == 0)
      for (uint32_t k = 0; k < (uint32_t) PERM_STAGES && k < my_tiles; k++)
        locate (k);
    __syncthreads ();
    for (int s = 0; s < PERM_STAGES - 1; s++)
      issue ((uint32_t) s);
    for (uint32_t k = 0; k < my_tiles; k++) {
      issue (k + PERM_STAGES - 1);
      asm volatile ("cp.async.wait_group %0;" :: "n" (PERM_STAGES - 1) : "memory");
      __syncthreads ();
      if (threadIdx.x == 0 && k + PERM_STAGES < my_tiles)
        locate (k + PERM_STAGES);       /* slot (k + S) % (S + 1): not read before the next barrier */
      const TileInfo &ti = info[k % INFO];
      const E *stage = ring + (size_t) (k % PERM_STAGES) * STAGE_ELEMS;
      E *dst = out + ti.out_base;
      if (ti.full && whole) {
        E w[PERM_EPT];
#pragma unroll
        for (int j = 0; j < PERM_EPT; j++)
          w[j] = stage[s_sm[j]];
#pragma unroll
        for (int j = 0; j < PERM_EPT; j++)
          st_stream<sizeof (E)> (dst + s_off[j], &w[j]);
      } else if (ti.full) {
#pragma unroll
        for (int j = 0; j < PERM_EPT; j++)
          if (threadIdx.x + j * PERM_THREADS < p.volume) {
            E w = stage[s_sm[j]];
            st_stream<sizeof (E)> (dst + s_off[j], &w);
          }
      } else {
        for (int j = 0; j < PERM_EPT; j++) {
          const uint32_t i = threadIdx.x + j * PERM_THREADS;
          if (i < p.volume && inside (i, ti.origin, false)) {
            E w = stage[s_sm[j]];
            st_stream<sizeof (E)> (dst + s_off[j], &w);
          }
        }
      }
      __syncthreads ();         /* the stage is overwritten by the issue of the next iteration */
    }
  } else {
    if (threadIdx.x == 0 && my_tiles > 0)
      locate (0);
    __syncthreads ();
    for (uint32_t k = 0; k < my_tiles; k++) {
      const TileInfo &ti = info[k % INFO];
      const E *src = in + ti.in_base;
      E *dst = out + ti.out_base;
      E v[PERM_EPT];
      if (ti.full) {
#pragma unroll
        for (int j = 0; j < PERM_EPT; j++)
          if (threadIdx.x + j * PERM_THREADS < p.volume)
            ld_stream<sizeof (E)> (&v[j], src + l_off[j]);
      } else {
#pragma unroll
        for (int j = 0; j < PERM_EPT; j++) {
          const uint32_t i = threadIdx.x + j * PERM_THREADS;
          if (i < p.volume && inside (i, ti.origin, true))
            ld_stream<sizeof (E)> (&v[j], src + l_off[j]);
        }
      }
#pragma unroll
      for (int j = 0; j < PERM_EPT; j++)
        if (threadIdx.x + j * PERM_THREADS < p.volume)
          ring[l_sm[j]] = v[j];
      if (threadIdx.x == 0 && k + 1 < my_tiles)
        locate (k + 1);
      __syncthreads ();
      if (ti.full) {
#pragma unroll
        for (int j = 0; j < PERM_EPT; j++)
          if (threadIdx.x + j * PERM_THREADS < p.volume) {
            E w = ring[s_sm[j]];
            st_stream<sizeof (E)> (dst + s_off[j], &w);
          }
      } else {
        for (int j = 0; j < PERM_EPT; j++) {
          const uint32_t i = threadIdx.x + j * PERM_THREADS;
          if (i < p.volume && inside (i, ti.origin, false)) {
            E w = ring[s_sm[j]];
            st_stream<sizeof (E)> (dst + s_off[j], &w);
          }
        }
      }
      __syncthreads ();
    }
  }
}

/* flat copy: 32 bytes per thread per step, 4 steps in flight */
__global__ void __launch_bounds__ (PERM_THREADS)
copy_kernel (const unsigned char *__restrict__ in, unsigned char *__restrict__ out, size_t bytes,
    size_t n_vec)
{
  const size_t tid = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
  const size_t nthreads = (size_t) gridDim.x * blockDim.x;
  for (size_t v = tid; v < n_vec; v += nthreads * 4) {
    Bytes<32> x[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const size_t vu = v + (size_t) u * nthreads;
      if (vu < n_vec)
        ld_stream<32> (&x[u], in + vu * 32);
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const size_t vu = v + (size_t) u * nthreads;
      if (vu < n_vec)
        st_stream<32> (out + vu * 32, &x[u]);
    }
  }
  for (size_t b = n_vec * 32 + tid; b < bytes; b += nthreads)
    out[b] = in[b];
}

static int
launch_copy (const void *in, void *out, size_t bytes, const LaunchCfg &cfg, cudaStream_t stream)
{
  if (bytes == 0)
    return 0;
  const bool aligned = ((uintptr_t) in % 32 == 0) && ((uintptr_t) out % 32 == 0);
  const size_t n_vec = aligned ? bytes / 32 : 0;
  const size_t work = n_vec ? n_vec : bytes;
  size_t blocks = (work + (size_t) PERM_THREADS * 4 - 1) / ((size_t) PERM_THREADS * 4);
  const size_t cap = (size_t) cfg.sm_count * (cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : 256);
  if (blocks > cap)
    blocks = cap;
  if (blocks == 0)
    blocks = 1;
  copy_kernel<<<(unsigned int) blocks, PERM_THREADS, 0, stream>>> ((const unsigned char *) in,
      (unsigned char *) out, bytes, n_vec);
  return (int) cudaGetLastError ();
}

template <typename E, bool ASYNC> static int
launch_permute_t (const PermDev &p, const void *in, void *out, const LaunchCfg &cfg,
    cudaStream_t stream)
{
  const size_t smem = (size_t) (ASYNC ? PERM_STAGES : 1) * (PERM_TILE_MAX + PERM_TILE_MAX / 32 + 1)
      * sizeof (E);
  /* per device: the attribute and the occupancy answer belong to a device's context */
  static int resident_of[64];
  int dev_id = 0;
  cudaGetDevice (&dev_id);
  int &resident = resident_of[(dev_id >= 0 && dev_id < 64) ? dev_id : 0];
  if (resident == 0 || dev_id >= 64) {
    if (smem > 48 * 1024) {
      cudaError_t e = cudaFuncSetAttribute (permute_kernel<E, ASYNC>,
          cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
      if (e != cudaSuccess)
        return (int) e;
    }
    resident = resident_blocks_per_sm (permute_kernel<E, ASYNC>, PERM_THREADS, smem);
  }
  /* persistent: every block amortises its coordinate set-up over many tiles */
  uint64_t grid = p.n_tiles;
  const uint64_t cap = (uint64_t) cfg.sm_count * (cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : resident);
  if (grid > cap)
    grid = cap;
  permute_kernel<E, ASYNC><<<(unsigned int) grid, PERM_THREADS, smem, stream>>> ((const E *) in,
      (E *) out, p);
  return (int) cudaGetLastError ();
}

/* which kernel a permutation of `batch` tensors at 256-byte aligned addresses would take */
const char *
permute_path_name (const PermDesc &d0, size_t frame_elems, uint32_t batch, const PermVecPlan *vec)
{
  CDim dims[PERM_DIMS];
  int es = d0.es;
  uint64_t alignment = 256;
  if (batch > 1)
    alignment |= (uint64_t) frame_elems * d0.es;
  const int n = canonicalise_perm (d0.size, d0.in_stride, d0.rank, frame_elems, batch, 8, alignment, dims, &es);
  if (n == 0 || (n == 1 && dims[0].in == 1))
    return "copy";
  if (launch_pixel_transpose (dims, n, es, (const void *) 256, (void *) 256, nullptr, true) == 0)
    return "pixel";
  const char *env = getenv ("NNSB_PERM_SLAB");
  const int mode = env ? atoi (env) : 1;
  LaunchCfg cfg;
  cfg.sm_count = 148;
  cfg.blocks_per_sm = 0;
  if (mode && n <= 8 && (mode == 2 || !vec || !vec->valid || vec->vb < 16)
      && launch_permute_slab (dims, n, es, (const void *) 256, (void *) 256,
          (uint64_t) frame_elems * batch * d0.es / es, cfg, nullptr, true) == 0)
    return "slab";
  return (vec && vec->valid) ? "vector" : "scalar";
}

int launch_permute (const PermDesc &d0, const void *in, void *out, size_t frame_elems,
    uint32_t batch, const LaunchCfg &cfg, cudaStream_t stream, const PermVecPlan *vec,
    const void *d_table)
{
  if (batch == 0 || frame_elems == 0)
    return 0;
  /* 1. canonical form: extent-1 dims dropped, batch appended, neighbours merged, an innermost
   *    block common to both sides folded into the element (RGBA bytes move as one word) */
  CDim dims[PERM_DIMS];
  int es = d0.es;
  uint64_t alignment = (uint64_t) (uintptr_t) in | (uint64_t) (uintptr_t) out;
  if (batch > 1)
    alignment |= (uint64_t) frame_elems * d0.es;
  int n = canonicalise_perm (d0.size, d0.in_stride, d0.rank, frame_elems, batch, 8, alignment, dims,
      &es);
  const size_t total = (size_t) frame_elems * batch;
  if (n == 0 || (n == 1 && dims[0].in == 1))
    return launch_copy (in, out, total * d0.es, cfg, stream);
  for (int i = 0; i < n; i++)
    if (dims[i].size > 0xffffffffull)
      return (int) cudaErrorInvalidValue;

  /* 3. video-shaped [P][C] <-> [C][P] (2..4 channels): the TMA pipeline (k_plane.cu) */
  if (!getenv ("NNSB_NO_PLANE")) {
    /* find X (out stride 1) and Y (in stride 1); everything else is an outer dim */
    int ix = -1, iy = -1;
    for (int i = 0; i < n; i++) {
      if (dims[i].out == 1) ix = i;
      if (dims[i].in == 1) iy = i;
    }
    if (ix >= 0 && iy >= 0 && ix != iy) {
      uint32_t o_size[PERM_DIMS];
      uint64_t o_in[PERM_DIMS], o_out[PERM_DIMS];
      int no = 0;
      for (int i = 0; i < n; i++)
        if (i != ix && i != iy) {
          o_size[no] = (uint32_t) dims[i].size;
          o_in[no] = dims[i].in;
          o_out[no] = dims[i].out;
          no++;
        }
      int r = -1;
      if (dims[iy].size >= 2 && dims[iy].size <= 4 && dims[ix].in == dims[iy].size)
        r = launch_plane (true, (int) dims[iy].size, es, dims[ix].size, dims[iy].out, no, o_size,
            o_in, o_out, in, out, cfg, stream);
      else if (dims[ix].size >= 2 && dims[ix].size <= 4 && dims[iy].out == dims[ix].size)
        r = launch_plane (false, (int) dims[ix].size, es, dims[iy].size, dims[ix].in, no, o_size,
            o_out, o_in, in, out, cfg, stream);
      if (r != -1)
        return r;
    }
  }

  /* 3a. the image transpose of packed RGB bytes (k_perm_pix.cu) */
  {
    const int r = launch_pixel_transpose (dims, n, es, in, out, stream);
    if (r != -1)
      return r;
  }

  /* 3b. rows that are not multiples of 16 bytes (the vector plan is 8 / 4 bytes wide or missing)
   *     but long output runs: contiguous input slabs through shared memory (k_perm_slab.cu) */
  {
    const char *env = getenv ("NNSB_PERM_SLAB");
    const int mode = env ? atoi (env) : 1;      /* 0 = never, 1 = when the vector path is narrow, 2 = whenever it fits */
    /* whenever the vector plan is narrower than 16 bytes: its per-vector overhead is that of a
     * 16-byte vector, the slab kernel's store phase works per element */
    if (mode && n <= 8 && (mode == 2 || !vec || !vec->valid || vec->vb < 16)) {
      const int r = launch_permute_slab (dims, n, es, in, out, total * d0.es / es, cfg, stream);
      if (r != -1)
        return r;
    }
  }

  /* 4. 4-byte and wider elements: the table-driven 16-byte vector kernel (k_perm_vec.cu), when
   *    the plan qualified */
  if (vec && vec->valid) {
    const int r = launch_permute_vec (*vec, d_table, in, out, frame_elems, batch, cfg, stream);
    if (r != -1)
      return r;
  }

  /* 5. the box: >= 32-element runs on both sides, then grown to ~2048 elements */
  PermDev p;
  memset (&p, 0, sizeof (p));
  p.n = n;
  int in_order[PERM_DIMS];
  for (int i = 0; i < n; i++)
    in_order[i] = i;
  for (int i = 1; i < n; i++)           /* insertion sort by input stride */
    for (int j = i; j > 0 && dims[in_order[j]].in < dims[in_order[j - 1]].in; j--) {
      const int t = in_order[j];
      in_order[j] = in_order[j - 1];
      in_order[j - 1] = t;
    }
  uint64_t T[PERM_DIMS];
  for (int i = 0; i < n; i++)
    T[i] = 1;
  const uint64_t RUN = getenv ("NNSB_PERM_RUN") ? (uint64_t) atoi (getenv ("NNSB_PERM_RUN")) : 32;
  {
    uint64_t prod = 1;
    for (int k = 0; k < n && prod < RUN; k++) {
      const int d = in_order[k];
      uint64_t want = (RUN + prod - 1) / prod;
      if (want > dims[d].size) want = dims[d].size;
      if (T[d] < want) T[d] = want;
      prod *= want;
    }
    prod = 1;
    for (int d = 0; d < n && prod < RUN; d++) {
      uint64_t want = (RUN + prod - 1) / prod;
      if (want > dims[d].size) want = dims[d].size;
      if (T[d] < want) T[d] = want;
      prod *= want;
    }
  }
  auto volume = [&] () { uint64_t v = 1; for (int i = 0; i < n; i++) v *= T[i]; return v; };
  /* too large (both runs needed many dims): shrink the largest extents */
  while (volume () > (uint64_t) PERM_TILE_MAX) {
    int big = 0;
    for (int i = 1; i < n; i++)
      if (T[i] > T[big]) big = i;
    T[big] = (T[big] + 1) / 2;
  }
  /* grow: alternately lengthen the output run and the input run */
  for (int round = 0; round < 64; round++) {
    bool grew = false;
    for (int side = 0; side < 2; side++) {
      for (int k = 0; k < n; k++) {
        const int d = side == 0 ? k : in_order[k];
        if (T[d] < dims[d].size) {
          uint64_t nt = T[d] * 2;
          if (nt > dims[d].size) nt = dims[d].size;
          if (volume () / T[d] * nt <= (uint64_t) PERM_TILE_MAX) {
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
  uint64_t tiles = 1;
  for (int i = 0; i < n; i++) {
    p.size[i] = (uint32_t) dims[i].size;
    p.tile[i] = make_fastdiv ((uint32_t) T[i]);
    p.ntile[i] = make_fastdiv ((uint32_t) ((dims[i].size + T[i] - 1) / T[i]));
    p.in_stride[i] = dims[i].in;
    p.out_stride[i] = dims[i].out;
    p.in_order[i] = in_order[i];
    tiles *= p.ntile[i].d;
  }
  {
    uint32_t ls = 1;
    for (int k = 0; k < n; k++) {
      p.lstride[in_order[k]] = ls;
      ls *= p.tile[in_order[k]].d;
    }
  }
  p.volume = (uint32_t) volume ();
  if (tiles >= (1ull << 31))
    return (int) cudaErrorInvalidValue;
  {
    /* in-box offsets are kept in 32 bits: shrink boxes whose span would not fit (huge strides) */
    uint64_t span_in = 0, span_out = 0;
    for (int i = 0; i < n; i++) {
      span_in += (uint64_t) (T[i] - 1) * dims[i].in;
      span_out += (uint64_t) (T[i] - 1) * dims[i].out;
    }
    if (span_in >= (1ull << 32) || span_out >= (1ull << 32))
      return (int) cudaErrorInvalidValue;
  }
  p.n_tiles = (uint32_t) tiles;

  const bool sync_only = getenv ("NNSB_PERM_SYNC") != nullptr;
  switch (es) {
    case 1: return launch_permute_t<uint8_t, false> (p, in, out, cfg, stream);
    case 2: return launch_permute_t<uint16_t, false> (p, in, out, cfg, stream);
    case 4: return sync_only ? launch_permute_t<uint32_t, false> (p, in, out, cfg, stream)
        : launch_permute_t<uint32_t, true> (p, in, out, cfg, stream);
    case 8: return sync_only ? launch_permute_t<uint64_t, false> (p, in, out, cfg, stream)
        : launch_permute_t<uint64_t, true> (p, in, out, cfg, stream);
    default: return (int) cudaErrorInvalidValue;
  }
}

/* ------------------------------------------------------------------------------------- */
/* strip / pad: rows re-pitched                                                            */
/* ------------------------------------------------------------------------------------- */
constexpr int STRIP_THREADS = 256;

/* `rows` rows of row_bytes bytes move from pitch ps to pitch pd >= row_bytes; bytes
 * [row_bytes, pd) of every destination row are written as zero.
 *   strip (gsttensor_converter.c:1087-1100): ps = RU4 (row_bytes), pd = row_bytes
 *   pad   (tensordec-directvideo.c:326-339): ps = row_bytes,        pd = RU4 (row_bytes)
 * A thread owns aligned 16-byte destination chunks; its (row, col) advances by a constant
 * step, so the loop holds no division.  A chunk that lies inside one row's payload is
 * assembled from five aligned source words with funnel shifts (the source is 0..3 bytes off
 * word alignment); the few chunks that touch a row end or the padding go byte by byte. */
__global__ void __launch_bounds__ (STRIP_THREADS)
repitch_kernel (const unsigned char *__restrict__ src, unsigned char *__restrict__ dst,
    uint32_t row_bytes, uint64_t ps, uint32_t pd, uint64_t total_dst, uint64_t src_bytes,
    uint32_t step_row, uint32_t step_col)
{
  const uint64_t tid = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t step = ((uint64_t) gridDim.x * blockDim.x) * 16;
  const uint64_t n_chunk_bytes = total_dst & ~15ull;
  uint64_t o = tid * 16;
  uint64_t row = o / pd;
  uint32_t col = (uint32_t) (o - row * pd);
  for (; o < n_chunk_bytes; o += step) {
    uint32_t w[4];
    const uint64_t s = row * ps + col;
    const uint64_t base = s & ~3ull;
    if (col + 16 <= row_bytes && base + 20 <= src_bytes) {
      const uint32_t *p = reinterpret_cast<const uint32_t *> (src + base);
      const uint32_t sh = (uint32_t) (s & 3) * 8;
      uint32_t v[5];
#pragma unroll
      for (int k = 0; k < 5; k++)
        v[k] = __ldg (p + k);
#pragma unroll
      for (int k = 0; k < 4; k++)
        w[k] = __funnelshift_r (v[k], v[k + 1], sh);    /* sh == 0: v[k] */
    } else {
      uint64_t r = row;
      uint32_t c = col;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        uint32_t x = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) {
          if (c < row_bytes)
            x |= (uint32_t) src[r * ps + c] << (8 * b);
          if (++c == pd) {
            c = 0;
            r++;
          }
        }
        w[k] = x;
      }
    }
    st_stream<16> (dst + o, w);
    col += step_col;
    row += step_row;
    if (col >= pd) {
      col -= pd;
      row++;
    }
  }
  /* last (total_dst % 16) bytes */
  for (uint64_t b = n_chunk_bytes + tid; b < total_dst; b += step / 16) {
    const uint64_t r = b / pd;
    const uint32_t c = (uint32_t) (b - r * pd);
    dst[b] = c < row_bytes ? src[r * ps + c] : (unsigned char) 0;
  }
}

/* any alignment */
__global__ void __launch_bounds__ (STRIP_THREADS)
repitch_kernel_b (const unsigned char *__restrict__ src, unsigned char *__restrict__ dst,
    uint64_t row_bytes, uint64_t ps, uint64_t pd, uint64_t total_dst)
{
  const uint64_t tid = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t nthreads = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t b = tid; b < total_dst; b += nthreads) {
    const uint64_t r = b / pd;
    const uint64_t c = b - r * pd;
    dst[b] = c < row_bytes ? src[r * ps + c] : (unsigned char) 0;
  }
}

static int
launch_repitch (const void *src, void *dst, size_t row_bytes, size_t ps, size_t pd, size_t rows,
    const LaunchCfg &cfg, cudaStream_t stream)
{
  const uint64_t total_dst = (uint64_t) pd * (rows - 1) + (pd > row_bytes ? pd : row_bytes);
  const uint64_t src_bytes = (uint64_t) ps * (rows - 1) + row_bytes;
  const bool word_ok = ((uintptr_t) dst % 16 == 0) && ((uintptr_t) src % 4 == 0)
      && row_bytes >= 16 && pd <= 0x7fffffffull;
  const uint64_t work = word_ok ? (total_dst + 15) / 16 : total_dst;
  uint64_t blocks = (work + STRIP_THREADS - 1) / STRIP_THREADS;
  const uint64_t cap = (uint64_t) cfg.sm_count * (cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : 64);
  if (blocks > cap)
    blocks = cap;
  if (blocks == 0)
    blocks = 1;
  if (word_ok) {
    const uint64_t step = blocks * STRIP_THREADS * 16;
    repitch_kernel<<<(unsigned int) blocks, STRIP_THREADS, 0, stream>>> ((const unsigned char *) src,
        (unsigned char *) dst, (uint32_t) row_bytes, ps, (uint32_t) pd, total_dst, src_bytes,
        (uint32_t) (step / pd), (uint32_t) (step % pd));
  } else {
    repitch_kernel_b<<<(unsigned int) blocks, STRIP_THREADS, 0, stream>>> ((const unsigned char *) src,
        (unsigned char *) dst, row_bytes, ps, pd, total_dst);
  }
  return (int) cudaGetLastError ();
}

int launch_strip (const void *src, void *dst, size_t row_bytes, size_t src_pitch, size_t rows,
    const LaunchCfg &cfg, cudaStream_t stream)
{
  if (rows == 0 || row_bytes == 0)
    return 0;
  if (src_pitch < row_bytes)
    return (int) cudaErrorInvalidValue;
  if (src_pitch == row_bytes)
    return launch_copy (src, dst, row_bytes * rows, cfg, stream);
  return launch_repitch (src, dst, row_bytes, src_pitch, row_bytes, rows, cfg, stream);
}

int launch_pad (const void *src, void *dst, size_t row_bytes, size_t dst_pitch, size_t rows,
    const LaunchCfg &cfg, cudaStream_t stream)
{
  if (rows == 0 || row_bytes == 0)
    return 0;
  if (dst_pitch < row_bytes)
    return (int) cudaErrorInvalidValue;
  if (dst_pitch == row_bytes)
    return launch_copy (src, dst, row_bytes * rows, cfg, stream);
  return launch_repitch (src, dst, row_bytes, row_bytes, dst_pitch, rows, cfg, stream);
}

/* ------------------------------------------------------------------------------------- */
/* merge: tensor_merge mode=linear (gsttensor_merge.c:506-586)                             */
/* ------------------------------------------------------------------------------------- */
/* out row r = [ in_0 block r | in_1 block r | ... ], block k is blk[k] bytes (the input's extent
 * along the merge axis and everything below it), rows = product of the dims above the axis.
 * Same scheme as repitch_kernel: a thread owns aligned 16-byte output chunks, (row, col) advance
 * without a division, a chunk inside one block is assembled from five aligned source words. */
struct MergeDev
{
  int n;
  const unsigned char *in[NNSB_MERGE_MAX];
  uint32_t blk[NNSB_MERGE_MAX];
  uint32_t off[NNSB_MERGE_MAX + 1];     /* prefix sums of blk; off[n] = row bytes */
  uint64_t in_bytes[NNSB_MERGE_MAX];
};

__device__ __forceinline__ int
merge_find (const MergeDev &m, uint32_t col)
{
  int k = 0;
  while (k + 1 < m.n && col >= m.off[k + 1])
    k++;
  return k;
}

template <bool WORDS>
__global__ void __launch_bounds__ (STRIP_THREADS)
merge_kernel (unsigned char *__restrict__ dst, const __grid_constant__ MergeDev m, uint64_t total,
    uint32_t step_row, uint32_t step_col)
{
  const uint32_t S = m.off[m.n];
  const uint64_t tid = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t nthreads = (uint64_t) gridDim.x * blockDim.x;
  const uint64_t n_chunk_bytes = WORDS ? (total & ~15ull) : 0;
  if (WORDS) {
    const uint64_t step = nthreads * 16;
    uint64_t o = tid * 16;
    uint64_t row = o / S;
    uint32_t col = (uint32_t) (o - row * S);
    for (; o < n_chunk_bytes; o += step) {
      uint32_t w[4];
      const int k = merge_find (m, col);
      const uint64_t s = row * m.blk[k] + (col - m.off[k]);
      const uint64_t base = s & ~3ull;
      if (col + 16 <= m.off[k + 1] && base + 20 <= m.in_bytes[k]) {
        const uint32_t *p = reinterpret_cast<const uint32_t *> (m.in[k] + base);
        const uint32_t sh = (uint32_t) (s & 3) * 8;
        uint32_t v[5];
#pragma unroll
        for (int j = 0; j < 5; j++)
          v[j] = __ldg (p + j);
#pragma unroll
        for (int j = 0; j < 4; j++)
          w[j] = __funnelshift_r (v[j], v[j + 1], sh);
      } else {
        uint64_t r = row;
        uint32_t c = col;
#pragma unroll
        for (int j = 0; j < 4; j++) {
          uint32_t x = 0;
#pragma unroll
          for (int b = 0; b < 4; b++) {
            const int kk = merge_find (m, c);
            x |= (uint32_t) m.in[kk][r * m.blk[kk] + (c - m.off[kk])] << (8 * b);
            if (++c == S) {
              c = 0;
              r++;
            }
          }
          w[j] = x;
        }
      }
      st_stream<16> (dst + o, w);
      col += step_col;
      row += step_row;
      if (col >= S) {
        col -= S;
        row++;
      }
    }
  }
  for (uint64_t b = n_chunk_bytes + tid; b < total; b += nthreads) {
    const uint64_t r = b / S;
    const uint32_t c = (uint32_t) (b - r * S);
    const int k = merge_find (m, c);
    dst[b] = m.in[k][r * m.blk[k] + (c - m.off[k])];
  }
}

/* Small blocks (RGB + alpha -> RGBA: 3 + 1 bytes per row): every 16-byte output chunk mixes
 * bytes of several inputs, and the generic byte path above spends its time searching for them.
 * But the pattern repeats with period P = lcm (row bytes, 16): the host lists, for each of the P
 * output bytes of a period, the input it comes from and its offset inside that input's share
 * of the period.  The grid stride is a multiple of P / 16 chunks, so a thread always produces
 * the same phase of the period: its 16 source addresses are set up once and then only advance
 * by a constant -- per byte one load, one add, one insert. */
constexpr int MERGE_PERIOD_MAX = 256;
struct MergePhaseDev
{
  int n;
  uint32_t chunks_per_period;
  const unsigned char *in[NNSB_MERGE_MAX];
  uint64_t in_bytes[NNSB_MERGE_MAX];
  uint32_t run[NNSB_MERGE_MAX];         /* bytes of input k consumed per period */
  uint16_t delta[MERGE_PERIOD_MAX];     /* per output byte of a period: offset inside that run */
  uint8_t src[MERGE_PERIOD_MAX];        /* ... and the input it comes from */
};

constexpr int MERGE_PHASE_J = 4;                /* chunks per thread and batch */
constexpr int MERGE_PHASE_SMEM = 2 * 16 * STRIP_THREADS * MERGE_PHASE_J / 2 + 32 * NNSB_MERGE_MAX;

/* The inputs' shares of a batch of periods are contiguous byte ranges: they are copied to shared
 * memory with 16-byte loads (aligned supersets, any alignment), and the byte gather runs on
 * ld.shared (byte loads from global memory were bound by L1 tag look-ups: 2.4 TB/s). */
__global__ void __launch_bounds__ (STRIP_THREADS)
merge_phase_kernel (unsigned char *__restrict__ dst, const __grid_constant__ MergePhaseDev m,
    uint64_t total, uint32_t threads_used, uint64_t n_batches)
{
  extern __shared__ __align__ (16) unsigned char stage[];
  __shared__ uint32_t region[NNSB_MERGE_MAX + 1];       /* where input k's bytes start in `stage` */
  const uint32_t cpp = m.chunks_per_period;
  const uint32_t tid = threadIdx.x;
  const bool worker = tid < threads_used;
  const uint32_t period0 = tid / cpp, phase = tid - period0 * cpp;
  const uint32_t periods_per_step = threads_used / cpp;         /* threads_used is a multiple of cpp */
  const uint32_t periods_per_batch = periods_per_step * MERGE_PHASE_J;
  const uint64_t batch_bytes = (uint64_t) periods_per_batch * cpp * 16;
  /* per thread: input and in-run offset of its 16 bytes (the phase never changes) */
  uint32_t off[16], adv[16];
  uint8_t src[16];
#pragma unroll
  for (int i = 0; i < 16; i++) {
    const uint32_t e = phase * 16 + i;
    src[i] = m.src[e];
    off[i] = period0 * m.run[src[i]] + m.delta[e];
    adv[i] = periods_per_step * m.run[src[i]];
  }
  for (uint64_t b = blockIdx.x; b < n_batches; b += gridDim.x) {
    const uint64_t first_period = b * periods_per_batch;
    /* ---- stage the inputs' shares of this batch ---- */
    if (tid == 0) {
      uint32_t pos = 0;
      for (int k = 0; k < m.n; k++) {
        const uintptr_t a0 = (uintptr_t) m.in[k] + first_period * m.run[k];
        region[k] = pos + (uint32_t) (a0 & 15);
        pos += (((uint32_t) (a0 & 15) + periods_per_batch * m.run[k] + 15) / 16) * 16;
      }
    }
    __syncthreads ();
    for (int k = 0; k < m.n; k++) {
      const uintptr_t lo = (uintptr_t) m.in[k], hi = lo + m.in_bytes[k];
      const uintptr_t a0 = lo + first_period * m.run[k];
      const uintptr_t ga = a0 & ~(uintptr_t) 15;
      const uint32_t n16 = ((uint32_t) (a0 - ga) + periods_per_batch * m.run[k] + 15) / 16;
      unsigned char *dstk = stage + (region[k] - (uint32_t) (a0 - ga));
      for (uint32_t c = tid; c < n16; c += STRIP_THREADS) {
        const uintptr_t g = ga + (uintptr_t) c * 16;
        if (g >= lo && g + 16 <= hi) {
          asm volatile ("cp.async.cg.shared.global [%0], [%1], 16;"
              :: "r" ((uint32_t) __cvta_generic_to_shared (dstk + c * 16)), "l" (g) : "memory");
        } else {
          for (int x = 0; x < 16; x++)
            if (g + x >= lo && g + x < hi)
              dstk[c * 16 + x] = *reinterpret_cast<const unsigned char *> (g + x);
        }
      }
    }
    asm volatile ("cp.async.commit_group;" ::: "memory");
    asm volatile ("cp.async.wait_group 0;" ::: "memory");
    __syncthreads ();
    /* ---- gather ---- */
    if (worker) {
      const unsigned char *base[16];
#pragma unroll
      for (int i = 0; i < 16; i++)
        base[i] = stage + region[src[i]] + off[i];
      uint64_t o = b * batch_bytes + (uint64_t) tid * 16;
#pragma unroll
      for (int j = 0; j < MERGE_PHASE_J; j++) {
        if (o + 16 <= total) {
          uint32_t w[4] = { 0, 0, 0, 0 };
#pragma unroll
          for (int i = 0; i < 16; i++)
            w[i >> 2] |= (uint32_t) base[i][j * adv[i]] << (8 * (i & 3));
          st_stream<16> (dst + o, w);
        } else if (o < total) {
#pragma unroll
          for (int i = 0; i < 16; i++)
            if (o + i < total)
              dst[o + i] = base[i][j * adv[i]];
        }
        o += (uint64_t) threads_used * 16;
      }
    }
    __syncthreads ();
  }
}

static bool
launch_merge_phase (int n, const void *const *in, const uint64_t *block_bytes, void *out,
    uint64_t rows, const LaunchCfg &cfg, cudaStream_t stream, int *err)
{
  uint64_t S = 0, smallest = ~0ull;
  for (int k = 0; k < n; k++) {
    S += block_bytes[k];
    if (block_bytes[k] < smallest)
      smallest = block_bytes[k];
  }
  if (smallest >= 16 || (uintptr_t) out % 16 || getenv ("NNSB_MERGE_NO_PHASE"))
    return false;
  uint64_t a = S, b = 16;
  while (b) { const uint64_t t = a % b; a = b; b = t; }
  const uint64_t P = S / a * 16;
  if (P > MERGE_PERIOD_MAX)
    return false;
  MergePhaseDev m;
  memset (&m, 0, sizeof (m));
  m.n = n;
  m.chunks_per_period = (uint32_t) (P / 16);
  const uint64_t rows_per_period = P / S;
  for (int k = 0; k < n; k++) {
    m.in[k] = (const unsigned char *) in[k];
    m.in_bytes[k] = block_bytes[k] * rows;
    m.run[k] = (uint32_t) (rows_per_period * block_bytes[k]);
  }
  for (uint64_t e = 0; e < P; e++) {
    const uint64_t r = e / S;
    uint64_t c = e % S;
    int k = 0;
    while (c >= block_bytes[k]) {
      c -= block_bytes[k];
      k++;
    }
    m.src[e] = (uint8_t) k;
    m.delta[e] = (uint16_t) (r * block_bytes[k] + c);
  }
  const uint64_t total = S * rows;
  const uint32_t cpp = m.chunks_per_period;
  const uint32_t threads_used = STRIP_THREADS / cpp * cpp;
  const uint64_t batch_bytes = (uint64_t) threads_used * MERGE_PHASE_J * 16;
  const uint64_t n_batches = (total + batch_bytes - 1) / batch_bytes;
  static int configured[64];
  int dev = 0;
  cudaGetDevice (&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaFuncSetAttribute (merge_phase_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, MERGE_PHASE_SMEM);
    configured[dev] = 1;
  }
  uint64_t blocks = n_batches;
  const uint64_t cap = (uint64_t) cfg.sm_count * (cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : 8);
  if (blocks > cap)
    blocks = cap;
  merge_phase_kernel<<<(unsigned int) blocks, STRIP_THREADS, MERGE_PHASE_SMEM, stream>>> (
      (unsigned char *) out, m, total, threads_used, n_batches);
  *err = (int) cudaGetLastError ();
  return true;
}

int launch_merge (int n, const void *const *in, const uint64_t *block_bytes, void *out,
    uint64_t rows, const LaunchCfg &cfg, cudaStream_t stream)
{
  if (n < 1 || n > NNSB_MERGE_MAX)
    return (int) cudaErrorInvalidValue;
  if (rows == 0)
    return 0;
  MergeDev m;
  memset (&m, 0, sizeof (m));
  m.n = n;
  uint64_t S = 0;
  bool words = (uintptr_t) out % 16 == 0;
  for (int k = 0; k < n; k++) {
    if (block_bytes[k] == 0 || block_bytes[k] > 0x7fffffffull)
      return (int) cudaErrorInvalidValue;
    m.in[k] = (const unsigned char *) in[k];
    m.blk[k] = (uint32_t) block_bytes[k];
    m.off[k] = (uint32_t) S;
    m.in_bytes[k] = block_bytes[k] * rows;
    S += block_bytes[k];
    words = words && (uintptr_t) in[k] % 4 == 0;
  }
  if (S > 0x7fffffffull)
    return (int) cudaErrorInvalidValue;
  m.off[n] = (uint32_t) S;
  if (n == 1)
    return launch_copy (in[0], out, S * rows, cfg, stream);
  {
    int err = 0;
    if (launch_merge_phase (n, in, block_bytes, out, rows, cfg, stream, &err))
      return err;
  }
  const uint64_t total = S * rows;
  const uint64_t work = words ? (total + 15) / 16 : total;
  uint64_t blocks = (work + STRIP_THREADS - 1) / STRIP_THREADS;
  const uint64_t cap = (uint64_t) cfg.sm_count * (cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : 64);
  if (blocks > cap)
    blocks = cap;
  if (blocks == 0)
    blocks = 1;
  const uint64_t step = blocks * STRIP_THREADS * 16;
  if (words)
    merge_kernel<true><<<(unsigned int) blocks, STRIP_THREADS, 0, stream>>> ((unsigned char *) out, m,
        total, (uint32_t) (step / S), (uint32_t) (step % S));
  else
    merge_kernel<false><<<(unsigned int) blocks, STRIP_THREADS, 0, stream>>> ((unsigned char *) out, m,
        total, 0, 0);
  return (int) cudaGetLastError ();
}

}  // namespace nnsb
