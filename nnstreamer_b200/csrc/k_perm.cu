/*
 * k_perm.cu -- K3/K4/K8: data-movement kernels.
 *
 *  permute : rank-N permutation copy through a padded shared-memory tile.  Replaces the
 *            per-element nns_memcpy nest of gst_tensor_transform_transpose
 *            (gsttensor_transform.c:1417-1500) and the block memcpy nest of
 *            gst_tensor_transform_dimchg (:1155-1178).  Both lower to the same problem after
 *            the plan canonicalises the dimensions (drop extent-1 dims, merge dims that are
 *            adjacent on both sides):
 *               I : innermost run, contiguous in input AND output (extent R, may be 1)
 *               X : the next output-contiguous dimension (out stride R)
 *               Y : the next input-contiguous dimension  (in  stride R)
 *            A block stages an RT x TX x TY tile: global reads walk (r, y, x) so that a warp
 *            reads dense runs of the input, global writes walk (r, x, y) so that a warp
 *            writes dense runs of the output; the shared tile has an odd pitch, so the
 *            transposed access is bank-conflict free.  HWC->CHW on video (C = 3 innermost)
 *            becomes Y = C (3), X = W*H: the whole input tile is one contiguous run.
 *  copy    : flat vector copy for permutations that canonicalise to the identity
 *            (dimchg from == to, :1133-1139; transpose 0:1:2:3, :1464-1469).
 *  strip   : tensor_converter's video row-padding removal (gsttensor_converter.c:1087-1100):
 *            rows of row_bytes at pitch RU4 (row_bytes) -> dense.  Rows are 1..3 bytes off
 *            word alignment, so each thread assembles aligned 16-byte output chunks from
 *            aligned 4-byte source words with funnel shifts.
 *
 * Algorithmic bytes: 2 * es per element (permute, copy), 2 * row_bytes per row (strip).
 */
#include <cstdlib>

#include "nnsb_device.cuh"
#include "nnsb_kernels.h"

namespace nnsb {

constexpr int PERM_THREADS = 256;
constexpr int PERM_MAX_OUTER = NNSB_PERM_MAX_RANK;

struct PermDev
{
  /* tile dims */
  uint32_t R, RT, nR;           /* inner run: extent, tile, tiles */
  uint32_t SX, TX, nX;
  uint32_t SY, TY, nY;
  uint64_t x_in_stride;         /* input stride of X (elements) */
  uint64_t y_out_stride;        /* output stride of Y (elements) */
  uint32_t pitch;               /* shared pitch along x, in (y) units: odd */
  uint32_t magic_rt, magic_ty, magic_tx;   /* ceil (2^32 / d) for the in-tile index split */
  /* outer dims (everything except I, X, Y; the batch is one of them) */
  int n_outer;
  uint32_t o_size[PERM_MAX_OUTER];
  uint64_t o_in[PERM_MAX_OUTER];
  uint64_t o_out[PERM_MAX_OUTER];
  uint64_t n_tiles;
};

__device__ __forceinline__ uint32_t
fast_div (uint32_t q, uint32_t d, uint32_t magic)
{
  /* exact for q < 2^32 / d; q is an in-tile index (< 64 Ki) */
  return d == 1 ? q : __umulhi (q, magic);
}

template <typename E>
__global__ void __launch_bounds__ (PERM_THREADS)
permute_kernel (const E *__restrict__ in, E *__restrict__ out, const __grid_constant__ PermDev p)
{
  extern __shared__ __align__ (16) unsigned char smem_raw[];
  E *tile = reinterpret_cast<E *> (smem_raw);

  for (uint64_t t = blockIdx.x; t < p.n_tiles; t += gridDim.x) {
    /* decode tile index: r-tile fastest, then x-tile, then y-tile, then outer dims */
    uint64_t q = t;
    const uint32_t tr = (uint32_t) (q % p.nR); q /= p.nR;
    const uint32_t tx = (uint32_t) (q % p.nX); q /= p.nX;
    const uint32_t ty = (uint32_t) (q % p.nY); q /= p.nY;
    uint64_t in_base = 0, out_base = 0;
    for (int i = 0; i < p.n_outer; i++) {
      const uint64_t idx = q % p.o_size[i];
      q /= p.o_size[i];
      in_base += idx * p.o_in[i];
      out_base += idx * p.o_out[i];
    }
    const uint32_t r0 = tr * p.RT, x0 = tx * p.TX, y0 = ty * p.TY;
    const uint32_t rn = min (p.RT, p.R - r0), xn = min (p.TX, p.SX - x0), yn = min (p.TY, p.SY - y0);
    in_base += (uint64_t) x0 * p.x_in_stride + (uint64_t) y0 * p.R + r0;
    out_base += (uint64_t) y0 * p.y_out_stride + (uint64_t) x0 * p.R + r0;

    const uint32_t count = p.RT * p.TX * p.TY;
    /* global -> shared: (r, y, x), r fastest */
    for (uint32_t i = threadIdx.x; i < count; i += PERM_THREADS) {
      const uint32_t a = fast_div (i, p.RT, p.magic_rt);
      const uint32_t r = i - a * p.RT;
      const uint32_t x = fast_div (a, p.TY, p.magic_ty);
      const uint32_t y = a - x * p.TY;
      if (r < rn && y < yn && x < xn) {
        E v;
        ld_stream<sizeof (E)> (&v, in + in_base + (uint64_t) x * p.x_in_stride + (uint64_t) y * p.R + r);
        tile[((size_t) x * p.pitch + y) * p.RT + r] = v;
      }
    }
    __syncthreads ();
    /* shared -> global: (r, x, y), r fastest */
    for (uint32_t i = threadIdx.x; i < count; i += PERM_THREADS) {
      const uint32_t a = fast_div (i, p.RT, p.magic_rt);
      const uint32_t r = i - a * p.RT;
      const uint32_t y = fast_div (a, p.TX, p.magic_tx);
      const uint32_t x = a - y * p.TX;
      if (r < rn && y < yn && x < xn) {
        E v = tile[((size_t) x * p.pitch + y) * p.RT + r];
        st_stream<sizeof (E)> (out + out_base + (uint64_t) y * p.y_out_stride + (uint64_t) x * p.R + r, &v);
      }
    }
    __syncthreads ();
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

static uint32_t
magic_of (uint32_t d)
{
  if (d <= 1)
    return 0;
  return (uint32_t) (((1ull << 32) + d - 1) / d);
}

struct uint128_e { uint32_t v[4]; };

template <typename E> static int
launch_permute_t (const PermDev &p, size_t smem, const void *in, void *out, unsigned int grid,
    cudaStream_t stream)
{
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute (permute_kernel<E>,
        cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e != cudaSuccess)
      return (int) e;
  }
  permute_kernel<E><<<grid, PERM_THREADS, smem, stream>>> ((const E *) in, (E *) out, p);
  return (int) cudaGetLastError ();
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

int launch_permute (const PermDesc &d0, const void *in, void *out, size_t frame_elems,
    uint32_t batch, const LaunchCfg &cfg, cudaStream_t stream)
{
  if (batch == 0 || frame_elems == 0)
    return 0;
  /* 1. collect (size, in_stride, out_stride), drop extent-1 dims, append the batch */
  struct Dim { uint64_t size, in, out; };
  Dim dims[NNSB_PERM_MAX_RANK + 1];
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
  if (batch > 1) {
    dims[n].size = batch;
    dims[n].in = frame_elems;
    dims[n].out = frame_elems;
    n++;
  }
  /* 2. merge neighbours that are adjacent on both sides */
  int m = 0;
  for (int i = 0; i < n; i++) {
    if (m > 0 && dims[m - 1].in * dims[m - 1].size == dims[i].in
        && dims[m - 1].out * dims[m - 1].size == dims[i].out) {
      dims[m - 1].size *= dims[i].size;
    } else {
      dims[m++] = dims[i];
    }
  }
  n = m;
  const size_t total = (size_t) frame_elems * batch;
  if (n == 0 || (n == 1 && dims[0].in == 1))
    return launch_copy (in, out, total * d0.es, cfg, stream);

  /* 3. pick I, X, Y */
  PermDev p;
  memset (&p, 0, sizeof (p));
  int ix = -1, iy = -1, ii = -1;
  uint64_t R = 1;
  if (dims[0].in == 1) {
    ii = 0;
    R = dims[0].size;
  }
  for (int i = 0; i < n; i++) {
    if (i == ii)
      continue;
    if (ix < 0 && dims[i].out == R)
      ix = i;
    if (iy < 0 && dims[i].in == R)
      iy = i;
  }
  if (ix < 0 || iy < 0 || ix == iy || R > 0xffffffffull || dims[ix].size > 0xffffffffull
      || dims[iy].size > 0xffffffffull)
    return (int) cudaErrorInvalidValue;   /* cannot happen for a canonical permutation */
  p.R = (uint32_t) R;
  p.SX = (uint32_t) dims[ix].size;
  p.SY = (uint32_t) dims[iy].size;
  p.x_in_stride = dims[ix].in;
  p.y_out_stride = dims[iy].out;
  for (int i = 0; i < n; i++) {
    if (i == ii || i == ix || i == iy)
      continue;
    p.o_size[p.n_outer] = (uint32_t) dims[i].size;
    p.o_in[p.n_outer] = dims[i].in;
    p.o_out[p.n_outer] = dims[i].out;
    p.n_outer++;
  }

  /* 3b. video-shaped cases go to the TMA pipeline (k_plane.cu) when they qualify */
  if (p.R == 1 && !getenv ("NNSB_NO_PLANE")) {
    int r = -1;
    if (p.SY >= 2 && p.SY <= 4 && p.x_in_stride == p.SY)
      /* [P][C] -> [C][P]: Y = channels (input-contiguous), X = pixels */
      r = launch_plane (true, (int) p.SY, d0.es, p.SX, p.y_out_stride, p.n_outer, p.o_size,
          p.o_in, p.o_out, in, out, cfg, stream);
    else if (p.SX >= 2 && p.SX <= 4 && p.y_out_stride == p.SX)
      /* [C][P] -> [P][C]: X = channels (output-contiguous), Y = pixels */
      r = launch_plane (false, (int) p.SX, d0.es, p.SY, p.x_in_stride, p.n_outer, p.o_size,
          p.o_out, p.o_in, in, out, cfg, stream);
    if (r != -1)
      return r;
  }

  /* 4. tile shape: ~8 KiB of payload, runs of >= 32 elements on both global sides */
  const uint32_t budget = (uint32_t) (d0.es <= 4 ? 2048 : 8192 / d0.es);
  p.RT = p.R < budget ? p.R : budget;
  uint32_t rem = budget / p.RT;
  if (rem < 1)
    rem = 1;
  const uint32_t run = (32 + p.RT - 1) / p.RT;       /* tiles of the other dims per 32-element run */
  p.TY = p.SY < run ? p.SY : run;
  if (p.TY > rem)
    p.TY = rem;
  if (p.TY < 1)
    p.TY = 1;
  uint32_t tx = rem / p.TY;
  if (tx < 1)
    tx = 1;
  if (tx > 32)
    tx = (tx / 32) * 32;
  p.TX = p.SX < tx ? p.SX : tx;
  /* give unused budget back to Y (e.g. small X) */
  if ((uint64_t) p.RT * p.TX * p.TY * 2 <= budget) {
    uint32_t ty = budget / (p.RT * p.TX);
    p.TY = p.SY < ty ? p.SY : ty;
  }
  p.nR = (p.R + p.RT - 1) / p.RT;
  p.nX = (p.SX + p.TX - 1) / p.TX;
  p.nY = (p.SY + p.TY - 1) / p.TY;
  p.pitch = p.TY | 1u;
  p.magic_rt = magic_of (p.RT);
  p.magic_ty = magic_of (p.TY);
  p.magic_tx = magic_of (p.TX);
  uint64_t tiles = (uint64_t) p.nR * p.nX * p.nY;
  for (int i = 0; i < p.n_outer; i++)
    tiles *= p.o_size[i];
  p.n_tiles = tiles;
  const size_t smem = (size_t) p.TX * p.pitch * p.RT * d0.es;

  uint64_t grid = tiles;
  const uint64_t cap = (uint64_t) cfg.sm_count * (cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : 16);
  if (grid > cap)
    grid = cap;

  switch (d0.es) {
    case 1: return launch_permute_t<uint8_t> (p, smem, in, out, (unsigned int) grid, stream);
    case 2: return launch_permute_t<uint16_t> (p, smem, in, out, (unsigned int) grid, stream);
    case 4: return launch_permute_t<uint32_t> (p, smem, in, out, (unsigned int) grid, stream);
    case 8: return launch_permute_t<Bytes<8> > (p, smem, in, out, (unsigned int) grid, stream);
    case 16: return launch_permute_t<Bytes<16> > (p, smem, in, out, (unsigned int) grid, stream);
    default: return (int) cudaErrorInvalidValue;
  }
}

/* ------------------------------------------------------------------------------------- */
/* strip                                                                                   */
/* ------------------------------------------------------------------------------------- */
constexpr int STRIP_THREADS = 256;

/* 4 source bytes starting at byte offset `s` (any alignment) from word-aligned memory */
__device__ __forceinline__ uint32_t
load_unaligned_u32 (const unsigned char *base, uint64_t s)
{
  const uint32_t *w = reinterpret_cast<const uint32_t *> (base + (s & ~3ull));
  const uint32_t sh = (uint32_t) (s & 3) * 8;
  const uint32_t lo = __ldg (w);
  if (sh == 0)
    return lo;
  const uint32_t hi = __ldg (w + 1);
  return __funnelshift_r (lo, hi, sh);
}

/* word path: dst 16-byte aligned, src 4-byte aligned, pitch % 4 == 0, row_bytes >= 4 */
__global__ void __launch_bounds__ (STRIP_THREADS)
strip_kernel_w (const unsigned char *__restrict__ src, unsigned char *__restrict__ dst,
    uint32_t row_bytes, uint64_t pitch, uint64_t total, uint64_t src_bytes)
{
  const uint64_t n_chunks = total / 16;
  const uint64_t tid = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t nthreads = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = tid; i < n_chunks; i += nthreads) {
    const uint64_t o = i * 16;
    uint64_t row = o / row_bytes;
    uint32_t col = (uint32_t) (o - row * row_bytes);
    uint32_t w[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      if (col + 4 <= row_bytes) {
        const uint64_t s = row * pitch + col;
        /* the second word of an unaligned read may be the padding after the last row */
        if (s + 8 <= src_bytes || (s & 3) == 0) {
          w[k] = load_unaligned_u32 (src, s);
        } else {
          w[k] = (uint32_t) src[s] | ((uint32_t) src[s + 1] << 8) | ((uint32_t) src[s + 2] << 16)
              | ((uint32_t) src[s + 3] << 24);
        }
        col += 4;
        if (col == row_bytes) {
          col = 0;
          row++;
        }
      } else {
        /* the word straddles a row boundary: byte by byte */
        uint32_t v = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) {
          v |= (uint32_t) src[row * pitch + col] << (8 * b);
          if (++col == row_bytes) {
            col = 0;
            row++;
          }
        }
        w[k] = v;
      }
    }
    st_stream<16> (dst + o, w);
  }
  /* last (total % 16) bytes */
  for (uint64_t b = n_chunks * 16 + tid; b < total; b += nthreads) {
    const uint64_t row = b / row_bytes;
    dst[b] = src[row * pitch + (b - row * row_bytes)];
  }
}

/* byte path: any alignment */
__global__ void __launch_bounds__ (STRIP_THREADS)
strip_kernel_b (const unsigned char *__restrict__ src, unsigned char *__restrict__ dst,
    uint64_t row_bytes, uint64_t pitch, uint64_t total)
{
  const uint64_t tid = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t nthreads = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t b = tid; b < total; b += nthreads) {
    const uint64_t row = b / row_bytes;
    dst[b] = src[row * pitch + (b - row * row_bytes)];
  }
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
  const uint64_t total = (uint64_t) row_bytes * rows;
  const uint64_t src_bytes = (uint64_t) src_pitch * (rows - 1) + row_bytes;
  const bool word_ok = ((uintptr_t) dst % 16 == 0) && ((uintptr_t) src % 4 == 0)
      && (src_pitch % 4 == 0) && row_bytes >= 4 && row_bytes <= 0xffffffffull;
  const uint64_t work = word_ok ? (total + 15) / 16 : total;
  uint64_t blocks = (work + STRIP_THREADS - 1) / STRIP_THREADS;
  const uint64_t cap = (uint64_t) cfg.sm_count * (cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : 8);
  if (blocks > cap)
    blocks = cap;
  if (blocks == 0)
    blocks = 1;
  if (word_ok)
    strip_kernel_w<<<(unsigned int) blocks, STRIP_THREADS, 0, stream>>> ((const unsigned char *) src,
        (unsigned char *) dst, (uint32_t) row_bytes, src_pitch, total, src_bytes);
  else
    strip_kernel_b<<<(unsigned int) blocks, STRIP_THREADS, 0, stream>>> ((const unsigned char *) src,
        (unsigned char *) dst, row_bytes, src_pitch, total);
  return (int) cudaGetLastError ();
}

}  // namespace nnsb
