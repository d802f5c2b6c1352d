/*
 * k_perm_pix.cu -- K3 pixel transpose: `transpose 0:2:1:3` of packed RGB / BGR bytes
 * ([3][W][H] -> [3][H][W]: the image is transposed, the three bytes of a pixel stay together).
 *
 * Same job as the other permute kernels (gst_tensor_transform_transpose,
 * gsttensor_transform.c:1417-1500), for THE video shape that suits none of them: the moved unit
 * is 3 bytes, so the canonicaliser cannot fold it into a power-of-two element, rows of 1080
 * pixels are 3240 bytes (8-byte vectors only), and a byte-wise gather spends one shared-memory
 * load per byte (2.9 TB/s on 1080p).  Here a pixel travels through shared memory as ONE 32-bit
 * word:
 *   load    a tile of 64 x 64 pixels; a thread reads 16 pixels of one input row with three
 *           16-byte loads, widens them to RGBx words (PRMT) and scatters them TRANSPOSED into
 *           shared memory (tile_t[w][h], pitch 65 words, columns swizzled so that neither the
 *           scatter nor the gather below has bank conflicts within a row);
 *   store   an output row of the tile is 64 pixels = 192 bytes = 24 pieces of 8 bytes; a thread
 *           builds a piece from the 4 pixels it touches (4 ld.shared.b32, 3 PRMT to re-pack to 12
 *           bytes, two funnel shifts to cut the 8 it wants) and consecutive threads store
 *           consecutive pieces: fully coalesced 8-byte stores (output rows are only 8-byte
 *           aligned when H is odd times 8).
 * Requirements: W % 16 == 0 (16-byte aligned input rows), H % 8 == 0, 16-byte aligned frames.
 * Algorithmic bytes: 2 per byte moved.
 */
#include <cstdlib>
#include <cstring>

#include "nnsb_device.cuh"
#include "nnsb_kernels.h"
#include "k_perm_common.cuh"

namespace nnsb {

namespace {

constexpr int PIX_T = 64;               /* tile edge, pixels */
constexpr int PIX_PITCH = PIX_T + 1;    /* words */
constexpr int PIX_THREADS = 256;

template <bool STAGE, bool SWZ>
__global__ void __launch_bounds__ (PIX_THREADS)
pixel_transpose_kernel (const unsigned char *__restrict__ in, unsigned char *__restrict__ out,
    uint32_t W, uint32_t H, uint32_t tiles_w, uint32_t tiles_h, uint64_t frame_bytes)
{
  __shared__ uint32_t tile_t[PIX_T * PIX_PITCH];
  const uint32_t t = threadIdx.x;
  const uint32_t tw = blockIdx.x % tiles_w, th = blockIdx.x / tiles_w;
  const uint32_t w0 = tw * PIX_T, h0 = th * PIX_T;
  const unsigned char *fin = in + (uint64_t) blockIdx.y * frame_bytes;
  unsigned char *fout = out + (uint64_t) blockIdx.y * frame_bytes;

  /* ---- stage: the tile's 64 rows of 192 bytes, global -> shared without passing registers.
   * Consecutive lanes copy consecutive 16-byte pieces, so a warp's copy touches 4 lines; the
   * direct loads below (lane stride 48 bytes) touch 12 per instruction, three times. ---- */
  __shared__ __align__ (16) unsigned char stage[STAGE ? PIX_T * 192 : 16];
  if constexpr (STAGE) {
#pragma unroll
    for (int i = 0; i < 3; i++) {
      const uint32_t g = t + PIX_THREADS * i;
      const uint32_t row = g / 12, pc = g - row * 12;
      if (h0 + row < H && (w0 * 3 + 16 * pc) < W * 3)
        asm volatile ("cp.async.cg.shared.global [%0], [%1], 16;"
            :: "r" ((uint32_t) __cvta_generic_to_shared (stage + row * 192 + pc * 16)),
               "l" (fin + ((uint64_t) (h0 + row) * W + w0) * 3 + 16 * pc) : "memory");
    }
    asm volatile ("cp.async.commit_group;" ::: "memory");
    asm volatile ("cp.async.wait_group 0;" ::: "memory");
    __syncthreads ();
  }

  /* ---- load: thread = (input row h, 16-pixel quarter) ---- */
  {
    const uint32_t h = t >> 2, q = t & 3;
    const uint32_t gw = w0 + 16 * q, gh = h0 + h;
    if (gh < H && gw < W) {             /* W % 16 == 0: a quarter is inside or outside as a whole */
      const uint4 *src = STAGE ? reinterpret_cast<const uint4 *> (stage + h * 192 + q * 48)
          : reinterpret_cast<const uint4 *> (fin + ((uint64_t) gh * W + gw) * 3);
      /* column of pixel (row 16q + p, h): h with bit 3 flipped in the upper half of the rows, and
       * the upper half of the columns shifted by one word.  The four quarters of a warp's store
       * then fall into four different groups of 8 banks, and the 24 pieces of a row that a warp
       * reads below hit 24 different banks (columns c and c + 32 no longer share one). */
      const uint32_t hx = SWZ ? h ^ ((q & 2) << 2) : h;
      const uint32_t hcol = SWZ ? hx + (hx >> 5) : h;
      uint32_t v[12];
#pragma unroll
      for (int i = 0; i < 3; i++) {
        uint32_t w4[4];
        if constexpr (STAGE) {
          const uint4 v4 = src[i];
          w4[0] = v4.x;
          w4[1] = v4.y;
          w4[2] = v4.z;
          w4[3] = v4.w;
        } else {
          ld_stream<16> (w4, src + i);
        }
        v[4 * i] = w4[0];
        v[4 * i + 1] = w4[1];
        v[4 * i + 2] = w4[2];
        v[4 * i + 3] = w4[3];
      }
#pragma unroll
      for (int p = 0; p < 16; p++) {
        /* pixel p = bytes [3p, 3p + 3) of the 48: from words 3p/4 and 3p/4 + 1 */
        const int b = 3 * p, wi = b >> 2, sh = b & 3;
        const uint32_t lo = v[wi], hi = wi + 1 < 12 ? v[wi + 1] : 0u;
        const uint32_t px = __funnelshift_r (lo, hi, 8 * sh) & 0x00ffffffu;
        tile_t[(16 * q + p) * PIX_PITCH + hcol] = px;
      }
    }
  }
  __syncthreads ();

  /* ---- store: piece g = (output row w, 8-byte piece p of its 192 bytes) ---- */
  const uint32_t rows_here = (W - w0 < (uint32_t) PIX_T) ? W - w0 : (uint32_t) PIX_T;
  const uint32_t h_here = (H - h0 < (uint32_t) PIX_T) ? H - h0 : (uint32_t) PIX_T;       /* a multiple of 8 */
  const uint32_t pieces_per_row = h_here * 3 / 8;
#pragma unroll
  for (int j = 0; j < 6; j++) {
    const uint32_t g = t + PIX_THREADS * j;
    const uint32_t w = g / 24, p = g - w * 24;
    if (w < rows_here && p < pieces_per_row) {
      /* bytes [8p, 8p + 8) of the row: pixels px0 .. px0 + 3, starting r bytes into px0 */
      const uint32_t px0 = (8 * p) / 3, r = 8 * p - 3 * px0;
      const uint32_t *row = tile_t + w * PIX_PITCH;
      uint32_t a, b, c, d;
      if constexpr (SWZ) {
        const uint32_t x = (w & 32) >> 2;        /* the swizzle of the load phase: q & 2 == (w >> 4) & 2 */
        const uint32_t h0x = px0 ^ x, h1x = (px0 + 1) ^ x, h2x = (px0 + 2) ^ x, h3x = (px0 + 3) ^ x;
        a = row[h0x + (h0x >> 5)];
        b = row[h1x + (h1x >> 5)];
        c = row[h2x + (h2x >> 5)];
        d = px0 + 3 < (uint32_t) PIX_T ? row[h3x + (h3x >> 5)] : 0u;
      } else {
        a = row[px0];
        b = row[px0 + 1];
        c = row[px0 + 2];
        d = px0 + 3 < (uint32_t) PIX_T ? row[px0 + 3] : 0u;
      }
      /* re-pack 4 x RGBx to 12 bytes */
      const uint32_t x0 = a | (b << 24);
      const uint32_t x1 = (b >> 8) | (c << 16);
      const uint32_t x2 = (c >> 16) | (d << 8);
      uint32_t o2[2];
      o2[0] = __funnelshift_r (x0, x1, 8 * r);
      o2[1] = __funnelshift_r (x1, x2, 8 * r);
      st_stream<8> (fout + ((uint64_t) (w0 + w) * H + h0) * 3 + 8 * p, o2);
    }
  }
}

}  // namespace

/* dims: canonical permutation (output order).  -1 = not this kernel's shape. */
int
launch_pixel_transpose (const CDim *dims, int n, int es, const void *in, void *out, cudaStream_t stream,
    bool dry_run)
{
  if (es != 1 || (n != 3 && n != 4) || getenv ("NNSB_PERM_NO_PIX"))
    return -1;
  /* [3 : in 1, out 1] [H : in 3 W, out 3] [W : in 3, out 3 H] (+ batch) */
  if (dims[0].size != 3 || dims[0].in != 1 || dims[0].out != 1)
    return -1;
  const uint64_t H = dims[1].size, W = dims[2].size;
  if (dims[1].out != 3 || dims[2].in != 3 || dims[1].in != 3 * W || dims[2].out != 3 * H)
    return -1;
  const uint64_t frame = 3 * W * H;
  uint64_t batch = 1;
  if (n == 4) {
    if (dims[3].in != frame || dims[3].out != frame)
      return -1;
    batch = dims[3].size;
  }
  if (W % 16 || H % 8 || frame % 16 || batch > 65535 || W >= (1u << 24) || H >= (1u << 24))
    return -1;
  if (((uintptr_t) in | (uintptr_t) out) & 15)
    return -1;
  const uint32_t tiles_w = (uint32_t) ((W + PIX_T - 1) / PIX_T), tiles_h = (uint32_t) ((H + PIX_T - 1) / PIX_T);
  if ((uint64_t) tiles_w * tiles_h >= (1ull << 31))
    return -1;
  if (dry_run)
    return 0;
  dim3 grid (tiles_w * tiles_h, (unsigned int) batch);
  static const int variant = [] {
    const char *e = getenv ("NNSB_PIX_STAGE"), *z = getenv ("NNSB_PIX_SWIZZLE");
    return (e ? atoi (e) != 0 : 1) * 2 + (z ? atoi (z) != 0 : 0);
  } ();
  const unsigned char *pi = (const unsigned char *) in;
  unsigned char *po = (unsigned char *) out;
  const uint32_t w32 = (uint32_t) W, h32 = (uint32_t) H;
  switch (variant) {
    case 0: pixel_transpose_kernel<false, false><<<grid, PIX_THREADS, 0, stream>>> (pi, po, w32, h32, tiles_w, tiles_h, frame); break;
    case 1: pixel_transpose_kernel<false, true><<<grid, PIX_THREADS, 0, stream>>> (pi, po, w32, h32, tiles_w, tiles_h, frame); break;
    case 2: pixel_transpose_kernel<true, false><<<grid, PIX_THREADS, 0, stream>>> (pi, po, w32, h32, tiles_w, tiles_h, frame); break;
    default: pixel_transpose_kernel<true, true><<<grid, PIX_THREADS, 0, stream>>> (pi, po, w32, h32, tiles_w, tiles_h, frame); break;
  }
  return (int) cudaGetLastError ();
}

}  // namespace nnsb
