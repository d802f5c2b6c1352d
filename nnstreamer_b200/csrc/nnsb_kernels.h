/*
 * nnsb_kernels.h -- host-callable launchers of the sm_100a kernels (internal).
 * Every launcher enqueues on `stream`, never synchronises, never allocates, and returns the
 * cudaError_t of the launch as int (0 = ok).
 */
#ifndef NNSB_KERNELS_H
#define NNSB_KERNELS_H

#include <cuda_runtime.h>
#include "nnsb_internal.h"

namespace nnsb {

struct LaunchCfg
{
  int sm_count;                 /* SMs of the device (148 on B200) */
  int blocks_per_sm;            /* grid = sm_count * blocks_per_sm (0 = kernel default) */
};

/* Blocks of `kernel` that are resident on one SM at the same time.  Grid-stride kernels are
 * launched with exactly sm_count * resident blocks: one full wave, no tail wave (a launch
 * with 1.6 waves leaves 40 % of the SMs idle during the second one). */
template <typename K> static inline int
resident_blocks_per_sm (K kernel, int threads, size_t dyn_smem)
{
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor (&n, kernel, threads, dyn_smem) != cudaSuccess
      || n < 1) {
    cudaGetLastError ();
    n = 1;
  }
  return n;
}

/* K1/K2: fused (typecast +) arithmetic chain.  typecast mode is the empty chain.
 * Replaces gst_tensor_transform_typecast (:1202) and _arithmetic (:1240). */
struct ChainArgs
{
  nnsb_chain_desc desc;
  const nnsb_cell *table;       /* device, [n_pos][nch]; used when nch > 1 */
  int fast_f32;                 /* float32 chain whose divisors all carry a reciprocal */
};
int launch_chain (int in_type, int out_type, bool orc, const void *in, void *out, size_t n,
    const ChainArgs & args, const LaunchCfg & cfg, cudaStream_t stream);

/* K5: clamp, gst_tensor_transform_clamp (:1632) */
int launch_clamp (int type, const void *in, void *out, size_t n, double lo, double hi,
    const LaunchCfg & cfg, cudaStream_t stream);

/* K6/K7: stand, gst_tensor_transform_stand (:1512) + tensor_data.c:314-493.
 * Per tensor f (of `batch`): statistics over frame_elems elements, per channel c = e % nch
 * (nch = 1 when not per-channel).
 *   scratch: device doubles, stand_scratch_doubles() of them, contents undefined on entry
 *   stats:   device doubles [batch][nch][2] = {mean, std}; written by the launch */
size_t stand_scratch_doubles (uint32_t batch, uint32_t nch, size_t frame_elems,
    const LaunchCfg & cfg);
int launch_stand (int in_type, int out_type, int stand_mode, const void *in, void *out,
    size_t frame_elems, uint32_t nch, uint32_t batch, double *scratch, double *stats,
    const LaunchCfg & cfg, cudaStream_t stream, int *n_launches);

/* Single-read stand for batches of float32 tensors whose per-SM slice fits in shared memory
 * (k_stand_fused.cu): one cooperative, persistent launch instead of three.  `scratch`:
 * stand_fused_scratch_bytes() of device memory (contents undefined); `stats`: batch * nch * 2
 * doubles {mean, std}. */
bool stand_fused_supported (int in_type, int out_type, const void *in, const void *out,
    size_t frame_elems, uint32_t nch, uint32_t batch, const LaunchCfg &cfg);
size_t stand_fused_scratch_bytes (const LaunchCfg &cfg);
int launch_stand_fused (int stand_mode, const void *in, void *out, size_t frame_elems, uint32_t nch,
    uint32_t batch, void *scratch, double *stats, const LaunchCfg &cfg, cudaStream_t stream,
    int *n_launches);

/* K3/K4: rank-N permutation copy (transpose :1443, dimchg :1120).
 * out index (innermost first) i has extent size[i] and input stride in_stride[i] (elements
 * of es bytes); the output is dense.  `batch` tensors of frame_elems elements each. */
#define NNSB_PERM_MAX_RANK 8
struct PermDesc
{
  int rank;
  uint32_t size[NNSB_PERM_MAX_RANK];
  uint64_t in_stride[NNSB_PERM_MAX_RANK];
  int es;                       /* element bytes: 1,2,4,8,16 */
};
/* K3/K4 vector path (k_perm_vec.cu): 16-byte accesses on both sides, driven by a table that
 * the host builds once per plan (per-vector global offsets and swizzled shared-memory slots
 * for one box) and the plan keeps in device memory.  4-, 8- and 16-byte elements (after an
 * innermost block common to both sides has been folded into the element).
 *   perm_vec_build : 0 = eligible, *host_table is malloc()ed (table_bytes long); -1 = not
 *   launch_permute_vec : -1 when this call does not qualify (unaligned pointers, too many
 *                        tiles), else the launch status */
struct PermVecPlan
{
  int valid;
  int es;                                       /* element bytes after folding (1 .. 16) */
  int vb;                                       /* vector bytes: 16, 8 or 4 */
  uint64_t frame_bytes;
  int n;                                        /* canonical dims, output order, no batch */
  uint32_t size[NNSB_PERM_MAX_RANK];
  uint32_t tile[NNSB_PERM_MAX_RANK];            /* box extent per dim */
  uint64_t in_stride[NNSB_PERM_MAX_RANK];       /* elements */
  uint64_t out_stride[NNSB_PERM_MAX_RANK];
  int box_slot[NNSB_PERM_MAX_RANK];             /* field of the packed coordinates, -1 = extent 1 */
  uint64_t limits_init;                         /* packed limits of the unused fields (1 each) */
  uint32_t n_vec;                               /* 16-byte vectors per box */
  uint64_t tiles_per_frame;
  size_t table_bytes;
  int swizzle_shift, swizzle_shift2;            /* chosen shared-memory swizzle (-1 = none) */
  int micro;                                    /* store side uses VE x VE register micro-transposes */
  uint64_t gather_wavefronts;                   /* simulated LDS wavefronts per box (ideal:
                                                 * one per 128 bytes gathered) */
};
int perm_vec_build (const PermDesc & d, size_t frame_elems, PermVecPlan * plan, void **host_table);
int launch_permute_vec (const PermVecPlan & plan, const void *d_table, const void *in, void *out,
    size_t frame_elems, uint32_t batch, const LaunchCfg & cfg, cudaStream_t stream);

/* vec/d_table: optional vector plan, tried after the interleaved<->planar fast path */
int launch_permute (const PermDesc & d, const void *in, void *out, size_t frame_elems,
    uint32_t batch, const LaunchCfg & cfg, cudaStream_t stream,
    const PermVecPlan * vec = nullptr, const void *d_table = nullptr);

/* "slab" / "vector" / "scalar" / "copy": the kernel a batch of such tensors takes when the
 * interleaved <-> planar fast path does not apply (for nnsb_plan_describe) */
const char *permute_path_name (const PermDesc & d, size_t frame_elems, uint32_t batch,
    const PermVecPlan * vec);

/* K3 fast path (k_plane.cu): interleaved <-> planar through TMA bulk copies.
 * Returns -1 when the problem does not qualify, else the launch status. */
int launch_plane (bool to_planar, int channels, int es, uint64_t pixels, uint64_t plane_stride,
    int n_outer, const uint32_t *o_size, const uint64_t *o_il, const uint64_t *o_pl,
    const void *in, void *out, const LaunchCfg & cfg, cudaStream_t stream);

/* K8: row-stride strip, gsttensor_converter.c:1087-1100 */
int launch_strip (const void *src, void *dst, size_t row_bytes, size_t src_pitch, size_t rows,
    const LaunchCfg & cfg, cudaStream_t stream);

/* K9: row padding, the inverse of K8 (tensordec-directvideo.c:326-339); the dst_pitch -
 * row_bytes padding bytes of every row are written as zero */
int launch_pad (const void *src, void *dst, size_t row_bytes, size_t dst_pitch, size_t rows,
    const LaunchCfg & cfg, cudaStream_t stream);

/* K10: tensor_merge mode=linear (gsttensor_merge.c:506-586): output row r is the concatenation
 * of block r of every input; block_bytes[k] = element size * prod (dim_k[0..axis]) */
#define NNSB_MERGE_MAX 16
int launch_merge (int n, const void *const *in, const uint64_t *block_bytes, void *out,
    uint64_t rows, const LaunchCfg & cfg, cudaStream_t stream);

}  // namespace nnsb
#endif
