/*
 * k_stand.cu -- type dispatch and launch sequencing for stand (see k_stand.cuh).
 */
#include "k_stand.cuh"

namespace nnsb {

#define NNSB_DECL(i) \
  stand_stats_fn stand_stats_row_##i (void); \
  stand_apply_fn stand_apply_row_##i (int);
NNSB_DECL (0) NNSB_DECL (1) NNSB_DECL (2) NNSB_DECL (3) NNSB_DECL (4) NNSB_DECL (5)
NNSB_DECL (6) NNSB_DECL (7) NNSB_DECL (8) NNSB_DECL (9) NNSB_DECL (10)
#undef NNSB_DECL

size_t stand_scratch_doubles (uint32_t batch, uint32_t nch, size_t frame_elems,
    const LaunchCfg &cfg)
{
  const StandGeom g = stand_geometry (batch, nch, frame_elems, cfg);
  return (size_t) g.batch * g.blocks_x * g.nch * 2;
}

int launch_stand (int in_type, int out_type, int stand_mode, const void *in, void *out,
    size_t frame_elems, uint32_t nch, uint32_t batch, double *scratch, double *stats,
    const LaunchCfg &cfg, cudaStream_t stream, int *n_launches)
{
  static stand_stats_fn (*const srows[NNSB_TYPE_END]) (void) = {
    stand_stats_row_0, stand_stats_row_1, stand_stats_row_2, stand_stats_row_3,
    stand_stats_row_4, stand_stats_row_5, stand_stats_row_6, stand_stats_row_7,
    stand_stats_row_8, stand_stats_row_9, stand_stats_row_10
  };
  static stand_apply_fn (*const arows[NNSB_TYPE_END]) (int) = {
    stand_apply_row_0, stand_apply_row_1, stand_apply_row_2, stand_apply_row_3,
    stand_apply_row_4, stand_apply_row_5, stand_apply_row_6, stand_apply_row_7,
    stand_apply_row_8, stand_apply_row_9, stand_apply_row_10
  };
  if (n_launches)
    *n_launches = 0;
  if (batch == 0 || frame_elems == 0)
    return 0;
  if (in_type < 0 || in_type >= NNSB_TYPE_END || nch == 0 || frame_elems % nch != 0)
    return (int) cudaErrorInvalidValue;
  stand_stats_fn sfn = srows[in_type] ();
  stand_apply_fn afn = arows[in_type] (out_type);
  if (!sfn || !afn)
    return (int) cudaErrorInvalidValue;

  const StandGeom g = stand_geometry (batch, nch, frame_elems, cfg);
  const int need_std = (stand_mode == NNSB_STAND_DEFAULT);

  /* statistics grid: blocks_x per tensor, except the column walk which spreads the
   * channels over as many blocks as needed */
  uint32_t stats_grid = g.blocks_x;
  if (g.path == 2) {
    uint64_t b = ((uint64_t) nch + STAND_THREADS - 1) / STAND_THREADS;
    const uint64_t cap = (uint64_t) cfg.sm_count * 8;
    stats_grid = (uint32_t) (b > cap ? cap : b);
  }
  int err = sfn (in, frame_elems, g, stats_grid, scratch, stats, need_std, stream);
  if (err)
    return err;

  /* apply grid: fill the machine; for the strided paths it must own at least nch threads */
  const int bps = cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : (batch > 1 ? 64 : 32);   /* as stand_geometry */
  uint64_t ax = ((uint64_t) cfg.sm_count * bps) / batch;
  if (ax < 8)
    ax = 8;
  uint64_t need = (frame_elems + (uint64_t) STAND_THREADS * 16 - 1) / ((uint64_t) STAND_THREADS * 16);
  if (need < 1)
    need = 1;
  if (ax > need)
    ax = need;
  if (g.path != 2) {
    const uint64_t min_blocks = ((uint64_t) nch + STAND_THREADS - 1) / STAND_THREADS;
    if (ax < min_blocks)
      ax = min_blocks;
  }
  err = afn (in, out, frame_elems, g, (uint32_t) ax, stats, need_std, stream);
  if (n_launches)
    *n_launches = 3;
  return err;
}

}  // namespace nnsb
