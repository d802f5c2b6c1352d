/*
 * k_stand_noalloc.cu -- stand (standardisation) for callers that must not allocate.
 *
 * nnsb_plan_launch() on a stand plan takes its scratch (block partial sums and the per-tensor
 * {mean, std} table) from the context's device pool, which calls cudaMalloc the first time
 * and whenever the batch grows.  Some users of this library cannot afford that call on the
 * path that processes frames:
 *   - a pipeline that shares the GPU with an inference runtime which has already reserved
 *     (almost) all device memory for its own pool: a late cudaMalloc may simply fail;
 *   - a streaming thread with a per-frame deadline: cudaMalloc takes driver-wide locks and
 *     may wait for other work on the device, so its latency is unbounded;
 *   - code that is only handed a CUDA stream and device pointers by its host application
 *     and owns no nnsb_ctx in which a pool could live.
 * For them this file offers ONE entry point that never calls cudaMalloc, cudaFree or any
 * other allocating routine: the scratch is the fixed array below, which is part of this
 * library's device image.  A GPU holds those NNSB_NOALLOC_SCRATCH_BYTES (1 MiB) from the
 * moment the image is loaded onto it until the process exits.
 *
 * When that moment is matters to exactly the users listed above.  CUDA loads device code
 * lazily by default (CUDA_MODULE_LOADING=LAZY since 12.2): the array is materialised by the
 * first cudaGetSymbolAddress on a device -- i.e. by the FIRST CALL of this entry point there --
 * and every kernel variant (input type x output type x channel-count class x mode) is loaded
 * by its first launch.  Those loads reserve device memory and take driver locks, which is
 * precisely what the frame path must avoid.  So: make one warm-up call per device and per
 * variant you will use during initialisation, before the other runtime reserves the rest of
 * the memory and before any stream capture; or start the process with
 * CUDA_MODULE_LOADING=EAGER, which moves all of it to library load time.  After that, calls
 * allocate nothing (tests/test_zz_no_alloc.py compares cudaMemGetInfo around warmed-up calls).
 *
 * The kernels are exactly those of the pooled path (k_stand.cuh, launch_stand).  With the
 * pooled path's launch geometry the results are bit-identical to nnsb_plan_launch(); when
 * that geometry's block partials do not fit the array the entry point uses a smaller grid
 * (four blocks per SM), i.e. another summation order, and the results then agree with
 * nnsb_plan_launch() only within stand's tolerance (both are checked against the CPU
 * reference in the tests).
 */
#include <cuda_runtime.h>

#include <mutex>

#include "nnsb_internal.h"
#include "nnsb_kernels.h"

/* The scratch itself.  NOTE for maintainers: in HOST code the name `g_noalloc_scratch` is only
 * a handle the runtime uses to find the symbol -- it is not a device address and must never
 * be passed to a kernel.  The device address is obtained once per device with
 * cudaGetSymbolAddress() below. */
static __device__ __align__ (256) unsigned char g_noalloc_scratch[NNSB_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;

struct PerDevice
{
  bool ready;
  double *scratch;              /* device address of g_noalloc_scratch on this device */
  int sm_count;
};

PerDevice g_dev[kMaxDevices];
std::mutex g_mu;

/* One-time set-up for the CURRENT device: resolve the device address of the array and read
 * the SM count.  Neither call allocates.  Runs on the first use of the entry point on a
 * device (under a lock; later calls only read the cached values). */
int
setup_current_device (int dev, PerDevice ** out)
{
  if (dev < 0 || dev >= kMaxDevices)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_stand_launch_noalloc: device index %d not supported",
        dev);
  std::lock_guard<std::mutex> lk (g_mu);
  PerDevice & pd = g_dev[dev];
  if (!pd.ready) {
    void *ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress (&ptr, g_noalloc_scratch);
    if (e != cudaSuccess)
      return nnsb_set_error (NNSB_ECUDA,
          "nnsb_stand_launch_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
          cudaGetErrorString (e));
    int sms = 0;
    e = cudaDeviceGetAttribute (&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess || sms <= 0)
      return nnsb_set_error (NNSB_ECUDA,
          "nnsb_stand_launch_noalloc: cannot read the SM count of device %d: %s", dev,
          cudaGetErrorString (e));
    pd.scratch = static_cast<double *> (ptr);
    pd.sm_count = sms;
    pd.ready = true;
  }
  *out = &pd;
  return NNSB_OK;
}

}  // namespace

/*
 * Standardise `batch` tensors without allocating anything (see include/nnstreamer_b200.h).
 *
 * The work runs on the calling thread's CURRENT CUDA device, on `cuda_stream` (a
 * cudaStream_t; NULL is the default stream); d_in / d_out must be device memory of that
 * device.  The first call on each device also performs the one-time set-up above, so make
 * that first call before starting any stream capture.
 *
 * All calls share one scratch array per device and it keeps whatever the previous call left
 * in it.  Nothing has to be cleared between calls: every slot a kernel reads is written
 * earlier in the same call.  For the same reason two calls that can overlap on the GPU (on
 * different streams, or from different threads) would overwrite each other's partial sums:
 * serialise them, e.g. by using one stream.
 */
extern "C" int
nnsb_stand_launch_noalloc (const void *d_in, void *d_out, int in_type, int out_type,
    const uint32_t *dim, int stand_mode, int per_channel, uint32_t batch, void *cuda_stream)
{
  if (!d_in || !d_out || !dim)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_stand_launch_noalloc: null argument");
  if (in_type < 0 || in_type >= NNSB_TYPE_END)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_stand_launch_noalloc: invalid input type %d", in_type);
  if (out_type == NNSB_TYPE_END)
    out_type = in_type;         /* the option's default: keep the input type */
  if (out_type < 0 || out_type >= NNSB_TYPE_END)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_stand_launch_noalloc: invalid output type %d", out_type);
  if (stand_mode != NNSB_STAND_DEFAULT && stand_mode != NNSB_STAND_DC_AVERAGE)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_stand_launch_noalloc: invalid stand mode %d", stand_mode);
  if (!nnsb_dim_is_valid (dim))
    return nnsb_set_error (NNSB_EINVAL, "nnsb_stand_launch_noalloc: invalid dimension "
        "(all NNSB_RANK_LIMIT entries are read; unused ones must be 1)");
  if (batch == 0)
    return NNSB_OK;
  if (batch > 65535)
    return nnsb_set_error (NNSB_ERANGE, "nnsb_stand_launch_noalloc: batch %u exceeds 65535 "
        "(one grid row per tensor); split the batch", batch);

  int dev = -1;
  cudaError_t e = cudaGetDevice (&dev);
  if (e != cudaSuccess)
    return nnsb_set_error (NNSB_ECUDA, "nnsb_stand_launch_noalloc: cudaGetDevice failed: %s",
        cudaGetErrorString (e));
  PerDevice *pd = nullptr;
  int r = setup_current_device (dev, &pd);
  if (r)
    return r;

  const uint64_t frame_elems = nnsb_element_count (dim);
  const uint32_t nch = per_channel ? dim[0] : 1;
  nnsb::LaunchCfg cfg;
  cfg.sm_count = pd->sm_count;
  cfg.blocks_per_sm = 0;        /* the grid nnsb_plan_launch uses: identical results */

  /* layout inside the array: [block partial sums][per-tensor {mean, std}] */
  size_t partial_doubles = nnsb::stand_scratch_doubles (batch, nch, frame_elems, cfg);
  const size_t stats_doubles = (size_t) batch * nch * 2;
  size_t need_bytes = (partial_doubles + stats_doubles) * sizeof (double);
  if (need_bytes > NNSB_NOALLOC_SCRATCH_BYTES) {
    /* the default grid keeps more block partials than the array holds (many channels): fall
     * back to four blocks per SM, a little slower, same arithmetic in another summation order */
    cfg.blocks_per_sm = 4;
    partial_doubles = nnsb::stand_scratch_doubles (batch, nch, frame_elems, cfg);
    need_bytes = (partial_doubles + stats_doubles) * sizeof (double);
  }
  if (need_bytes > NNSB_NOALLOC_SCRATCH_BYTES)
    return nnsb_set_error (NNSB_ERANGE,
        "nnsb_stand_launch_noalloc: batch %u x %u channel(s) needs %zu bytes of scratch, the "
        "built-in array holds %u; split the batch or use nnsb_plan_launch()", batch, nch,
        need_bytes, (unsigned) NNSB_NOALLOC_SCRATCH_BYTES);

  double *partial = pd->scratch;
  double *stats = pd->scratch + partial_doubles;
  int n_launches = 0;
  const int err = nnsb::launch_stand (in_type, out_type, stand_mode, d_in, d_out, frame_elems, nch,
      batch, partial, stats, cfg, static_cast<cudaStream_t> (cuda_stream), &n_launches);
  if (err)
    return nnsb_set_error (NNSB_ECUDA, "nnsb_stand_launch_noalloc: kernel launch failed: %s",
        cudaGetErrorString ((cudaError_t) err));
  return NNSB_OK;
}
