/*
 * nnsb_aggregator.cpp -- device-side mirror of GstTensorAggregator
 * (gst/nnstreamer/elements/gsttensor_aggregator.c; semantics gsttensor_aggregator.md:7-75).
 *
 * The reference keeps incoming buffers in a host GstAdapter, takes frames_out frames off its
 * head for every output (gst_adapter_get_buffer, :903), re-orders them with a host memcpy
 * loop when the frames are concatenated along an inner axis (:768-787) and flushes
 * frames_flush frames (:921-938).  Here the adapter is a byte range [head, tail) of one pooled,
 * reference-counted device block (the ring):
 *   push    a producer that asked for its output memory with nnsb_aggregator_acquire writes
 *           straight into the tail (nothing is moved); any other buffer is copied to the tail
 *           (H2D for host buffers, D2D for device ones);
 *   output  without concatenation an output is a SHARE of the ring at the head (nothing is
 *           moved; overlapping windows of frames_flush < frames_out simply overlap); with it,
 *           one launch of the permutation kernels (the block interleave is a rank-3 strided
 *           copy: out[block][frame][byte-in-block] <- in[frame][block][byte-in-block]);
 *   flush   pointer arithmetic.  When the tail reaches the end of the block while outputs still
 *           reference it, the few live bytes move to the front of a fresh block from the pool
 *           and the old one lives on until its last share is released.
 * Everything is stream-ordered (nnsb_memory): chain() does not wait for the GPU unless the
 * input is host memory (the caller may reuse that on return).
 *
 * Same property names and defaults (:276-281), the same caps arithmetic (:1026-1081), the same
 * timestamp interpolation (:873-900).  Not mirrored: the per-client adapter table keyed by the
 * tensor_query meta (:520-531) -- one stream per instance.
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "nnsb_caps.h"
#include "nnsb_internal.h"

using namespace nnsb;

namespace {

struct StampedOffset
{
  uint64_t offset;              /* position of the buffer's first byte in the pushed stream */
  uint64_t pts, dts;
};

int
copy_str (const std::string & s, char *out, size_t outlen)
{
  if (!out || s.size () + 1 > outlen)
    return nnsb_set_error (NNSB_EINVAL, "output buffer too small (%zu needed)", s.size () + 1);
  memcpy (out, s.c_str (), s.size () + 1);
  return (int) s.size ();
}

bool
parse_uint (const char *v, uint32_t * out)
{
  if (!v || !*v)
    return false;
  char *end = nullptr;
  unsigned long long x = strtoull (v, &end, 10);
  if (end == v || *end || x > 0xffffffffull || v[0] == '-')
    return false;
  *out = (uint32_t) x;
  return true;
}

/* gst_util_uint64_scale_int: val * num / denom, rounded down, without overflow */
uint64_t
scale_u64 (uint64_t val, uint64_t num, uint64_t denom)
{
  return (uint64_t) (((unsigned __int128) val * num) / denom);
}

}  // namespace

struct nnsb_aggregator
{
  int device_id = 0;
  nnsb_ctx *ctx = nullptr;
  nnsb_stream *stream = nullptr;

  bool silent = true;                   /* DEFAULT_SILENT, :76 */
  uint32_t frames_in = 1;               /* :81 */
  uint32_t frames_out = 1;              /* :86 */
  uint32_t frames_flush = 0;            /* :91 */
  uint32_t frames_dim = NNSB_RANK_LIMIT - 1;    /* :96 */
  bool concat = true;                   /* :101 */

  bool configured = false;
  nnsb_tensors_config in_config, out_config;

  /* the adapter: bytes [head, tail) of the ring hold the stream positions
   * [stream_pos, stream_pos + tail - head) */
  nnsb_memory ring;                     /* data == NULL: none yet */
  size_t capacity = 0, head = 0, tail = 0;
  uint64_t copied_in = 0, copied_out = 0;       /* bytes this element moved itself */
  uint64_t stream_pos = 0;              /* stream position of d_buf[head] */
  uint64_t pushed = 0;                  /* stream position of the next pushed byte */
  std::vector<StampedOffset> stamps;

  /* concatenation plan (depends on caps + properties) */
  nnsb_plan *concat_plan = nullptr;
  size_t concat_frame_size = 0;
  uint64_t launches = 0;
};

static void
drop_plan (nnsb_aggregator * a)
{
  if (a->concat_plan)
    nnsb_plan_destroy (a->concat_plan);
  a->concat_plan = nullptr;
  a->concat_frame_size = 0;
}

extern "C" int
nnsb_aggregator_new (int device_id, nnsb_aggregator ** out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_aggregator_new: null argument");
  nnsb_aggregator *a = new (std::nothrow) nnsb_aggregator ();
  if (!a)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_aggregator_new: out of memory");
  a->device_id = device_id;
  memset (&a->ring, 0, sizeof (a->ring));
  config_init (&a->in_config);
  config_init (&a->out_config);
  *out = a;
  return NNSB_OK;
}

/* gst_tensor_aggregator_reset, :979-990: clears the adapter */
extern "C" int
nnsb_aggregator_reset (nnsb_aggregator * a)
{
  if (!a)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_aggregator_reset: null argument");
  if (a->stream)
    nnsb_stream_sync (a->stream);
  /* outputs handed out earlier keep the old block; start over in a fresh one */
  nnsb_memory_release (&a->ring);
  memset (&a->ring, 0, sizeof (a->ring));
  a->capacity = 0;
  a->head = a->tail = 0;
  a->stream_pos = a->pushed = 0;
  a->stamps.clear ();
  return NNSB_OK;
}

extern "C" int
nnsb_aggregator_free (nnsb_aggregator * a)
{
  if (!a)
    return NNSB_OK;
  drop_plan (a);
  if (a->stream)
    nnsb_stream_sync (a->stream);
  nnsb_memory_release (&a->ring);
  if (a->stream)
    nnsb_stream_destroy (a->stream);
  if (a->ctx)
    nnsb_ctx_destroy (a->ctx);
  delete a;
  return NNSB_OK;
}

extern "C" int
nnsb_aggregator_set_property (nnsb_aggregator * a, const char *name, const char *value)
{
  if (!a || !name || !value)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_aggregator_set_property: null argument");
  const std::string n (name);
  uint32_t v;
  if (n == "frames-in" || n == "frames-out") {
    /* g_param_spec_uint (..., 1, G_MAXUINT, ...), :177-191 */
    if (!parse_uint (value, &v) || v < 1)
      return nnsb_set_error (NNSB_EINVAL, "%s: expected an unsigned integer >= 1", name);
    (n == "frames-in" ? a->frames_in : a->frames_out) = v;
    drop_plan (a);
    return NNSB_OK;
  }
  if (n == "frames-flush") {
    if (!parse_uint (value, &v))
      return nnsb_set_error (NNSB_EINVAL, "frames-flush: expected an unsigned integer");
    a->frames_flush = v;
    return NNSB_OK;
  }
  if (n == "frames-dim") {
    /* 0 .. NNS_TENSOR_RANK_LIMIT - 1, :212-216 */
    if (!parse_uint (value, &v) || v >= NNSB_RANK_LIMIT)
      return nnsb_set_error (NNSB_EINVAL, "frames-dim: expected 0..%d", NNSB_RANK_LIMIT - 1);
    a->frames_dim = v;
    drop_plan (a);
    return NNSB_OK;
  }
  if (n == "concat" || n == "silent") {
    std::string s (value);
    for (auto & c : s)
      if (c >= 'A' && c <= 'Z')
        c = (char) (c - 'A' + 'a');
    bool b;
    if (s == "true" || s == "1" || s == "yes")
      b = true;
    else if (s == "false" || s == "0" || s == "no")
      b = false;
    else
      return nnsb_set_error (NNSB_EINVAL, "%s: expected a boolean", name);
    if (n == "concat") {
      a->concat = b;
      drop_plan (a);
    } else {
      a->silent = b;
    }
    return NNSB_OK;
  }
  return nnsb_set_error (NNSB_EINVAL, "tensor_aggregator has no property '%s'", name);
}

extern "C" int
nnsb_aggregator_get_property (nnsb_aggregator * a, const char *name, char *buf, size_t buflen)
{
  if (!a || !name)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_aggregator_get_property: null argument");
  const std::string n (name);
  if (n == "frames-in")
    return copy_str (std::to_string (a->frames_in), buf, buflen);
  if (n == "frames-out")
    return copy_str (std::to_string (a->frames_out), buf, buflen);
  if (n == "frames-flush")
    return copy_str (std::to_string (a->frames_flush), buf, buflen);
  if (n == "frames-dim")
    return copy_str (std::to_string (a->frames_dim), buf, buflen);
  if (n == "concat")
    return copy_str (a->concat ? "true" : "false", buf, buflen);
  if (n == "silent")
    return copy_str (a->silent ? "true" : "false", buf, buflen);
  return nnsb_set_error (NNSB_EINVAL, "tensor_aggregator has no property '%s'", name);
}

/* gst_tensor_aggregator_parse_caps, :1026-1081.  Writes the source-pad caps to `out`. */
extern "C" int
nnsb_aggregator_set_caps (nnsb_aggregator * a, const char *incaps, char *out, size_t outlen)
{
  if (!a || !incaps)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_aggregator_set_caps: null argument");
  Caps caps;
  std::string err;
  if (!caps_parse (incaps, &caps, &err) || caps.structures.empty ())
    return nnsb_set_error (NNSB_EINVAL, "cannot parse caps '%s': %s", incaps, err.c_str ());
  if (!caps.is_fixed ())
    return nnsb_set_error (NNSB_EINVAL, "caps must be fixed");
  nnsb_tensors_config config;
  if (!config_from_structure (&config, caps.structures[0]) || !config_validate (config))
    return nnsb_set_error (NNSB_EINVAL, "Cannot configure tensor info");
  /* "Check properties to detect invalid case", :1051-1055 */
  const uint32_t count = (a->frames_out + a->frames_in - 1) / a->frames_in;
  if ((uint64_t) a->frames_in * count < a->frames_flush)
    return nnsb_set_error (NNSB_EINVAL, "Cannot flush frames");
  /* tensor-aggregator handles a single tensor (:1058; its pads are other/tensor, :111-125) */
  nnsb_tensor_info *info = &config.info[0];
  if (a->frames_dim >= NNSB_RANK_LIMIT || (info->dim[a->frames_dim] % a->frames_in) != 0)
    return nnsb_set_error (NNSB_EINVAL, "Cannot update dimension in output tensor");
  a->in_config = config;
  const uint32_t per_frame = info->dim[a->frames_dim] / a->frames_in;
  info->dim[a->frames_dim] = per_frame * a->frames_out;
  a->out_config = config;
  a->configured = true;
  drop_plan (a);
  if (out)
    return copy_str (caps_to_string (tensor_caps_from_config (a->out_config)), out, outlen);
  return NNSB_OK;
}

extern "C" int
nnsb_aggregator_get_config (nnsb_aggregator * a, int direction, nnsb_tensors_config * cfg)
{
  if (!a || !cfg || !a->configured)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_aggregator_get_config: not configured");
  *cfg = direction == NNSB_PAD_SINK ? a->in_config : a->out_config;
  return NNSB_OK;
}

extern "C" int
nnsb_aggregator_kernel_launches (nnsb_aggregator * a, uint64_t * count)
{
  if (!a || !count)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_aggregator_kernel_launches: null argument");
  *count = a->launches;
  return NNSB_OK;
}

/* outputs the next chain call with a buffer of in_size bytes will produce (:865-941) */
extern "C" int
nnsb_aggregator_outputs_for (nnsb_aggregator * a, size_t in_size)
{
  if (!a || in_size == 0)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_aggregator_outputs_for: invalid argument");
  if (a->frames_in == a->frames_out)
    return 1;
  const size_t frame_size = in_size / a->frames_in;
  const size_t out_size = frame_size * a->frames_out;
  if (out_size == 0)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_aggregator_outputs_for: buffer smaller than frames-in");
  size_t avail = (a->tail - a->head) + in_size;
  int n = 0;
  while (avail >= out_size) {
    size_t flush = a->frames_flush > 0 ? frame_size * a->frames_flush : out_size;
    if (flush > avail)
      flush = avail;
    avail -= flush;
    n++;
    if (flush == 0)
      break;
  }
  return n;
}

static int
ensure_device (nnsb_aggregator * a)
{
  if (a->ctx)
    return NNSB_OK;
  int r = nnsb_ctx_create (a->device_id, &a->ctx);
  if (r == NNSB_OK)
    r = nnsb_stream_create (a->ctx, &a->stream);
  return r;
}

/* gst_tensor_aggregator_check_concat_axis, :539-559 */
static bool
needs_concat (const nnsb_aggregator * a, const nnsb_tensor_info & frame)
{
  if (a->concat && a->frames_out > 1)
    for (uint32_t i = a->frames_dim + 1; i < NNSB_RANK_LIMIT; i++)
      if (frame.dim[i] > 1)
        return true;
  return false;
}

/* The block interleave of gst_tensor_aggregator_concat (:768-787) as a strided-copy plan:
 *   block = es * prod (frame.dim[0..frames_dim]) bytes, n_blocks = frame_size / block,
 *   out[b][f][x] = in[f][b][x]. */
static int
ensure_concat_plan (nnsb_aggregator * a, const nnsb_tensor_info & frame, size_t frame_size)
{
  if (a->concat_plan && a->concat_frame_size == frame_size)
    return NNSB_OK;
  drop_plan (a);
  const size_t es = nnsb_element_size (frame.type);
  uint64_t block_elems = 1;
  for (uint32_t i = 0; i <= a->frames_dim; i++)
    block_elems *= frame.dim[i];
  const uint64_t frame_elems = frame_size / es;
  const uint64_t n_blocks = frame_elems / block_elems;
  if (block_elems > 0xffffffffull || n_blocks > 0xffffffffull)
    return nnsb_set_error (NNSB_ERANGE, "tensor_aggregator: frame too large to concatenate");
  const uint32_t size[3] = { (uint32_t) block_elems, a->frames_out, (uint32_t) n_blocks };
  const uint64_t in_stride[3] = { 1, frame_elems, block_elems };
  int r = nnsb_plan_create_permute (a->ctx, (int) es, 3, size, in_stride, &a->concat_plan);
  if (r == NNSB_OK)
    a->concat_frame_size = frame_size;
  return r;
}

/* gst_tensor_aggregator_push, :805-830: `src` (a device memory: the caller's buffer, a staged
 * copy of it, or the ring) holds frames_out frames back to back at byte `off`; the result
 * (concatenated if needed) becomes *out. */
static int
produce (nnsb_aggregator * a, const nnsb_memory * src, size_t off, size_t frame_size, nnsb_memory * out)
{
  nnsb_tensor_info frame = a->out_config.info[0];
  frame.dim[a->frames_dim] /= a->frames_out;
  if (frame_size != nnsb_info_size (&frame) || frame_size == 0) {
    nnsb_set_error (NNSB_FLOW_ERROR,
        "Invalid output capability of tensor_aggregator. Frame size = %zu", frame_size);
    return NNSB_FLOW_ERROR;
  }
  const size_t out_size = frame_size * a->frames_out;
  if (!needs_concat (a, frame)) {
    /* gst_pad_push of (a sub-buffer of) what is already there: one more reference, no copy */
    return nnsb_memory_share (src, off, out_size, out) == NNSB_OK ? NNSB_FLOW_OK : NNSB_FLOW_ERROR;
  }
  if (nnsb_memory_alloc (a->ctx, out_size, 1, a->stream, out) != NNSB_OK)
    return NNSB_FLOW_ERROR;
  int r = ensure_concat_plan (a, frame, frame_size);
  if (r == NNSB_OK)
    r = nnsb_memory_wait (src, a->stream);
  if (r == NNSB_OK)
    r = nnsb_plan_launch (a->concat_plan, (const char *) src->data + off, out->data, 1, a->stream);
  if (r == NNSB_OK) {
    a->launches += (uint64_t) nnsb_plan_last_launch_count (a->concat_plan);
    a->copied_out += out_size;
    r = nnsb_memory_mark_io (a->stream, out, 1, src, 1);
  }
  if (r != NNSB_OK) {
    nnsb_stream_sync (a->stream);
    nnsb_memory_release (out);
    return NNSB_FLOW_ERROR;
  }
  return NNSB_FLOW_OK;
}

/* room for `need` more bytes behind the tail.  An empty adapter rewinds only when nobody else
 * references the block (outputs are shares of it); otherwise, and when the tail reaches the end,
 * the live bytes move to the front of a fresh block from the pool. */
static int
make_room (nnsb_aggregator * a, size_t need, size_t out_size)
{
  if (a->ring.data && a->head == a->tail && nnsb_memory_is_unique (&a->ring))
    a->head = a->tail = 0;
  if (a->ring.data && a->tail + need <= a->capacity)
    return NNSB_OK;
  const size_t live = a->tail - a->head;
  size_t cap = a->capacity;
  /* Room for a few outputs, a whole number of them: when the inputs tile the outputs (frames-in
   * divides frames-out, no overlap) the block then runs full exactly at an output boundary and
   * nothing has to move.  Blocks of one size cycle through the pool as their outputs are released. */
  size_t want = (live + need + 2 * out_size + out_size - 1) / out_size * out_size;
  if (want < 3 * out_size)
    want = 3 * out_size;
  if (cap < want)
    cap = want;
  nnsb_memory fresh;
  int r = nnsb_memory_alloc (a->ctx, cap, 1, a->stream, &fresh);
  if (r != NNSB_OK)
    return r;
  if (live) {
    r = nnsb_memory_wait (&a->ring, a->stream);
    if (r == NNSB_OK)
      r = nnsb_d2d_async (a->stream, fresh.data, (char *) a->ring.data + a->head, live);
    if (r == NNSB_OK)
      r = nnsb_memory_mark_io (a->stream, &fresh, 1, &a->ring, 1);
    if (r != NNSB_OK) {
      nnsb_stream_sync (a->stream);
      nnsb_memory_release (&fresh);
      return r;
    }
    a->copied_in += live;
  }
  nnsb_memory_release (&a->ring);       /* lives on while outputs share it */
  a->ring = fresh;
  a->capacity = cap;
  a->head = 0;
  a->tail = live;
  return NNSB_OK;
}

/* The memory the NEXT buffer should be written into: a share of the ring at the tail.  A
 * producer that fills it (and marks it ready on its stream) and passes it to chain() has
 * pushed its buffer without a copy -- what a downstream-proposed buffer pool does in GStreamer
 * (gst_base_transform decide_allocation).  Same signature as nnsb_output_provider_fn. */
extern "C" int
nnsb_aggregator_acquire (void *aggregator, size_t size, nnsb_memory * out)
{
  nnsb_aggregator *a = (nnsb_aggregator *) aggregator;
  if (!a || !out || size == 0)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_aggregator_acquire: invalid argument");
  if (!a->configured || a->frames_in == a->frames_out)
    return nnsb_set_error (NNSB_ENOTSUP, "nnsb_aggregator_acquire: nothing is accumulated in this configuration");
  int r = ensure_device (a);
  if (r)
    return r;
  const size_t frame_size = size / a->frames_in;
  r = make_room (a, size, frame_size * a->frames_out);
  if (r)
    return r;
  return nnsb_memory_share (&a->ring, a->tail, size, out);
}

extern "C" int
nnsb_aggregator_copy_counters (nnsb_aggregator * a, uint64_t * copied_in, uint64_t * copied_out)
{
  if (!a)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_aggregator_copy_counters: null argument");
  if (copied_in)
    *copied_in = a->copied_in;
  if (copied_out)
    *copied_out = a->copied_out;
  return NNSB_OK;
}

/* gst_adapter_prev_pts / _prev_dts at the head of the adapter: the last valid timestamp of a
 * buffer that starts at or before the head, and the distance in bytes from that buffer's start */
static void
prev_stamp (const nnsb_aggregator * a, bool want_dts, uint64_t * ts, uint64_t * dist)
{
  *ts = NNSB_CLOCK_TIME_NONE;
  *dist = 0;
  for (size_t i = a->stamps.size (); i-- > 0;) {
    const StampedOffset & s = a->stamps[i];
    const uint64_t v = want_dts ? s.dts : s.pts;
    if (s.offset <= a->stream_pos && v != NNSB_CLOCK_TIME_NONE) {
      *ts = v;
      *dist = a->stream_pos - s.offset;
      return;
    }
  }
  /* no timestamp seen yet: GstAdapter reports the distance from the start of the stream */
  *dist = a->stream_pos;
}

/* gst_tensor_aggregator_chain, :835-941 */
extern "C" int
nnsb_aggregator_chain (nnsb_aggregator * a, const nnsb_memory * in, const nnsb_buffer_time * t_in,
    nnsb_memory * outs, nnsb_buffer_time * t_outs, uint32_t max_out, uint32_t * n_out)
{
  if (!a || !in || !outs || !n_out)
    return NNSB_FLOW_ERROR;
  *n_out = 0;
  if (!a->configured) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_aggregator: caps are not set");
    return NNSB_FLOW_ERROR;
  }
  if (in->size == 0 || !in->data) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_aggregator: empty buffer");
    return NNSB_FLOW_ERROR;
  }
  if (ensure_device (a) != NNSB_OK)
    return NNSB_FLOW_ERROR;
  /* a device input may still be in flight on its producer's stream */
  if (in->on_device && nnsb_memory_wait (in, a->stream) != NNSB_OK)
    return NNSB_FLOW_ERROR;
  const uint32_t frames_in = a->frames_in, frames_out = a->frames_out;
  const size_t buf_size = in->size;
  const size_t frame_size = buf_size / frames_in;
  const int need = nnsb_aggregator_outputs_for (a, buf_size);
  if (need < 0)
    return NNSB_FLOW_ERROR;
  if ((uint32_t) need > max_out) {
    nnsb_set_error (NNSB_ERANGE, "tensor_aggregator: this buffer yields %d outputs, room for %u",
        need, max_out);
    return NNSB_FLOW_ERROR;
  }
  nnsb_buffer_time none = { NNSB_CLOCK_TIME_NONE, NNSB_CLOCK_TIME_NONE, NNSB_CLOCK_TIME_NONE };
  const nnsb_buffer_time tin = t_in ? *t_in : none;

  if (frames_in == frames_out) {
    /* push the incoming buffer (do concat if needed), :856-859 */
    nnsb_memory staged;
    memset (&staged, 0, sizeof (staged));
    const nnsb_memory *src = in;
    if (!in->on_device) {
      if (nnsb_memory_alloc (a->ctx, buf_size, 1, a->stream, &staged) != NNSB_OK)
        return NNSB_FLOW_ERROR;
      if (nnsb_h2d_async (a->stream, staged.data, in->data, buf_size) != NNSB_OK) {
        nnsb_stream_sync (a->stream);
        nnsb_memory_release (&staged);
        return NNSB_FLOW_ERROR;
      }
      a->copied_in += buf_size;
      src = &staged;
    }
    int r = produce (a, src, 0, frame_size, &outs[0]);
    if (staged.data) {
      nnsb_stream_sync (a->stream);     /* the caller may reuse its host buffer */
      nnsb_memory_release (&staged);    /* the output holds its own reference if it is a share */
    }
    if (r != NNSB_FLOW_OK)
      return r;
    if (t_outs)
      t_outs[0] = tin;
    *n_out = 1;
    return NNSB_FLOW_OK;
  }

  uint64_t duration = tin.duration;
  if (duration != NNSB_CLOCK_TIME_NONE)
    duration = scale_u64 (duration, frames_out, frames_in);     /* :868-871 */

  /* gst_adapter_push, :873 */
  const size_t out_size = frame_size * frames_out;
  const bool in_place = a->ring.data && in->on_device && in->block && in->block == a->ring.block
      && (const char *) in->data == (const char *) a->ring.data + a->tail
      && a->tail + buf_size <= a->capacity;
  if (!in_place) {
    if (make_room (a, buf_size, out_size) != NNSB_OK)
      return NNSB_FLOW_ERROR;
    char *dst = (char *) a->ring.data + a->tail;
    int r = in->on_device ? nnsb_d2d_async (a->stream, dst, in->data, buf_size)
        : nnsb_h2d_async (a->stream, dst, in->data, buf_size);
    if (r == NNSB_OK && in->on_device)
      r = nnsb_memory_mark_io (a->stream, &a->ring, 1, in, 1);
    if (r != NNSB_OK)
      return NNSB_FLOW_ERROR;
    a->copied_in += buf_size;
  }
  /* in place: the producer wrote the bytes where they belong and marked the block ready */
  a->tail += buf_size;
  a->stamps.push_back ({ a->pushed, tin.pts, tin.dts });
  a->pushed += buf_size;

  int ret = NNSB_FLOW_OK;
  size_t avail;
  while ((avail = a->tail - a->head) >= out_size && ret == NNSB_FLOW_OK) {
    uint64_t pts, dts, pts_dist, dts_dist;
    prev_stamp (a, false, &pts, &pts_dist);
    prev_stamp (a, true, &dts, &dts_dist);
    /* "If frames-in is larger then frames-out, the same timestamp would be returned", :886-906 */
    if (frames_in > 1) {
      const int64_t fn = a->in_config.rate_n, fd = a->in_config.rate_d;
      if (fn > 0 && fd > 0) {
        if (pts != NNSB_CLOCK_TIME_NONE)
          pts += scale_u64 (pts_dist * (uint64_t) fd, 1000000000ull, (uint64_t) fn * frame_size);
        if (dts != NNSB_CLOCK_TIME_NONE)
          dts += scale_u64 (dts_dist * (uint64_t) fd, 1000000000ull, (uint64_t) fn * frame_size);
      }
    }
    ret = produce (a, &a->ring, a->head, frame_size, &outs[*n_out]);
    if (ret != NNSB_FLOW_OK)
      break;
    if (t_outs) {
      t_outs[*n_out].pts = pts;
      t_outs[*n_out].dts = dts;
      t_outs[*n_out].duration = duration;
    }
    (*n_out)++;
    /* flush data, :921-938 */
    size_t flush = a->frames_flush > 0 ? frame_size * a->frames_flush : out_size;
    if (flush > avail)
      flush = avail;
    a->head += flush;
    a->stream_pos += flush;
    if (flush == 0)
      break;
  }
  /* timestamps of buffers that lie entirely before the head are no longer reachable */
  while (a->stamps.size () > 1 && a->stamps[1].offset <= a->stream_pos)
    a->stamps.erase (a->stamps.begin ());
  /* a host buffer, or device memory the caller manages itself (no block to hold on to), may be
   * reused by the caller once we return; for pooled blocks the device work just stays queued */
  if ((!in->on_device || !in->block) && !in_place && nnsb_stream_sync (a->stream) != NNSB_OK)
    return NNSB_FLOW_ERROR;
  if (ret != NNSB_FLOW_OK) {
    for (uint32_t k = 0; k < *n_out; k++)
      nnsb_memory_release (&outs[k]);
    *n_out = 0;
  }
  return ret;
}
