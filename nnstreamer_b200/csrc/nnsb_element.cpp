/*
 * nnsb_element.cpp -- host-side mirrors of the two reference elements on top of the C-ABI
 * runtime (no GLib / GStreamer; compiled with g++, calls only nnsb_* functions).
 *
 *   nnsb_transform  <->  GstTensorTransform   (gst/nnstreamer/elements/gsttensor_transform.c)
 *   nnsb_converter  <->  GstTensorConverter, video branch (gsttensor_converter.c)
 *   nnsb_queue           device-side batching queue (north_star; replaces the host GstAdapter
 *                        concat of _gst_tensor_converter_chain_chunk, :945-1009)
 *
 * Same property names / value strings, same caps strings, same flow-return conventions; each
 * function cites the reference lines it follows.
 */
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "nnsb_caps.h"
#include "nnsb_internal.h"

using namespace nnsb;

/* ------------------------------------------------------------------------------------- */
/* small helpers                                                                           */
/* ------------------------------------------------------------------------------------- */
static int
copy_out (const std::string & s, char *out, size_t outlen)
{
  if (!out || s.size () + 1 > outlen)
    return nnsb_set_error (NNSB_EINVAL, "output buffer too small (%zu needed)", s.size () + 1);
  memcpy (out, s.c_str (), s.size () + 1);
  return (int) s.size ();
}

static bool
parse_bool (const char *v, bool * out)
{
  if (!v)
    return false;
  std::string s (v);
  for (auto & c : s)
    if (c >= 'A' && c <= 'Z')
      c = (char) (c - 'A' + 'a');
  if (s == "true" || s == "1" || s == "yes") { *out = true; return true; }
  if (s == "false" || s == "0" || s == "no") { *out = false; return true; }
  return false;
}

/* ------------------------------------------------------------------------------------- */
/* tensor_transform                                                                        */
/* ------------------------------------------------------------------------------------- */
struct PlanSlot
{
  nnsb_plan *plan = nullptr;
  nnsb_tensor_info info;
  int accel = -1;
};

struct nnsb_transform
{
  int device_id = 0;
  nnsb_ctx *ctx = nullptr;
  nnsb_stream *stream[2] = { nullptr, nullptr };
  nnsb_event *ev[2] = { nullptr, nullptr };

  bool silent = true;                   /* :291 */
  int mode = NNSB_MODE_UNKNOWN;         /* :292 */
  bool has_option = false;              /* option == NULL, :293 */
  std::string option;
  nnsb_option opt;
  bool loaded = false;                  /* :294 */
  bool acceleration = false;            /* DEFAULT_ACCELERATION without ORC, :105-109 */
  std::vector<int> apply;               /* :297 */

  nnsb_tensors_config in_config, out_config;
  bool caps_set = false;
  PlanSlot slots[NNSB_SIZE_LIMIT];
  uint64_t launches = 0;
  nnsb_output_provider_fn provider = nullptr;   /* downstream hands out the device outputs */
  void *provider_user = nullptr;
};

static bool
in_apply (const nnsb_transform * t, unsigned idx)
{
  if (t->apply.empty ())
    return true;
  for (int v : t->apply)
    if (v == (int) idx)
      return true;
  return false;
}

/* gst_tensor_transform_set_option_data, :652-958: TRUE if OK or skipped */
static bool
apply_option (nnsb_transform * t)
{
  if (t->mode == NNSB_MODE_UNKNOWN || !t->has_option)
    return true;
  nnsb_option parsed;
  if (nnsb_option_parse (t->mode, t->option.c_str (), &parsed) != NNSB_OK)
    return false;
  t->opt = parsed;
  t->loaded = true;
  /* plans depend on the option */
  for (auto & s : t->slots) {
    if (s.plan)
      nnsb_plan_destroy (s.plan);
    s.plan = nullptr;
  }
  return true;
}

extern "C" int
nnsb_transform_new (int device_id, nnsb_transform ** out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_transform_new: null argument");
  nnsb_transform *t = new (std::nothrow) nnsb_transform ();
  if (!t)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_transform_new: out of memory");
  t->device_id = device_id;
  memset (&t->opt, 0, sizeof (t->opt));
  t->opt.mode = NNSB_MODE_UNKNOWN;
  config_init (&t->in_config);
  config_init (&t->out_config);
  *out = t;
  return NNSB_OK;
}

extern "C" int
nnsb_transform_free (nnsb_transform * t)
{
  if (!t)
    return NNSB_OK;
  for (auto & s : t->slots)
    if (s.plan)
      nnsb_plan_destroy (s.plan);
  for (int i = 0; i < 2; i++) {
    if (t->ev[i])
      nnsb_event_destroy (t->ev[i]);
    if (t->stream[i])
      nnsb_stream_destroy (t->stream[i]);
  }
  if (t->ctx)
    nnsb_ctx_destroy (t->ctx);
  delete t;
  return NNSB_OK;
}

static int
ensure_device (nnsb_transform * t)
{
  if (t->ctx)
    return NNSB_OK;
  int r = nnsb_ctx_create (t->device_id, &t->ctx);
  if (r)
    return r;
  /* device-resident buffers need one stream; the second one (host-buffer path: chunks alternate
   * between two streams) is created on first use -- idle streams still occupy connections */
  return nnsb_stream_create (t->ctx, &t->stream[0]);
}

static int
ensure_second_stream (nnsb_transform * t)
{
  if (t->stream[1])
    return NNSB_OK;
  return nnsb_stream_create (t->ctx, &t->stream[1]);
}

extern "C" int
nnsb_transform_set_property (nnsb_transform * t, const char *name, const char *value)
{
  if (!t || !name)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_transform_set_property: null argument");
  const std::string n (name);
  if (n == "silent") {
    bool b;
    if (!parse_bool (value, &b))
      return nnsb_set_error (NNSB_EINVAL, "silent: expected a boolean");
    t->silent = b;
    return NNSB_OK;
  }
  if (n == "mode") {
    /* :973-976: g_value_get_enum; accepts the nick or the number */
    if (!value)
      return nnsb_set_error (NNSB_EINVAL, "mode: null value");
    int m = nnsb_mode_from_string (value);
    if (m == NNSB_MODE_UNKNOWN) {
      char *end = nullptr;
      long v = strtol (value, &end, 10);
      if (end != value && *end == '\0' && v >= -1 && v <= NNSB_MODE_CLAMP)
        m = (int) v;
      else if (strcmp (value, "unknown") != 0)
        return nnsb_set_error (NNSB_EINVAL, "mode: '%s' is not a value of gtt_mode_type", value);
    }
    t->mode = m;
    apply_option (t);
    return NNSB_OK;
  }
  if (n == "option") {
    /* :977-992 */
    const bool had = t->has_option;
    const std::string backup = t->option;
    t->has_option = value != nullptr;
    t->option = value ? value : "";
    if (apply_option (t))
      return NNSB_OK;
    /* ERROR! Revert the change! */
    t->has_option = had;
    t->option = backup;
    apply_option (t);
    return NNSB_EINVAL;         /* nnsb_last_error() holds the parser's message */
  }
  if (n == "acceleration") {
    bool b;
    if (!parse_bool (value, &b))
      return nnsb_set_error (NNSB_EINVAL, "acceleration: expected a boolean");
    t->acceleration = b;
    return NNSB_OK;
  }
  if (n == "apply") {
    /* :1002-1020: comma separated indices, appended to the list */
    if (!value)
      return nnsb_set_error (NNSB_EINVAL, "apply: null value");
    std::string v (value);
    size_t start = 0;
    while (true) {
      size_t p = v.find (',', start);
      std::string tok = v.substr (start, p == std::string::npos ? std::string::npos : p - start);
      t->apply.push_back ((int) strtoll (tok.c_str (), nullptr, 10));
      if (p == std::string::npos)
        break;
      start = p + 1;
    }
    return NNSB_OK;
  }
  if (n == "transpose-rank-limit")
    return nnsb_set_error (NNSB_EINVAL, "transpose-rank-limit is read-only");
  return nnsb_set_error (NNSB_EINVAL, "tensor_transform has no property '%s'", name);
}

extern "C" int
nnsb_transform_get_property (nnsb_transform * t, const char *name, char *buf, size_t buflen)
{
  if (!t || !name)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_transform_get_property: null argument");
  const std::string n (name);
  if (n == "silent")
    return copy_out (t->silent ? "true" : "false", buf, buflen);
  if (n == "mode")
    return copy_out (nnsb_mode_to_string (t->mode), buf, buflen);
  if (n == "option") {
    if (!t->has_option) {
      if (buf && buflen)
        buf[0] = '\0';
      return NNSB_ENOENT;
    }
    return copy_out (t->option, buf, buflen);
  }
  if (n == "acceleration")
    return copy_out (t->acceleration ? "true" : "false", buf, buflen);
  if (n == "apply") {
    std::string s;                      /* :1049-1073 */
    for (size_t i = 0; i < t->apply.size (); i++) {
      if (i)
        s += ",";
      s += std::to_string (t->apply[i]);
    }
    return copy_out (s, buf, buflen);
  }
  if (n == "transpose-rank-limit")
    return copy_out (std::to_string (NNSB_TRANSPOSE_RANK_LIMIT), buf, buflen);
  return nnsb_set_error (NNSB_EINVAL, "tensor_transform has no property '%s'", name);
}

/* gst_tensor_transform_convert_dimension incl. the `apply` filter, :1857-1861 */
static bool
convert_info (const nnsb_transform * t, int direction, unsigned idx, const nnsb_tensor_info * in,
    nnsb_tensor_info * out)
{
  *out = *in;
  if (!in_apply (t, idx))
    return true;
  if (t->mode == NNSB_MODE_UNKNOWN)
    return false;               /* default: return FALSE, :1954-1955 */
  nnsb_option opt = t->opt;
  opt.mode = t->mode;           /* the reference switches on filter->mode */
  return nnsb_convert_dimension (&opt, direction, in, out) == NNSB_OK;
}

static Caps
transform_caps_impl (nnsb_transform * t, int direction, const Caps & caps, const Caps * filter)
{
  Caps result;
  for (const auto & st : caps.structures) {
    nnsb_tensors_config in_config, out_config;
    config_init (&out_config);
    bool types_not_fixed = false;
    config_from_structure (&in_config, st);
    if (in_config.format == NNSB_FORMAT_FLEXIBLE) {
      out_config.format = NNSB_FORMAT_FLEXIBLE;         /* :1999-2001 */
    } else {
      for (uint32_t j = 0; j < in_config.num_tensors; j++) {
        convert_info (t, direction, j, &in_config.info[j], &out_config.info[j]);
        if (out_config.info[j].type == NNSB_TYPE_END)
          types_not_fixed = true;
      }
    }
    out_config.rate_d = in_config.rate_d;
    out_config.rate_n = in_config.rate_n;
    out_config.num_tensors = in_config.num_tensors;
    Caps aux = (st.name == "other/tensor") ? tensor_caps_from_config (out_config)
        : tensors_caps_from_config (out_config);
    if (types_not_fixed)
      for (auto & s : aux.structures)
        s.remove ("types");                             /* :2023-2030 */
    for (auto & s : aux.structures)
      result.structures.push_back (s);
  }
  if (filter && !filter->empty () && !filter->any) {
    caps_update_dimension (&result, *filter);           /* :2040 */
    result = caps_intersect (result, *filter);          /* GST_CAPS_INTERSECT_FIRST */
  }
  return result;
}

extern "C" int
nnsb_transform_transform_caps (nnsb_transform * t, int direction, const char *caps,
    const char *filter, char *out, size_t outlen)
{
  if (!t || !caps)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_transform_transform_caps: null argument");
  Caps c, f;
  std::string err;
  if (!caps_parse (caps, &c, &err))
    return nnsb_set_error (NNSB_EINVAL, "cannot parse caps '%s': %s", caps, err.c_str ());
  if (filter && !caps_parse (filter, &f, &err))
    return nnsb_set_error (NNSB_EINVAL, "cannot parse filter caps '%s': %s", filter, err.c_str ());
  Caps r = transform_caps_impl (t, direction, c, filter ? &f : nullptr);
  return copy_out (caps_to_string (r), out, outlen);
}

extern "C" int
nnsb_transform_fixate_caps (nnsb_transform * t, int direction, const char *caps,
    const char *othercaps, char *out, size_t outlen)
{
  if (!t || !caps || !othercaps)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_transform_fixate_caps: null argument");
  Caps c, o;
  std::string err;
  if (!caps_parse (caps, &c, &err) || !caps_parse (othercaps, &o, &err))
    return nnsb_set_error (NNSB_EINVAL, "cannot parse caps: %s", err.c_str ());
  Caps r = caps_fixate (transform_caps_impl (t, direction, c, &o));     /* :2079-2084 */
  return copy_out (caps_to_string (r), out, outlen);
}

extern "C" int
nnsb_transform_set_caps (nnsb_transform * t, const char *incaps, const char *outcaps)
{
  if (!t || !incaps || !outcaps)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_transform_set_caps: null argument");
  Caps ci, co;
  std::string err;
  if (!caps_parse (incaps, &ci, &err) || ci.structures.empty ())
    return nnsb_set_error (NNSB_EINVAL, "Cannot read cap of incaps");
  if (!caps_parse (outcaps, &co, &err) || co.structures.empty ())
    return nnsb_set_error (NNSB_EINVAL, "Cannot read cap of outcaps");
  nnsb_tensors_config in_config, out_config, config;
  /* gst_tensor_transform_read_caps, :1823-1839 */
  if (!config_from_structure (&in_config, ci.structures[0]) || !config_validate (in_config))
    return nnsb_set_error (NNSB_EINVAL, "Cannot read cap of incaps");
  if (!config_from_structure (&out_config, co.structures[0]) || !config_validate (out_config))
    return nnsb_set_error (NNSB_EINVAL, "Cannot read cap of outcaps");
  const bool in_flexible = in_config.format == NNSB_FORMAT_FLEXIBLE;
  const bool out_flexible = out_config.format == NNSB_FORMAT_FLEXIBLE;
  config_init (&config);
  config.format = out_config.format;
  config.rate_n = in_config.rate_n;
  config.rate_d = in_config.rate_d;
  config.num_tensors = in_config.num_tensors;
  if (!in_flexible) {
    for (uint32_t i = 0; i < in_config.num_tensors; i++)
      if (!convert_info (t, NNSB_PAD_SINK, i, &in_config.info[i], &config.info[i]))
        return nnsb_set_error (NNSB_EINVAL, "Tensor info is not matched with given properties.");
  }
  if (out_flexible) {
    if (!in_flexible)
      out_config = config;
  } else if (!config_is_equal (out_config, config)) {
    return nnsb_set_error (NNSB_EINVAL, "Tensor info is not matched with given properties.");
  }
  t->in_config = in_config;
  t->out_config = out_config;
  t->caps_set = true;
  return NNSB_OK;
}

extern "C" int
nnsb_transform_transform_size (nnsb_transform * t, int direction, const char *caps, size_t size,
    const char *othercaps, size_t *othersize)
{
  (void) t; (void) direction; (void) caps; (void) size; (void) othercaps;
  if (othersize)
    *othersize = 0;
  return NNSB_OK;
}

extern "C" int
nnsb_transform_get_config (nnsb_transform * t, int direction, nnsb_tensors_config * cfg)
{
  if (!t || !cfg)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_transform_get_config: null argument");
  *cfg = (direction == NNSB_PAD_SINK) ? t->in_config : t->out_config;
  return NNSB_OK;
}

extern "C" int
nnsb_transform_alloc (nnsb_transform * t, size_t size, int on_device, nnsb_memory * out)
{
  if (!t || !out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_transform_alloc: null argument");
  int r = ensure_device (t);
  if (r)
    return r;
  return nnsb_memory_alloc (t->ctx, size, on_device, t->stream[0], out);
}

extern "C" int
nnsb_transform_set_output_provider (nnsb_transform * t, nnsb_output_provider_fn fn, void *user)
{
  if (!t)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_transform_set_output_provider: null argument");
  t->provider = fn;
  t->provider_user = fn ? user : nullptr;
  return NNSB_OK;
}

/* the memory a transformed device tensor is written to */
static int
alloc_device_output (nnsb_transform * t, size_t size, nnsb_memory * out)
{
  if (t->provider) {
    nnsb_memory m;
    memset (&m, 0, sizeof (m));
    if (t->provider (t->provider_user, size, &m) == NNSB_OK) {
      if (m.data && m.on_device && m.size == size) {
        /* the block may still be read by the provider's earlier consumers */
        if (nnsb_memory_wait (&m, t->stream[0]) == NNSB_OK) {
          *out = m;
          return NNSB_OK;
        }
      }
      nnsb_memory_release (&m);
    }
  }
  return nnsb_memory_alloc (t->ctx, size, 1, t->stream[0], out);
}

extern "C" int
nnsb_transform_download (nnsb_transform * t, void *host_dst, const nnsb_memory * dev_src)
{
  if (!t || !host_dst || !dev_src || !dev_src->on_device)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_transform_download: invalid argument");
  int r = ensure_device (t);
  if (!r)
    r = nnsb_memory_wait (dev_src, t->stream[0]);
  if (!r)
    r = nnsb_d2h_async (t->stream[0], host_dst, dev_src->data, dev_src->size);
  if (!r)
    r = nnsb_stream_sync (t->stream[0]);
  return r;
}

extern "C" int
nnsb_transform_upload (nnsb_transform * t, const nnsb_memory * dev_dst, const void *host_src)
{
  if (!t || !host_src || !dev_dst || !dev_dst->on_device)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_transform_upload: invalid argument");
  int r = ensure_device (t);
  if (!r)
    r = nnsb_memory_wait (dev_dst, t->stream[0]);
  if (!r)
    r = nnsb_h2d_async (t->stream[0], dev_dst->data, host_src, dev_dst->size);
  if (!r)
    r = nnsb_stream_sync (t->stream[0]);
  return r;
}

extern "C" int
nnsb_transform_kernel_launches (nnsb_transform * t, uint64_t * count)
{
  if (!t || !count)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_transform_kernel_launches: null argument");
  *count = t->launches;
  return NNSB_OK;
}

static nnsb_plan *
plan_for (nnsb_transform * t, unsigned idx, const nnsb_tensor_info * in_info)
{
  PlanSlot & s = t->slots[idx];
  if (s.plan && s.accel == (int) t->acceleration && memcmp (&s.info, in_info, sizeof (*in_info)) == 0)
    return s.plan;
  if (s.plan)
    nnsb_plan_destroy (s.plan);
  s.plan = nullptr;
  nnsb_option opt = t->opt;
  if (nnsb_plan_create (t->ctx, &opt, in_info, t->acceleration, &s.plan) != NNSB_OK)
    return nullptr;
  s.info = *in_info;
  s.accel = (int) t->acceleration;
  return s.plan;
}

/* Pinned staging for PAGEABLE host inputs (north_star: "pinned-host staging with
 * cudaMemcpyAsync").  An upstream GstMemory is ordinary malloc'd memory: cudaMemcpyAsync from it
 * is staged by the driver and blocks the caller, so the H2D || kernel || D2H pipeline below
 * would collapse into a sequence.  Instead helper threads memcpy chunk k+1.. of the source into
 * a small ring of pinned slots while the copy engine moves chunk k; a slot is refilled only
 * after the H2D that read it has completed (one event per slot). */
struct Stager
{
  static const int RING = 4;
  nnsb_ctx *ctx = nullptr;
  const char *src = nullptr;
  size_t total = 0, chunk = 0, n_chunks = 0;
  void *slot[RING] = { nullptr, nullptr, nullptr, nullptr };
  nnsb_event *ev[RING] = { nullptr, nullptr, nullptr, nullptr };
  std::mutex mu;
  std::condition_variable cv;
  std::vector<char> staged;             /* chunk k sits in slot k % RING */
  size_t issued = 0;                    /* chunks whose H2D has been queued (event recorded) */
  bool abort = false;
  std::vector<std::thread> workers;

  size_t bytes_of (size_t k) const { return (k + 1) * chunk <= total ? chunk : total - k * chunk; }

  void work (size_t first, size_t step)
  {
    for (size_t k = first; k < n_chunks; k += step) {
      if (k >= RING) {
        /* the H2D of chunk k - RING read this slot: wait until it has been queued, then done */
        {
          std::unique_lock<std::mutex> lk (mu);
          cv.wait (lk, [&] { return abort || issued > k - RING; });
          if (abort)
            return;
        }
        if (nnsb_event_sync (ev[k % RING]) != NNSB_OK) {
          std::lock_guard<std::mutex> lk (mu);
          abort = true;
          cv.notify_all ();
          return;
        }
      }
      memcpy (slot[k % RING], src + k * chunk, bytes_of (k));
      {
        std::lock_guard<std::mutex> lk (mu);
        staged[k] = 1;
      }
      cv.notify_all ();
    }
  }

  int start (nnsb_ctx * c, const char *source, size_t total_bytes, size_t chunk_bytes)
  {
    ctx = c;
    src = source;
    total = total_bytes;
    chunk = chunk_bytes;
    n_chunks = (total + chunk - 1) / chunk;
    staged.assign (n_chunks, 0);
    for (int i = 0; i < RING; i++) {
      int r = nnsb_pinned_alloc (ctx, chunk, &slot[i]);
      if (!r)
        r = nnsb_event_create (ctx, &ev[i]);
      if (r)
        return r;
    }
    size_t n_threads = 2;
    const char *env = getenv ("NNSB_STAGING_THREADS");
    if (env && atoi (env) > 0)
      n_threads = (size_t) atoi (env);
    if (n_threads > RING)
      n_threads = RING;
    if (n_threads > n_chunks)
      n_threads = n_chunks;
    for (size_t i = 0; i < n_threads; i++)
      workers.emplace_back (&Stager::work, this, i, n_threads);
    return NNSB_OK;
  }

  /* blocks until chunk k is in its slot; nullptr after a failure */
  const void *wait_chunk (size_t k)
  {
    std::unique_lock<std::mutex> lk (mu);
    cv.wait (lk, [&] { return abort || staged[k]; });
    return abort ? nullptr : slot[k % RING];
  }

  /* the H2D of chunk k has just been queued on `s` */
  int chunk_issued (size_t k, nnsb_stream * s)
  {
    int r = nnsb_event_record (ev[k % RING], s);
    std::lock_guard<std::mutex> lk (mu);
    if (r)
      abort = true;
    issued = k + 1;
    cv.notify_all ();
    return r;
  }

  void finish (bool failed)
  {
    {
      std::lock_guard<std::mutex> lk (mu);
      if (failed)
        abort = true;
    }
    cv.notify_all ();
    for (auto & w : workers)
      w.join ();
    workers.clear ();
    for (int i = 0; i < RING; i++) {
      if (ev[i])
        nnsb_event_destroy (ev[i]);
      if (slot[i])
        nnsb_pinned_free (ctx, slot[i]);
      ev[i] = nullptr;
      slot[i] = nullptr;
    }
  }
};

/* One tensor whose input and output live in host memory: upload, transform, download.
 * Element-wise modes are split into chunks handled alternately by two streams, so the H2D
 * copy of chunk k+1, the kernel of chunk k and the D2H copy of chunk k-1 overlap (PCIe is
 * full duplex: wall time ~ max (H2D, D2H) instead of their sum).  A pageable source goes
 * through the pinned staging ring above. */
static int
run_host_tensor (nnsb_transform * t, nnsb_plan * plan, const char *h_in, char *h_out)
{
  const size_t in_size = nnsb_plan_in_size (plan), out_size = nnsb_plan_out_size (plan);
  void *d_in = nullptr, *d_out = nullptr;
  int r = nnsb_dev_alloc (t->ctx, in_size, &d_in);
  if (r)
    return r;
  r = nnsb_dev_alloc (t->ctx, out_size, &d_out);
  if (r) {
    nnsb_dev_free (t->ctx, d_in);
    return r;
  }
  nnsb_tensor_info oi;
  nnsb_plan_out_info (plan, &oi);
  const size_t es_out = nnsb_element_size (oi.type);
  const uint64_t n = es_out ? out_size / es_out : 0;
  const size_t es_in = n ? in_size / n : 1;
  const size_t pipeline_min = 8u << 20;
  /* small tensors: the driver's own staging of a pageable source is as good as ours */
  const bool stage = in_size >= (4u << 20) && !nnsb_host_is_pinned (h_in);
  Stager stager;

  if (nnsb_plan_is_elementwise (plan) && in_size + out_size >= pipeline_min
      && ensure_second_stream (t) == NNSB_OK) {
    const uint64_t g = nnsb_plan_chunk_granule (plan);
    /* ~16 chunks, each a multiple of the granule, at least 1 MiB of traffic */
    uint64_t chunk = (n / 16 + g - 1) / g * g;
    const uint64_t min_chunk = ((1u << 20) / (es_in + es_out) + g - 1) / g * g;
    if (chunk < min_chunk)
      chunk = min_chunk;
    if (stage)
      r = stager.start (t->ctx, h_in, in_size, chunk * es_in);
    int k = 0;
    size_t ci = 0;
    for (uint64_t first = 0; first < n && !r; first += chunk, k ^= 1, ci++) {
      const uint64_t cnt = (n - first < chunk) ? n - first : chunk;
      nnsb_stream *s = t->stream[k];
      const void *src = h_in + first * es_in;
      if (stage) {
        src = stager.wait_chunk (ci);
        if (!src) {
          r = nnsb_set_error (NNSB_ECUDA, "tensor_transform: staging of a pageable input failed");
          break;
        }
      }
      r = nnsb_h2d_async (s, (char *) d_in + first * es_in, src, cnt * es_in);
      if (!r && stage)
        r = stager.chunk_issued (ci, s);
      if (!r)
        r = nnsb_plan_launch_elems (plan, d_in, d_out, first, cnt, s);
      if (!r)
        r = nnsb_d2h_async (s, h_out + first * es_out, (char *) d_out + first * es_out, cnt * es_out);
      t->launches += 1;
    }
    int r2 = nnsb_stream_sync (t->stream[0]);
    int r3 = nnsb_stream_sync (t->stream[1]);
    if (!r)
      r = r2 ? r2 : r3;
  } else {
    nnsb_stream *s = t->stream[0];
    if (stage) {
      const size_t piece = 4u << 20;
      r = stager.start (t->ctx, h_in, in_size, piece);
      for (size_t ci = 0; ci * piece < in_size && !r; ci++) {
        const void *src = stager.wait_chunk (ci);
        if (!src) {
          r = nnsb_set_error (NNSB_ECUDA, "tensor_transform: staging of a pageable input failed");
          break;
        }
        r = nnsb_h2d_async (s, (char *) d_in + ci * piece, src, stager.bytes_of (ci));
        if (!r)
          r = stager.chunk_issued (ci, s);
      }
    } else {
      r = nnsb_h2d_async (s, d_in, h_in, in_size);
    }
    if (!r)
      r = nnsb_plan_launch (plan, d_in, d_out, 1, s);
    if (!r) {
      t->launches += (uint64_t) nnsb_plan_last_launch_count (plan);
      r = nnsb_d2h_async (s, h_out, d_out, out_size);
    }
    int r2 = nnsb_stream_sync (s);
    if (!r)
      r = r2;
  }
  if (stage)
    stager.finish (r != NNSB_OK);
  nnsb_dev_free (t->ctx, d_in);
  nnsb_dev_free (t->ctx, d_out);
  return r;
}

extern "C" int
nnsb_transform_transform (nnsb_transform * t, const nnsb_memory * in, uint32_t n_in,
    nnsb_memory * out, uint32_t * n_out)
{
  if (!t || !in || !out || !n_out)
    return NNSB_FLOW_ERROR;
  *n_out = 0;
  if (!t->loaded) {             /* g_return_val_if_fail (filter->loaded, GST_FLOW_ERROR), :1687 */
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_transform: mode/option are not loaded");
    return NNSB_FLOW_ERROR;
  }
  if (!t->caps_set) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_transform: caps are not negotiated");
    return NNSB_FLOW_ERROR;
  }
  if (ensure_device (t) != NNSB_OK)
    return NNSB_FLOW_ERROR;

  const bool in_flexible = t->in_config.format == NNSB_FORMAT_FLEXIBLE;
  const bool out_flexible = t->out_config.format == NNSB_FORMAT_FLEXIBLE;

  /* gst_tensor_buffer_from_config (nnstreamer_plugin_api_impl.c:452-556): one blob holding
   * several static tensors is split into per-tensor memories */
  std::vector<nnsb_memory> mems (in, in + n_in);
  if (!in_flexible && n_in == 1 && t->in_config.num_tensors > 1) {
    size_t off = 0;
    mems.clear ();
    for (uint32_t i = 0; i < t->in_config.num_tensors; i++) {
      nnsb_memory m = in[0];            /* a borrowed view of the blob (same block) */
      const size_t sz = nnsb_info_size (&t->in_config.info[i]);
      m.data = (char *) in[0].data + off;
      m.size = sz;
      m.owned = 0;
      off += sz;
      mems.push_back (m);
    }
    if (off > in[0].size) {
      nnsb_set_error (NNSB_FLOW_ERROR, "tensor_transform: input buffer is smaller than the configured tensors");
      return NNSB_FLOW_ERROR;
    }
  }

  uint32_t num_tensors;
  if (in_flexible) {
    num_tensors = (uint32_t) mems.size ();
    if (!out_flexible)
      return NNSB_FLOW_ERROR;   /* :1697 */
  } else {
    num_tensors = t->in_config.num_tensors;
    if (mems.size () != num_tensors) {
      nnsb_set_error (NNSB_FLOW_ERROR, "tensor_transform: %zu memories for %u tensors", mems.size (), num_tensors);
      return NNSB_FLOW_ERROR;
    }
  }
  if (num_tensors > NNSB_SIZE_LIMIT)
    return NNSB_FLOW_ERROR;

  int res = NNSB_FLOW_OK;
  uint32_t produced = 0;
  /* device tensors of this buffer: everything queued on stream[0] is covered by ONE event after
   * the loop (outputs ready at it, inputs not recycled before it) -- no host synchronisation */
  nnsb_memory dev_ins[NNSB_SIZE_LIMIT];
  uint32_t dev_outs[NNSB_SIZE_LIMIT];
  uint32_t n_dev = 0;
  for (uint32_t i = 0; i < num_tensors && res == NNSB_FLOW_OK; i++) {
    nnsb_tensor_info in_info = t->in_config.info[i], out_info = t->out_config.info[i];
    const nnsb_memory & im = mems[i];

    if (im.on_device && nnsb_memory_wait (&im, t->stream[0]) != NNSB_OK) {
      res = NNSB_FLOW_ERROR;
      break;
    }

    if (!in_apply (t, i)) {
      /* :1708-1721 */
      if (!in_flexible && out_flexible) {
        /* gst_tensor_meta_info_append_header: header + copy of the payload */
        nnsb_meta_info meta;
        nnsb_memory om;
        if (nnsb_meta_from_info (&out_info, &meta) != NNSB_OK
            || nnsb_transform_alloc (t, im.size + NNSB_META_HEADER_SIZE, im.on_device, &om) != NNSB_OK) {
          res = NNSB_FLOW_ERROR;
          break;
        }
        unsigned char hdr[NNSB_META_HEADER_SIZE];
        nnsb_meta_write_header (&meta, hdr);
        if (im.on_device) {
          nnsb_h2d_async (t->stream[0], om.data, hdr, sizeof (hdr));
          nnsb_d2d_async (t->stream[0], (char *) om.data + sizeof (hdr), im.data, im.size);
          nnsb_stream_sync (t->stream[0]);
        } else {
          memcpy (om.data, hdr, sizeof (hdr));
          memcpy ((char *) om.data + sizeof (hdr), im.data, im.size);
        }
        out[produced++] = om;
      } else {
        /* gst_memory_ref (:1716): another reference to the same block -- for a tensor that sits
         * inside a multi-tensor blob, a share of its byte range */
        nnsb_memory shared;
        if (nnsb_memory_share (&im, 0, im.size, &shared) != NNSB_OK) {
          res = NNSB_FLOW_ERROR;
          break;
        }
        out[produced++] = shared;
      }
      continue;
    }

    const char *inptr = (const char *) im.data;
    size_t in_avail = im.size;
    unsigned char hdr_in[NNSB_META_HEADER_SIZE];
    if (in_flexible) {
      /* :1732-1748 */
      if (im.size < NNSB_META_HEADER_SIZE) {
        res = NNSB_FLOW_ERROR;
        break;
      }
      if (im.on_device) {
        if (nnsb_d2h_async (t->stream[0], hdr_in, im.data, sizeof (hdr_in)) != NNSB_OK
            || nnsb_stream_sync (t->stream[0]) != NNSB_OK) {
          res = NNSB_FLOW_ERROR;
          break;
        }
      } else {
        memcpy (hdr_in, im.data, sizeof (hdr_in));
      }
      nnsb_meta_info meta;
      if (nnsb_meta_parse_header (&meta, hdr_in) != NNSB_OK || nnsb_meta_to_info (&meta, &in_info) != NNSB_OK) {
        res = NNSB_FLOW_ERROR;
        break;
      }
      convert_info (t, NNSB_PAD_SINK, i, &in_info, &out_info);
      inptr += NNSB_META_HEADER_SIZE;
      in_avail -= NNSB_META_HEADER_SIZE;
    }

    nnsb_plan *plan = plan_for (t, i, &in_info);
    if (!plan) {
      /* dimchg from > to (:1179-1188) ends up here: flow error at transform time */
      res = NNSB_FLOW_ERROR;
      break;
    }
    nnsb_plan_out_info (plan, &out_info);
    const size_t in_size = nnsb_plan_in_size (plan), out_size = nnsb_plan_out_size (plan);
    if (in_avail < in_size) {
      nnsb_set_error (NNSB_FLOW_ERROR, "tensor_transform: tensor %u has %zu bytes, %zu expected", i, in_avail, in_size);
      res = NNSB_FLOW_ERROR;
      break;
    }
    const size_t hsize = out_flexible ? NNSB_META_HEADER_SIZE : 0;
    nnsb_memory om;
    if ((im.on_device ? alloc_device_output (t, out_size + hsize, &om)
            : nnsb_transform_alloc (t, out_size + hsize, 0, &om)) != NNSB_OK) {
      res = NNSB_FLOW_ERROR;
      break;
    }
    out[produced++] = om;
    char *outptr = (char *) om.data;
    if (out_flexible) {
      /* :1752-1771 */
      nnsb_meta_info meta;
      unsigned char hdr[NNSB_META_HEADER_SIZE];
      if (nnsb_meta_from_info (&out_info, &meta) != NNSB_OK || nnsb_meta_write_header (&meta, hdr) != NNSB_OK) {
        res = NNSB_FLOW_ERROR;
        break;
      }
      if (im.on_device) {
        nnsb_h2d_async (t->stream[0], outptr, hdr, sizeof (hdr));
        nnsb_stream_sync (t->stream[0]);
      } else {
        memcpy (outptr, hdr, sizeof (hdr));
      }
      outptr += hsize;
    }

    int r;
    if (im.on_device) {
      r = nnsb_plan_launch (plan, inptr, outptr, 1, t->stream[0]);
      if (!r) {
        t->launches += (uint64_t) nnsb_plan_last_launch_count (plan);
        dev_ins[n_dev] = im;
        dev_outs[n_dev] = produced - 1;
        n_dev++;
      }
    } else {
      r = run_host_tensor (t, plan, inptr, outptr);
    }
    if (r != NNSB_OK)
      res = NNSB_FLOW_ERROR;
  }
  if (n_dev) {
    nnsb_memory marked[NNSB_SIZE_LIMIT];
    for (uint32_t k = 0; k < n_dev; k++)
      marked[k] = out[dev_outs[k]];
    if (nnsb_memory_mark_io (t->stream[0], marked, n_dev, dev_ins, n_dev) != NNSB_OK)
      res = NNSB_FLOW_ERROR;
  }
  *n_out = produced;
  if (res != NNSB_FLOW_OK) {
    /* the failed buffer's blocks must be idle before they go back to the pool */
    nnsb_stream_sync (t->stream[0]);
    for (uint32_t i = 0; i < produced; i++)
      nnsb_memory_release (&out[i]);
    *n_out = 0;
  }
  return res;
}

/* ------------------------------------------------------------------------------------- */
/* tensor_converter (video)                                                                */
/* ------------------------------------------------------------------------------------- */
struct nnsb_converter
{
  int device_id = 0;
  nnsb_ctx *ctx = nullptr;
  nnsb_stream *stream = nullptr;
  bool silent = true;
  uint32_t frames_per_tensor = 1;
  bool configured = false;
  nnsb_tensors_config config;
  bool remove_padding = false;
  size_t frame_size = 0;                /* GST_VIDEO_INFO_SIZE: padded */
  size_t row_bytes = 0, pitch = 0, rows = 0;
  size_t dense_frame = 0;               /* C*W*H */
  nnsb_memory staging;                  /* padded host frames land here before the strip kernel */
  nnsb_memory batch;                    /* current batch buffer (device block), data == NULL: none */
  uint32_t batch_cap = 0;               /* frames the block was allocated for */
  uint32_t filled = 0;
};

static void
converter_drop_batch (nnsb_converter * c)
{
  if (c->batch.data) {
    if (c->stream)
      nnsb_stream_sync (c->stream);
    nnsb_memory_release (&c->batch);
  }
  memset (&c->batch, 0, sizeof (c->batch));
  c->batch_cap = 0;
  c->filled = 0;
}

extern "C" int
nnsb_converter_new (int device_id, nnsb_converter ** out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_converter_new: null argument");
  nnsb_converter *c = new (std::nothrow) nnsb_converter ();
  if (!c)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_converter_new: out of memory");
  c->device_id = device_id;
  config_init (&c->config);
  memset (&c->batch, 0, sizeof (c->batch));
  memset (&c->staging, 0, sizeof (c->staging));
  *out = c;
  return NNSB_OK;
}

extern "C" int
nnsb_converter_free (nnsb_converter * c)
{
  if (!c)
    return NNSB_OK;
  converter_drop_batch (c);
  if (c->stream)
    nnsb_stream_sync (c->stream);
  nnsb_memory_release (&c->staging);
  if (c->stream)
    nnsb_stream_destroy (c->stream);
  if (c->ctx)
    nnsb_ctx_destroy (c->ctx);
  delete c;
  return NNSB_OK;
}

extern "C" int
nnsb_converter_set_property (nnsb_converter * c, const char *name, const char *value)
{
  if (!c || !name || !value)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_converter_set_property: null argument");
  const std::string n (name);
  if (n == "frames-per-tensor") {
    char *end = nullptr;
    long v = strtol (value, &end, 10);
    if (end == value || *end || v < 1)
      return nnsb_set_error (NNSB_EINVAL, "frames-per-tensor: expected an unsigned integer >= 1");
    const uint32_t fpt = (uint32_t) v;
    if (fpt != c->frames_per_tensor && c->batch.data && c->filled) {
      /* A batch is partly filled and was sized for the old count.  The reference's GstAdapter
       * just keeps the bytes (gsttensor_converter.c:945-1009); here the frames move into a block
       * of the new size.  If more frames are waiting than the new tensor can take together with
       * the next one, the oldest fpt-1 are kept (the reference would emit several tensors from
       * one chain call, which this one-output-per-call interface cannot). */
      const uint32_t keep = c->filled < fpt ? c->filled : fpt - 1;
      nnsb_memory fresh;
      memset (&fresh, 0, sizeof (fresh));
      int r = NNSB_OK;
      if (keep) {
        r = nnsb_memory_alloc (c->ctx, c->dense_frame * fpt, 1, c->stream, &fresh);
        if (!r)
          r = nnsb_d2d_async (c->stream, fresh.data, c->batch.data, c->dense_frame * keep);
        if (!r)
          r = nnsb_stream_sync (c->stream);
        if (r) {
          nnsb_memory_release (&fresh);
          return r;
        }
      }
      converter_drop_batch (c);
      c->batch = fresh;
      c->batch_cap = keep ? fpt : 0;
      c->filled = keep;
    } else if (fpt != c->frames_per_tensor && c->batch.data) {
      converter_drop_batch (c);
    }
    c->frames_per_tensor = fpt;
    if (c->configured)
      c->config.info[0].dim[3] = c->frames_per_tensor;
    return NNSB_OK;
  }
  if (n == "silent") {
    bool b;
    if (!parse_bool (value, &b))
      return nnsb_set_error (NNSB_EINVAL, "silent: expected a boolean");
    c->silent = b;
    return NNSB_OK;
  }
  return nnsb_set_error (NNSB_EINVAL, "tensor_converter (video) has no property '%s'", name);
}

extern "C" int
nnsb_converter_get_property (nnsb_converter * c, const char *name, char *buf, size_t buflen)
{
  if (!c || !name)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_converter_get_property: null argument");
  const std::string n (name);
  if (n == "frames-per-tensor")
    return copy_out (std::to_string (c->frames_per_tensor), buf, buflen);
  if (n == "silent")
    return copy_out (c->silent ? "true" : "false", buf, buflen);
  return nnsb_set_error (NNSB_EINVAL, "tensor_converter (video) has no property '%s'", name);
}

extern "C" int
nnsb_converter_set_caps (nnsb_converter * c, const char *video_caps, char *out, size_t outlen)
{
  if (!c || !video_caps)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_converter_set_caps: null argument");
  Caps caps;
  std::string err;
  if (!caps_parse (video_caps, &caps, &err) || caps.structures.empty ())
    return nnsb_set_error (NNSB_EINVAL, "cannot parse caps '%s': %s", video_caps, err.c_str ());
  if (!caps.is_fixed ())
    return nnsb_set_error (NNSB_EINVAL, "caps must be fixed");
  const CapsStructure & st = caps.structures[0];
  if (st.name != "video/x-raw")
    return nnsb_set_error (NNSB_ENOTSUP, "only video/x-raw is handled by this converter (got %s)", st.name.c_str ());
  const char *fmt = st.get_string ("format");
  const CapsValue *w = st.get ("width"), *h = st.get ("height");
  if (!fmt || !w || !h || w->kind != CapsValue::INT || h->kind != CapsValue::INT || w->i0 < 1 || h->i0 < 1)
    return nnsb_set_error (NNSB_EINVAL, "Failed to get video info from caps '%s'", video_caps);
  /* gst_tensor_converter_parse_video, :1421-1519 */
  const std::string f (fmt);
  uint32_t color = 0;
  if (f == "GRAY8")
    color = 1;
  else if (f == "RGB" || f == "BGR")
    color = 3;
  else if (f == "RGBx" || f == "BGRx" || f == "xRGB" || f == "xBGR" || f == "RGBA" || f == "BGRA"
      || f == "ARGB" || f == "ABGR")
    color = 4;
  else
    return nnsb_set_error (NNSB_ENOTSUP,
        "The given video caps with format \"%s\" is not supported. Please use GRAY8, RGB, BGR, RGBx, BGRx, xRGB, xBGR, RGBA, BGRA, ARGB, or ABGR.",
        fmt);
  config_init (&c->config);
  c->config.num_tensors = 1;
  c->config.info[0].type = NNSB_UINT8;
  c->config.info[0].dim[0] = color;
  c->config.info[0].dim[1] = (uint32_t) w->i0;
  c->config.info[0].dim[2] = (uint32_t) h->i0;
  for (int i = 3; i < NNSB_RANK_LIMIT; i++)
    c->config.info[0].dim[i] = 1;
  c->config.info[0].dim[3] = c->frames_per_tensor;      /* :2132,:2204-2206 */
  const CapsValue *fr = st.get ("framerate");
  if (fr && fr->kind == CapsValue::FRACTION) {
    c->config.rate_n = (int32_t) fr->i0;
    c->config.rate_d = (int32_t) fr->i1;
  } else {
    c->config.rate_n = 0;
    c->config.rate_d = 1;
  }
  /* gst_tensor_converter_video_stride, :1390-1411 (I420 is rejected above) */
  c->remove_padding = (color == 1 || color == 3) && ((uint32_t) w->i0 % 4 != 0);
  c->row_bytes = (size_t) color * (size_t) w->i0;
  c->pitch = (c->row_bytes + 3) / 4 * 4;
  c->rows = (size_t) h->i0;
  c->frame_size = c->pitch * c->rows;                   /* GST_VIDEO_INFO_SIZE for packed RGB / GRAY8 */
  c->dense_frame = c->row_bytes * c->rows;
  c->configured = true;
  converter_drop_batch (c);             /* new caps: frames of the old geometry are gone */
  if (!config_validate (c->config))
    return nnsb_set_error (NNSB_EINVAL, "Failed to configure tensors info with gst cap (%s)", video_caps);
  if (out)
    return copy_out (caps_to_string (tensors_caps_from_config (c->config)), out, outlen);
  return NNSB_OK;
}

extern "C" int
nnsb_converter_get_config (nnsb_converter * c, nnsb_tensors_config * cfg)
{
  if (!c || !cfg || !c->configured)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_converter_get_config: not configured");
  *cfg = c->config;
  return NNSB_OK;
}

extern "C" int
nnsb_converter_remove_padding (nnsb_converter * c)
{
  return (c && c->configured && c->remove_padding) ? 1 : 0;
}

extern "C" size_t
nnsb_converter_frame_size (nnsb_converter * c)
{
  return (c && c->configured) ? c->frame_size : 0;
}

extern "C" int
nnsb_converter_chain (nnsb_converter * c, const nnsb_memory * frame, nnsb_memory * out, int *produced)
{
  if (!c || !frame || !out || !produced)
    return NNSB_FLOW_ERROR;
  *produced = 0;
  if (!c->configured) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_converter: caps are not set");
    return NNSB_FLOW_ERROR;
  }
  if (!c->ctx) {
    if (nnsb_ctx_create (c->device_id, &c->ctx) != NNSB_OK || nnsb_stream_create (c->ctx, &c->stream) != NNSB_OK)
      return NNSB_FLOW_ERROR;
  }
  /* "supposed 1 frame in buffer", :1060 */
  if (frame->size != c->frame_size) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_converter: frame of %zu bytes, %zu expected", frame->size, c->frame_size);
    return NNSB_FLOW_ERROR;
  }
  if (c->frames_per_tensor == 1 && frame->on_device && !c->remove_padding) {
    /* nothing to strip, nothing to batch: the frame IS the tensor (the reference pushes the
     * incoming buffer itself when remove_padding is FALSE, gsttensor_converter.c:1062,1107-1109) */
    if (nnsb_memory_ref (frame, out) != NNSB_OK)
      return NNSB_FLOW_ERROR;
    *produced = 1;
    return NNSB_FLOW_OK;
  }
  const size_t batch_bytes = c->dense_frame * c->frames_per_tensor;
  if (!c->batch.data) {
    if (nnsb_memory_alloc (c->ctx, batch_bytes, 1, c->stream, &c->batch) != NNSB_OK)
      return NNSB_FLOW_ERROR;
    c->batch_cap = c->frames_per_tensor;
    c->filled = 0;
  }
  if (c->filled >= c->batch_cap || c->batch_cap != c->frames_per_tensor) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_converter: batch buffer of %u frames cannot take frame %u",
        c->batch_cap, c->filled);
    return NNSB_FLOW_ERROR;
  }
  char *dst = (char *) c->batch.data + (size_t) c->filled * c->dense_frame;
  int r = NNSB_OK;
  if (frame->on_device)
    r = nnsb_memory_wait (frame, c->stream);
  if (r != NNSB_OK)
    return NNSB_FLOW_ERROR;
  if (c->remove_padding) {
    /* the row loop of :1087-1100, done by the copy engine (host frame) or K8 (device frame) */
    if (frame->on_device) {
      r = nnsb_strip_launch (c->stream, frame->data, dst, c->row_bytes, c->pitch, c->rows);
    } else {
      /* a pitched H2D copy runs at ~60 % of the link (one DMA descriptor per row); a linear copy
       * of the padded frame followed by the strip kernel keeps PCIe at full speed and costs only
       * HBM traffic.  The staging block is reused frame after frame on this one stream. */
      if (c->staging.data && c->staging.size != c->frame_size) {
        nnsb_stream_sync (c->stream);           /* new caps: the old block must be idle */
        nnsb_memory_release (&c->staging);
      }
      if (!c->staging.data)
        r = nnsb_memory_alloc (c->ctx, c->frame_size, 1, c->stream, &c->staging);
      if (r == NNSB_OK)
        r = nnsb_h2d_async (c->stream, c->staging.data, frame->data, c->frame_size);
      if (r == NNSB_OK)
        r = nnsb_strip_launch (c->stream, c->staging.data, dst, c->row_bytes, c->pitch, c->rows);
    }
  } else {
    r = frame->on_device ? nnsb_d2d_async (c->stream, dst, frame->data, c->dense_frame)
        : nnsb_h2d_async (c->stream, dst, frame->data, c->dense_frame);
  }
  if (r != NNSB_OK)
    return NNSB_FLOW_ERROR;
  c->filled++;
  const bool complete = c->filled == c->frames_per_tensor;
  if (frame->on_device) {
    /* stream-ordered: the frame's block is not recycled before the copy has read it, and the
     * batch becomes ready with its last frame */
    if (nnsb_memory_mark_io (c->stream, &c->batch, complete ? 1 : 0, frame, 1) != NNSB_OK)
      return NNSB_FLOW_ERROR;
    if (!frame->block && nnsb_stream_sync (c->stream) != NNSB_OK)
      return NNSB_FLOW_ERROR;   /* caller-managed device frame: it may be reused on return */
  } else {
    /* a host frame may be reused by the caller as soon as we return */
    if (nnsb_stream_sync (c->stream) != NNSB_OK)
      return NNSB_FLOW_ERROR;
  }
  if (complete) {
    *out = c->batch;
    out->size = batch_bytes;
    memset (&c->batch, 0, sizeof (c->batch));
    c->batch_cap = 0;
    c->filled = 0;
    *produced = 1;
  }
  return NNSB_FLOW_OK;
}

/* ------------------------------------------------------------------------------------- */
/* batching queue                                                                          */
/* ------------------------------------------------------------------------------------- */
struct nnsb_queue
{
  nnsb_ctx *ctx;
  nnsb_plan *plan;
  uint32_t batch;
  size_t in_size, out_size;
  void *d_in[2], *d_out[2];
  int cur;                              /* buffer being filled */
  uint32_t filled;
  nnsb_stream *copy_stream, *compute_stream;
  nnsb_event *copied[2], *computed[2];
  bool computed_valid[2];
};

extern "C" int
nnsb_queue_destroy (nnsb_queue * q)
{
  if (!q)
    return NNSB_OK;
  if (q->copy_stream)
    nnsb_stream_sync (q->copy_stream);
  if (q->compute_stream)
    nnsb_stream_sync (q->compute_stream);
  for (int i = 0; i < 2; i++) {
    if (q->d_in[i])
      nnsb_dev_free (q->ctx, q->d_in[i]);
    if (q->d_out[i])
      nnsb_dev_free (q->ctx, q->d_out[i]);
    if (q->copied[i])
      nnsb_event_destroy (q->copied[i]);
    if (q->computed[i])
      nnsb_event_destroy (q->computed[i]);
  }
  if (q->copy_stream)
    nnsb_stream_destroy (q->copy_stream);
  if (q->compute_stream)
    nnsb_stream_destroy (q->compute_stream);
  delete q;
  return NNSB_OK;
}

extern "C" int
nnsb_queue_create (nnsb_ctx * ctx, nnsb_plan * plan, uint32_t batch, nnsb_queue ** out)
{
  if (!ctx || !plan || !out || batch == 0)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_queue_create: invalid argument");
  nnsb_queue *q = new (std::nothrow) nnsb_queue ();
  if (!q)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_queue_create: out of memory");
  memset (q, 0, sizeof (*q));
  q->ctx = ctx;
  q->plan = plan;
  q->batch = batch;
  q->in_size = nnsb_plan_in_size (plan);
  q->out_size = nnsb_plan_out_size (plan);
  int r = nnsb_stream_create (ctx, &q->copy_stream);
  if (!r)
    r = nnsb_stream_create (ctx, &q->compute_stream);
  for (int i = 0; i < 2 && !r; i++) {
    r = nnsb_dev_alloc (ctx, q->in_size * batch, &q->d_in[i]);
    if (!r)
      r = nnsb_dev_alloc (ctx, q->out_size * batch, &q->d_out[i]);
    if (!r)
      r = nnsb_event_create (ctx, &q->copied[i]);
    if (!r)
      r = nnsb_event_create (ctx, &q->computed[i]);
  }
  if (r) {
    nnsb_queue_destroy (q);
    return r;
  }
  *out = q;
  return NNSB_OK;
}

extern "C" int
nnsb_queue_push (nnsb_queue * q, const void *frame, int on_device, int *full)
{
  if (!q || !frame)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_queue_push: null argument");
  if (q->filled >= q->batch)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_queue_push: queue is full, flush it first");
  if (q->filled == 0 && q->computed_valid[q->cur]) {
    /* the launch that last read this input buffer must be done before it is overwritten */
    int r = nnsb_stream_wait_event (q->copy_stream, q->computed[q->cur]);
    if (r)
      return r;
  }
  char *dst = (char *) q->d_in[q->cur] + (size_t) q->filled * q->in_size;
  int r = on_device ? nnsb_d2d_async (q->copy_stream, dst, frame, q->in_size)
      : nnsb_h2d_async (q->copy_stream, dst, frame, q->in_size);
  if (r)
    return r;
  q->filled++;
  if (full)
    *full = q->filled == q->batch;
  return NNSB_OK;
}

extern "C" int
nnsb_queue_flush (nnsb_queue * q, void **d_out, uint32_t * n_frames, nnsb_stream ** stream)
{
  if (!q || !d_out || !n_frames)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_queue_flush: null argument");
  *d_out = nullptr;
  *n_frames = 0;
  if (stream)
    *stream = q->compute_stream;
  if (q->filled == 0)
    return NNSB_OK;
  const int b = q->cur;
  int r = nnsb_event_record (q->copied[b], q->copy_stream);
  if (!r)
    r = nnsb_stream_wait_event (q->compute_stream, q->copied[b]);
  if (!r)
    r = nnsb_plan_launch (q->plan, q->d_in[b], q->d_out[b], q->filled, q->compute_stream);
  if (!r)
    r = nnsb_event_record (q->computed[b], q->compute_stream);
  if (r)
    return r;
  q->computed_valid[b] = true;
  *d_out = q->d_out[b];
  *n_frames = q->filled;
  q->cur ^= 1;
  q->filled = 0;
  return NNSB_OK;
}
