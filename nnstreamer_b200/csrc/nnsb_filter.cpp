/*
 * nnsb_filter.cpp -- the elements either side of the hot path that only hand memories on
 * (SURVEY.md 8f ranks 2 and 4): nothing here computes; what matters is that tensors stay in
 * HBM and stay stream-ordered while they change hands.
 *
 *   nnsb_filter  <->  GstTensorFilter's transform (gst/nnstreamer/tensor_filter/
 *                     tensor_filter.c:632-890) with a CUDA-aware sub-plugin behind it: the
 *                     sub-plugin's v1 `invoke` (nnstreamer_plugin_api_filter.h:419-421) receives
 *                     DEVICE pointers and the CUDA stream to enqueue on, instead of the host
 *                     pointers gst_memory_map hands out at :675 / :759.  This is what
 *                     ext/nnstreamer/tensor_filter/tensor_filter_tensorrt.cc wants: it stages
 *                     through cudaMallocManaged (:390) only because its inputs arrive on the host.
 *   nnsb_mux     <->  GstTensorMux (elements/gsttensor_mux.c): the memories of all sink pads,
 *                     in pad order, become one other/tensors buffer -- references, no copies
 *                     (gst_buffer_append_memory of the collected memories, :467,482).
 *   nnsb_demux   <->  GstTensorDemux (elements/gsttensor_demux.c): `tensorpick` groups
 *                     ("0,1:2,2+0", :600-620) -> one buffer per source pad holding references
 *                     to the picked memories (:516-522); a single blob carrying several static
 *                     tensors is first cut into shares (gst_tensor_buffer_from_config,
 *                     nnstreamer_plugin_api_impl.c:452-556, gst_memory_share :534).
 *
 * Not mirrored: tensor_mux's collect-pads time synchronisation (sync-mode / sync-option are
 * stored and returned); the filter's model loading, framework registry, combination
 * properties, latency/throughput statistics and shared-model table.
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

int
copy_str (const std::string & s, char *out, size_t outlen)
{
  if (!out || s.size () + 1 > outlen)
    return nnsb_set_error (NNSB_EINVAL, "output buffer too small (%zu needed)", s.size () + 1);
  memcpy (out, s.c_str (), s.size () + 1);
  return (int) s.size ();
}

bool
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

/* g_strsplit_set (param, set, -1) */
std::vector<std::string>
split_set (const std::string & s, const char *set)
{
  std::vector<std::string> out;
  size_t start = 0;
  for (;;) {
    const size_t p = s.find_first_of (set, start);
    out.push_back (s.substr (start, p == std::string::npos ? std::string::npos : p - start));
    if (p == std::string::npos)
      break;
    start = p + 1;
  }
  return out;
}

Caps
pad_caps (const nnsb_tensors_config & c)
{
  /* gst_tensor_pad_caps_from_config without a peer: other/tensor for one static tensor */
  if (c.format == NNSB_FORMAT_STATIC && c.num_tensors == 1)
    return tensor_caps_from_config (c);
  return tensors_caps_from_config (c);
}

}  // namespace

/* ------------------------------------------------------------------------------------- */
/* tensor_filter hand-off                                                                  */
/* ------------------------------------------------------------------------------------- */
struct nnsb_filter
{
  int device_id = 0;
  nnsb_ctx *ctx = nullptr;
  nnsb_stream *stream = nullptr;
  bool silent = true;
  nnsb_filter_invoke_fn invoke = nullptr;
  void *private_data = nullptr;
  bool configured = false;
  nnsb_tensors_config in_config, out_config;    /* prop->input_meta / output_meta */
  uint64_t h2d_bytes = 0, d2h_bytes = 0, invokes = 0;
};

extern "C" int
nnsb_filter_new (int device_id, nnsb_filter ** out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_filter_new: null argument");
  nnsb_filter *f = new (std::nothrow) nnsb_filter ();
  if (!f)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_filter_new: out of memory");
  f->device_id = device_id;
  config_init (&f->in_config);
  config_init (&f->out_config);
  *out = f;
  return NNSB_OK;
}

extern "C" int
nnsb_filter_free (nnsb_filter * f)
{
  if (!f)
    return NNSB_OK;
  if (f->stream)
    nnsb_stream_destroy (f->stream);
  if (f->ctx)
    nnsb_ctx_destroy (f->ctx);
  delete f;
  return NNSB_OK;
}

extern "C" int
nnsb_filter_set_framework (nnsb_filter * f, nnsb_filter_invoke_fn invoke, void *private_data)
{
  if (!f || !invoke)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_filter_set_framework: null argument");
  f->invoke = invoke;
  f->private_data = private_data;
  return NNSB_OK;
}

extern "C" int
nnsb_filter_set_info (nnsb_filter * f, const nnsb_tensors_config * in, const nnsb_tensors_config * out)
{
  if (!f || !in || !out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_filter_set_info: null argument");
  if (in->format != NNSB_FORMAT_STATIC || out->format != NNSB_FORMAT_STATIC)
    return nnsb_set_error (NNSB_ENOTSUP, "nnsb_filter_set_info: only static tensors are handed over on the device");
  if (!config_validate (*in) || !config_validate (*out))
    return nnsb_set_error (NNSB_EINVAL, "nnsb_filter_set_info: invalid tensors info");
  f->in_config = *in;
  f->out_config = *out;
  f->configured = true;
  return NNSB_OK;
}

/* gst_tensor_filter_transform, tensor_filter.c:632-890 */
extern "C" int
nnsb_filter_transform (nnsb_filter * f, const nnsb_memory * in, uint32_t n_in, nnsb_memory * out,
    uint32_t * n_out)
{
  if (!f || !in || !out || !n_out)
    return NNSB_FLOW_ERROR;
  *n_out = 0;
  /* 0. _gst_tensor_filter_transform_validate, :560-626 */
  if (!f->invoke) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_filter: no framework (invoke callback) is set");
    return NNSB_FLOW_ERROR;
  }
  if (!f->configured) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_filter: input / output tensors info is not set");
    return NNSB_FLOW_ERROR;
  }
  if (!f->ctx) {
    if (nnsb_ctx_create (f->device_id, &f->ctx) != NNSB_OK
        || nnsb_stream_create (f->ctx, &f->stream) != NNSB_OK)
      return NNSB_FLOW_ERROR;
  }
  /* 1. input tensors, :671-735 */
  if (n_in != f->in_config.num_tensors) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_filter: Input buffer has invalid number of memory "
        "blocks (%u), which is expected to be %u (the number of tensors).", n_in,
        f->in_config.num_tensors);
    return NNSB_FLOW_ERROR;
  }
  nnsb_tensor_memory in_t[NNSB_SIZE_LIMIT], out_t[NNSB_SIZE_LIMIT];
  nnsb_memory staged[NNSB_SIZE_LIMIT];
  nnsb_memory used[NNSB_SIZE_LIMIT];
  memset (staged, 0, sizeof (staged));
  int res = NNSB_FLOW_OK;
  for (uint32_t i = 0; i < n_in && res == NNSB_FLOW_OK; i++) {
    const size_t expected = nnsb_info_size (&f->in_config.info[i]);
    if (in[i].size != expected || !in[i].data) {
      nnsb_set_error (NNSB_FLOW_ERROR, "tensor_filter: Input buffer size (%u'th memory chunk: %zu) "
          "is invalid, which is expected to be %zu", i, in[i].size, expected);
      res = NNSB_FLOW_ERROR;
      break;
    }
    if (in[i].on_device) {
      /* the hand-off: no map, no copy -- the filter's stream waits for the producer's event */
      if (nnsb_memory_wait (&in[i], f->stream) != NNSB_OK)
        res = NNSB_FLOW_ERROR;
      in_t[i].data = in[i].data;
      used[i] = in[i];
    } else {
      /* a host tensor is uploaded once; the copy is counted */
      if (nnsb_memory_alloc (f->ctx, expected, 1, f->stream, &staged[i]) != NNSB_OK
          || nnsb_h2d_async (f->stream, staged[i].data, in[i].data, expected) != NNSB_OK)
        res = NNSB_FLOW_ERROR;
      f->h2d_bytes += expected;
      in_t[i].data = staged[i].data;
      used[i] = staged[i];
    }
    in_t[i].size = expected;
  }
  /* 2. output tensors, :737-778 (allocate_in_invoke == FALSE) */
  uint32_t produced = 0;
  for (uint32_t i = 0; i < f->out_config.num_tensors && res == NNSB_FLOW_OK; i++) {
    const size_t size = nnsb_info_size (&f->out_config.info[i]);
    if (nnsb_memory_alloc (f->ctx, size, 1, f->stream, &out[i]) != NNSB_OK) {
      res = NNSB_FLOW_ERROR;
      break;
    }
    produced++;
    out_t[i].data = out[i].data;
    out_t[i].size = size;
  }
  /* 3. invoke, :786 */
  int ret = 0;
  if (res == NNSB_FLOW_OK) {
    ret = f->invoke (f->private_data, in_t, n_in, out_t, f->out_config.num_tensors,
        nnsb_stream_handle (f->stream));
    f->invokes++;
    /* whatever the sub-plugin queued on the stream: outputs ready after it, inputs held until it */
    if (nnsb_memory_mark_io (f->stream, out, produced, used, n_in) != NNSB_OK)
      res = NNSB_FLOW_ERROR;
  }
  /* host inputs may be reused by the caller on return */
  bool any_staged = false;
  for (uint32_t i = 0; i < n_in; i++)
    any_staged = any_staged || staged[i].data;
  if (any_staged)
    nnsb_stream_sync (f->stream);
  for (uint32_t i = 0; i < n_in; i++)
    nnsb_memory_release (&staged[i]);
  if (res == NNSB_FLOW_OK && ret < 0) {
    nnsb_set_error (NNSB_FLOW_ERROR, "Calling invoke function (inference instance) of the "
        "tensor-filter subplugin has failed with error code (%d).", ret);
    res = NNSB_FLOW_ERROR;
  }
  if (res != NNSB_FLOW_OK || ret > 0) {
    /* error, or "drop this buffer" (:812-814) */
    nnsb_stream_sync (f->stream);
    for (uint32_t i = 0; i < produced; i++)
      nnsb_memory_release (&out[i]);
    return res != NNSB_FLOW_OK ? res : NNSB_FLOW_DROPPED;
  }
  *n_out = produced;
  return NNSB_FLOW_OK;
}

extern "C" int
nnsb_filter_copy_counters (nnsb_filter * f, uint64_t * h2d_bytes, uint64_t * d2h_bytes, uint64_t * invokes)
{
  if (!f)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_filter_copy_counters: null argument");
  if (h2d_bytes)
    *h2d_bytes = f->h2d_bytes;
  if (d2h_bytes)
    *d2h_bytes = f->d2h_bytes;
  if (invokes)
    *invokes = f->invokes;
  return NNSB_OK;
}

/* ------------------------------------------------------------------------------------- */
/* tensor_mux                                                                              */
/* ------------------------------------------------------------------------------------- */
struct nnsb_mux
{
  bool silent = true;
  std::string sync_mode = "nosync", sync_option;
  bool has_sync_option = false;
  bool configured = false;
  nnsb_tensors_config config;
};

extern "C" int
nnsb_mux_new (nnsb_mux ** out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_mux_new: null argument");
  nnsb_mux *m = new (std::nothrow) nnsb_mux ();
  if (!m)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_mux_new: out of memory");
  config_init (&m->config);
  *out = m;
  return NNSB_OK;
}

extern "C" int
nnsb_mux_free (nnsb_mux * m)
{
  delete m;
  return NNSB_OK;
}

/* gst_tensor_mux_set_property, gsttensor_mux.c:602-632 */
extern "C" int
nnsb_mux_set_property (nnsb_mux * m, const char *name, const char *value)
{
  if (!m || !name || !value)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_mux_set_property: null argument");
  const std::string n (name);
  if (n == "silent") {
    bool b;
    if (!parse_bool (value, &b))
      return nnsb_set_error (NNSB_EINVAL, "silent: expected a boolean");
    m->silent = b;
    return NNSB_OK;
  }
  if (n == "sync-mode") {
    const std::string v (value);
    if (v != "nosync" && v != "slowest" && v != "basepad" && v != "refresh")
      return nnsb_set_error (NNSB_EINVAL, "sync-mode: '%s' is not one of nosync, slowest, basepad, refresh", value);
    m->sync_mode = v;
    return NNSB_OK;
  }
  if (n == "sync-option") {
    m->sync_option = value;
    m->has_sync_option = true;
    return NNSB_OK;
  }
  return nnsb_set_error (NNSB_EINVAL, "tensor_mux has no property '%s'", name);
}

extern "C" int
nnsb_mux_get_property (nnsb_mux * m, const char *name, char *buf, size_t buflen)
{
  if (!m || !name)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_mux_get_property: null argument");
  const std::string n (name);
  if (n == "silent")
    return copy_str (m->silent ? "true" : "false", buf, buflen);
  if (n == "sync-mode")
    return copy_str (m->sync_mode, buf, buflen);
  if (n == "sync-option") {
    if (!m->has_sync_option) {
      if (buf && buflen)
        buf[0] = '\0';
      return NNSB_ENOENT;
    }
    return copy_str (m->sync_option, buf, buflen);
  }
  return nnsb_set_error (NNSB_EINVAL, "tensor_mux has no property '%s'", name);
}

/* the merged config the collect function builds (gst_tensor_time_sync_buffer_from_collectpad
 * appends every pad's tensors info in pad order) and gst_tensor_mux_set_src_caps, :390-412 */
extern "C" int
nnsb_mux_set_caps (nnsb_mux * m, uint32_t n, const char *const *caps, char *out, size_t outlen)
{
  if (!m || !caps || n < 1)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_mux_set_caps: invalid argument");
  nnsb_tensors_config merged;
  config_init (&merged);
  bool flexible = false;
  for (uint32_t k = 0; k < n; k++) {
    Caps c;
    std::string err;
    if (!caps[k] || !caps_parse (caps[k], &c, &err) || c.structures.empty () || !c.is_fixed ())
      return nnsb_set_error (NNSB_EINVAL, "cannot parse fixed caps of sink_%u", k);
    nnsb_tensors_config pc;
    if (!config_from_structure (&pc, c.structures[0]))
      return nnsb_set_error (NNSB_EINVAL, "sink_%u: not a tensor stream", k);
    if (pc.format == NNSB_FORMAT_FLEXIBLE)
      flexible = true;
    else if (!config_validate (pc))
      return nnsb_set_error (NNSB_EINVAL, "sink_%u: invalid tensors info", k);
    if (merged.num_tensors + pc.num_tensors > NNSB_SIZE_LIMIT)
      return nnsb_set_error (NNSB_ERANGE, "tensor_mux: more than %d tensors in total", NNSB_SIZE_LIMIT);
    for (uint32_t i = 0; i < pc.num_tensors; i++)
      merged.info[merged.num_tensors++] = pc.info[i];
    /* gst_tensor_time_sync_buffer_from_collectpad, nnstreamer_plugin_api_impl.c:345-348,433-434:
     * the smallest numerator and the smallest denominator seen, each on its own */
    if (k == 0 || pc.rate_d < merged.rate_d)
      merged.rate_d = pc.rate_d;
    if (k == 0 || pc.rate_n < merged.rate_n)
      merged.rate_n = pc.rate_n;
  }
  merged.format = flexible ? NNSB_FORMAT_FLEXIBLE : NNSB_FORMAT_STATIC;
  m->config = merged;
  m->configured = true;
  if (out)
    return copy_str (caps_to_string (tensors_caps_from_config (merged)), out, outlen);
  return NNSB_OK;
}

extern "C" int
nnsb_mux_get_config (nnsb_mux * m, nnsb_tensors_config * cfg)
{
  if (!m || !cfg || !m->configured)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_mux_get_config: not configured");
  *cfg = m->config;
  return NNSB_OK;
}

/* one collected set of buffers -> one buffer: a reference per memory, in pad order */
extern "C" int
nnsb_mux_chain (nnsb_mux * m, const nnsb_memory * in, uint32_t n_in, nnsb_memory * out, uint32_t * n_out)
{
  if (!m || !in || !out || !n_out)
    return NNSB_FLOW_ERROR;
  *n_out = 0;
  if (!m->configured) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_mux: caps are not set");
    return NNSB_FLOW_ERROR;
  }
  if (n_in != m->config.num_tensors && m->config.format == NNSB_FORMAT_STATIC) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_mux: %u memories collected, %u tensors configured",
        n_in, m->config.num_tensors);
    return NNSB_FLOW_ERROR;
  }
  if (n_in > NNSB_SIZE_LIMIT)
    return NNSB_FLOW_ERROR;
  for (uint32_t i = 0; i < n_in; i++) {
    if (nnsb_memory_ref (&in[i], &out[i]) != NNSB_OK) {
      for (uint32_t k = 0; k < i; k++)
        nnsb_memory_release (&out[k]);
      return NNSB_FLOW_ERROR;
    }
  }
  *n_out = n_in;
  return NNSB_FLOW_OK;
}

/* ------------------------------------------------------------------------------------- */
/* tensor_demux                                                                            */
/* ------------------------------------------------------------------------------------- */
struct nnsb_demux
{
  bool silent = true;
  bool has_pick = false;
  std::string pick_text;
  std::vector<std::vector<int64_t> > pick;      /* one index group per source pad */
  bool configured = false;
  nnsb_tensors_config config;
};

extern "C" int
nnsb_demux_new (nnsb_demux ** out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_demux_new: null argument");
  nnsb_demux *d = new (std::nothrow) nnsb_demux ();
  if (!d)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_demux_new: out of memory");
  config_init (&d->config);
  *out = d;
  return NNSB_OK;
}

extern "C" int
nnsb_demux_free (nnsb_demux * d)
{
  delete d;
  return NNSB_OK;
}

/* gst_tensor_demux_set_property, gsttensor_demux.c:590-626 */
extern "C" int
nnsb_demux_set_property (nnsb_demux * d, const char *name, const char *value)
{
  if (!d || !name || !value)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_demux_set_property: null argument");
  const std::string n (name);
  if (n == "silent") {
    bool b;
    if (!parse_bool (value, &b))
      return nnsb_set_error (NNSB_EINVAL, "silent: expected a boolean");
    d->silent = b;
    return NNSB_OK;
  }
  if (n == "tensorpick") {
    d->pick.clear ();
    d->pick_text = value;
    d->has_pick = true;
    for (const std::string & grp : split_set (value, ",.;/")) {
      std::vector<int64_t> idx;
      for (const std::string & tok : split_set (grp, ":+"))
        idx.push_back (strtoll (tok.c_str (), nullptr, 10));    /* g_ascii_strtoll */
      d->pick.push_back (idx);
    }
    return NNSB_OK;
  }
  return nnsb_set_error (NNSB_EINVAL, "tensor_demux has no property '%s'", name);
}

extern "C" int
nnsb_demux_get_property (nnsb_demux * d, const char *name, char *buf, size_t buflen)
{
  if (!d || !name)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_demux_get_property: null argument");
  const std::string n (name);
  if (n == "silent")
    return copy_str (d->silent ? "true" : "false", buf, buflen);
  if (n == "tensorpick") {
    /* :637-660: the groups joined with ',' */
    std::string s;
    for (size_t i = 0; i < d->pick.size (); i++) {
      if (i)
        s += ",";
      const std::vector<std::string> grp = split_set (d->pick_text, ",.;/");
      s += i < grp.size () ? grp[i] : "";
    }
    return copy_str (s, buf, buflen);
  }
  return nnsb_set_error (NNSB_EINVAL, "tensor_demux has no property '%s'", name);
}

/* sink caps (gst_tensor_demux_event CAPS, :246-272) */
extern "C" int
nnsb_demux_set_caps (nnsb_demux * d, const char *caps)
{
  if (!d || !caps)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_demux_set_caps: null argument");
  Caps c;
  std::string err;
  if (!caps_parse (caps, &c, &err) || c.structures.empty () || !c.is_fixed ())
    return nnsb_set_error (NNSB_EINVAL, "cannot parse fixed caps '%s'", caps);
  nnsb_tensors_config cfg;
  if (!config_from_structure (&cfg, c.structures[0]))
    return nnsb_set_error (NNSB_EINVAL, "tensor_demux: not a tensor stream");
  if (cfg.format == NNSB_FORMAT_STATIC && !config_validate (cfg))
    return nnsb_set_error (NNSB_EINVAL, "tensor_demux: invalid tensors info");
  d->config = cfg;
  d->configured = true;
  return NNSB_OK;
}

extern "C" int
nnsb_demux_num_outputs (nnsb_demux * d)
{
  if (!d || !d->configured)
    return 0;
  return d->has_pick ? (int) d->pick.size () : (int) d->config.num_tensors;
}

/* gst_tensor_demux_get_tensor_config, :288-330 */
static bool
demux_pad_config (const nnsb_demux * d, uint32_t nth, uint32_t total, nnsb_tensors_config * out)
{
  config_init (out);
  if (d->has_pick) {
    if (nth >= d->pick.size ())
      return false;
    const std::vector<int64_t> & grp = d->pick[nth];
    if (grp.size () > NNSB_SIZE_LIMIT)
      return false;
    for (size_t i = 0; i < grp.size (); i++) {
      if (grp[i] < 0 || (uint64_t) grp[i] >= total)
        return false;           /* "Internal error, handle invalid index." */
      out->info[i] = d->config.info[grp[i]];
    }
    out->num_tensors = (uint32_t) grp.size ();
  } else {
    if (nth >= total)
      return false;
    out->num_tensors = 1;
    out->info[0] = d->config.info[nth];
  }
  out->format = d->config.format;
  out->rate_n = d->config.rate_n;
  out->rate_d = d->config.rate_d;
  return true;
}

extern "C" int
nnsb_demux_get_out_caps (nnsb_demux * d, uint32_t idx, char *out, size_t outlen)
{
  if (!d || !d->configured)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_demux_get_out_caps: not configured");
  nnsb_tensors_config pc;
  if (!demux_pad_config (d, idx, d->config.num_tensors, &pc))
    return nnsb_set_error (NNSB_EINVAL, "nnsb_demux_get_out_caps: no such pad / invalid tensorpick index");
  return copy_str (caps_to_string (pad_caps (pc)), out, outlen);
}

/* gst_tensor_demux_chain, :468-560 */
extern "C" int
nnsb_demux_chain (nnsb_demux * d, const nnsb_memory * in, uint32_t n_in, nnsb_memory * outs,
    uint32_t max_out, uint32_t * counts, uint32_t max_src, uint32_t * n_src)
{
  if (!d || !in || !outs || !counts || !n_src)
    return NNSB_FLOW_ERROR;
  *n_src = 0;
  if (!d->configured) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_demux: caps are not set");
    return NNSB_FLOW_ERROR;
  }
  /* gst_tensor_buffer_from_config: one blob with several static tensors -> shares */
  std::vector<nnsb_memory> mems;
  bool cut = false;
  if (d->config.format == NNSB_FORMAT_STATIC && n_in == 1 && d->config.num_tensors > 1) {
    size_t off = 0;
    for (uint32_t i = 0; i < d->config.num_tensors; i++) {
      const size_t sz = nnsb_info_size (&d->config.info[i]);
      nnsb_memory m;
      if (nnsb_memory_share (&in[0], off, sz, &m) != NNSB_OK) {
        for (auto & x : mems)
          nnsb_memory_release (&x);
        return NNSB_FLOW_ERROR;
      }
      mems.push_back (m);
      off += sz;
    }
    cut = true;
  } else {
    mems.assign (in, in + n_in);
  }
  const uint32_t num_tensors = d->config.format == NNSB_FORMAT_FLEXIBLE ? (uint32_t) mems.size ()
      : d->config.num_tensors;
  int res = NNSB_FLOW_OK;
  if (mems.size () != num_tensors) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_demux: %zu memories for %u tensors", mems.size (), num_tensors);
    res = NNSB_FLOW_ERROR;
  }
  const uint32_t num_srcs = d->has_pick ? (uint32_t) d->pick.size () : num_tensors;
  if (res == NNSB_FLOW_OK && num_srcs > max_src) {
    nnsb_set_error (NNSB_ERANGE, "tensor_demux: %u source pads, room for %u", num_srcs, max_src);
    res = NNSB_FLOW_ERROR;
  }
  uint32_t total = 0;
  for (uint32_t i = 0; i < num_srcs && res == NNSB_FLOW_OK; i++) {
    std::vector<int64_t> one (1, (int64_t) i);
    const std::vector<int64_t> & grp = d->has_pick ? d->pick[i] : one;
    counts[i] = 0;
    for (int64_t idx : grp) {
      if (idx < 0 || (uint64_t) idx >= mems.size () || total >= max_out) {
        nnsb_set_error (NNSB_FLOW_ERROR, "tensor_demux: tensorpick index %lld is out of range (or no room)",
            (long long) idx);
        res = NNSB_FLOW_ERROR;
        break;
      }
      /* gst_buffer_get_memory: a new reference */
      if (nnsb_memory_ref (&mems[idx], &outs[total]) != NNSB_OK) {
        res = NNSB_FLOW_ERROR;
        break;
      }
      total++;
      counts[i]++;
    }
    if (res == NNSB_FLOW_OK)
      (*n_src)++;
  }
  if (cut)
    for (auto & x : mems)
      nnsb_memory_release (&x);
  if (res != NNSB_FLOW_OK) {
    for (uint32_t k = 0; k < total; k++)
      nnsb_memory_release (&outs[k]);
    *n_src = 0;
  }
  return res;
}

/* ------------------------------------------------------------------------------------------
 * nnsb_branch: converter -> transform ... -> transform pushed by one call (gst_pad_push chain)
 * ------------------------------------------------------------------------------------------ */
struct nnsb_branch
{
  nnsb_converter *conv = nullptr;
  std::vector<nnsb_transform *> t;
};

extern "C" int
nnsb_branch_new (nnsb_converter * conv, nnsb_transform * const *t, uint32_t n_t, nnsb_branch ** out)
{
  if (!out || (n_t && !t) || (!conv && !n_t))
    return nnsb_set_error (NNSB_EINVAL, "nnsb_branch_new: no elements");
  for (uint32_t i = 0; i < n_t; i++)
    if (!t[i])
      return nnsb_set_error (NNSB_EINVAL, "nnsb_branch_new: element %u is NULL", i);
  nnsb_branch *b = new (std::nothrow) nnsb_branch ();
  if (!b)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_branch_new: out of memory");
  b->conv = conv;
  b->t.assign (t, t + n_t);
  *out = b;
  return NNSB_OK;
}

extern "C" int
nnsb_branch_free (nnsb_branch * b)
{
  delete b;
  return NNSB_OK;
}

extern "C" int
nnsb_branch_push (nnsb_branch * b, const nnsb_memory * in, uint32_t n_in, nnsb_memory * outs,
    uint32_t *n_out)
{
  if (!b || !in || !outs || !n_out || n_in == 0 || n_in > NNSB_SIZE_LIMIT) {
    nnsb_set_error (NNSB_EINVAL, "nnsb_branch_push: bad arguments");
    return NNSB_FLOW_ERROR;
  }
  *n_out = 0;
  nnsb_memory a[NNSB_SIZE_LIMIT], c[NNSB_SIZE_LIMIT];
  nnsb_memory *cur = a, *next = c;
  const nnsb_memory *src = in;
  uint32_t n = n_in;
  bool own = false;             /* cur[] holds references of ours */
  if (b->conv) {
    int produced = 0;
    if (n_in != 1) {
      nnsb_set_error (NNSB_EINVAL, "nnsb_branch_push: the converter takes one frame per push");
      return NNSB_FLOW_ERROR;
    }
    if (nnsb_converter_chain (b->conv, in, cur, &produced) != NNSB_FLOW_OK)
      return NNSB_FLOW_ERROR;
    if (!produced)
      return NNSB_FLOW_OK;
    src = cur;
    n = 1;
    own = true;
  }
  for (size_t k = 0; k < b->t.size (); k++) {
    uint32_t m = 0;
    nnsb_memory *dst = (k + 1 == b->t.size ()) ? outs : next;
    const int r = nnsb_transform_transform (b->t[k], src, n, dst, &m);
    if (own)
      for (uint32_t i = 0; i < n; i++)
        nnsb_memory_release (const_cast<nnsb_memory *> (&src[i]));
    if (r != NNSB_FLOW_OK)
      return NNSB_FLOW_ERROR;
    if (dst == outs) {
      *n_out = m;
      return NNSB_FLOW_OK;
    }
    src = dst;
    n = m;
    own = true;
    std::swap (cur, next);
  }
  /* converter only */
  for (uint32_t i = 0; i < n; i++)
    outs[i] = src[i];
  *n_out = n;
  return NNSB_FLOW_OK;
}
