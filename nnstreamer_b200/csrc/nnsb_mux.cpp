/*
 * nnsb_mux.cpp -- data paths of tensor_merge and tensor_split on the device (SURVEY.md 8f
 * rank 4).
 *
 *   nnsb_merge  <->  GstTensorMerge, mode=linear (gst/nnstreamer/elements/gsttensor_merge.c):
 *                    option "0".."3" = the axis along which the input tensors are concatenated
 *                    (:91-97,779-800); output dims / type checks of get_merged_config
 *                    (:380-427); the nested memcpy loops of generate_mem (:506-586) are one
 *                    launch of the merge kernel (K10).
 *   nnsb_split  <->  GstTensorSplit (gsttensor_split.c): tensorseg / tensorpick parsing
 *                    (:577-621), one output memory per picked segment, each a copy of a
 *                    contiguous byte range (get_splited, :413-450) -- D2D / H2D copies.
 *
 * Not mirrored: the collect-pads time synchronisation of tensor_merge (sync-mode / sync-option,
 * tensor_common's gst_tensor_time_sync_*): `chain` takes one already collected set of buffers.
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <strings.h>
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

int
ensure_device (int device_id, nnsb_ctx ** ctx, nnsb_stream ** stream)
{
  if (*ctx)
    return NNSB_OK;
  int r = nnsb_ctx_create (device_id, ctx);
  if (r == NNSB_OK)
    r = nnsb_stream_create (*ctx, stream);
  return r;
}

}  // namespace

/* ------------------------------------------------------------------------------------- */
/* tensor_merge                                                                            */
/* ------------------------------------------------------------------------------------- */
struct nnsb_merge
{
  int device_id = 0;
  nnsb_ctx *ctx = nullptr;
  nnsb_stream *stream = nullptr;
  bool silent = true;
  bool mode_linear = false;             /* GTT_END until mode=linear is set */
  bool has_option = false;
  std::string option;
  int direction = -1;                   /* LINEAR_FIRST .. LINEAR_FOURTH */
  std::string sync_mode = "nosync", sync_option;
  bool configured = false;
  nnsb_tensors_config in_config, out_config;
};

extern "C" int
nnsb_merge_new (int device_id, nnsb_merge ** out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_merge_new: null argument");
  nnsb_merge *m = new (std::nothrow) nnsb_merge ();
  if (!m)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_merge_new: out of memory");
  m->device_id = device_id;
  config_init (&m->in_config);
  config_init (&m->out_config);
  *out = m;
  return NNSB_OK;
}

extern "C" int
nnsb_merge_free (nnsb_merge * m)
{
  if (!m)
    return NNSB_OK;
  if (m->stream) {
    nnsb_stream_sync (m->stream);
    nnsb_stream_destroy (m->stream);
  }
  if (m->ctx)
    nnsb_ctx_destroy (m->ctx);
  delete m;
  return NNSB_OK;
}

/* gst_tensor_merge_set_option_data, :778-803 */
static bool
merge_apply_option (nnsb_merge * m)
{
  if (!m->mode_linear || !m->has_option)
    return true;
  static const char *const k_linear[] = { "0", "1", "2", "3" };
  for (int i = 0; i < 4; i++)
    if (m->option == k_linear[i]) {
      m->direction = i;
      return true;
    }
  return false;
}

extern "C" int
nnsb_merge_set_property (nnsb_merge * m, const char *name, const char *value)
{
  if (!m || !name || !value)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_merge_set_property: null argument");
  const std::string n (name);
  if (n == "mode") {
    /* gst_tensor_merge_get_mode, :244-253: only "linear" exists, compared without case */
    if (strcasecmp (value, "linear") != 0)
      return nnsb_set_error (NNSB_EINVAL, "mode: '%s' is not a merge mode", value);
    m->mode_linear = true;
    if (!merge_apply_option (m))
      return nnsb_set_error (NNSB_EINVAL, "option: '%s' is not an axis of mode=linear", m->option.c_str ());
    return NNSB_OK;
  }
  if (n == "option") {
    m->has_option = true;
    m->option = value;
    if (!merge_apply_option (m)) {
      m->direction = -1;
      return nnsb_set_error (NNSB_EINVAL, "option: '%s' is not an axis of mode=linear", value);
    }
    return NNSB_OK;
  }
  if (n == "silent") {
    m->silent = strcmp (value, "false") != 0 && strcmp (value, "0") != 0;
    return NNSB_OK;
  }
  if (n == "sync-mode") {
    m->sync_mode = value;
    return NNSB_OK;
  }
  if (n == "sync-option") {
    m->sync_option = value;
    return NNSB_OK;
  }
  return nnsb_set_error (NNSB_EINVAL, "tensor_merge has no property '%s'", name);
}

extern "C" int
nnsb_merge_get_property (nnsb_merge * m, const char *name, char *buf, size_t buflen)
{
  if (!m || !name)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_merge_get_property: null argument");
  const std::string n (name);
  if (n == "mode")
    return copy_str (m->mode_linear ? "linear" : "", buf, buflen);
  if (n == "option")
    return m->has_option ? copy_str (m->option, buf, buflen) : NNSB_ENOENT;
  if (n == "silent")
    return copy_str (m->silent ? "true" : "false", buf, buflen);
  if (n == "sync-mode")
    return copy_str (m->sync_mode, buf, buflen);
  if (n == "sync-option")
    return copy_str (m->sync_option, buf, buflen);
  return nnsb_set_error (NNSB_EINVAL, "tensor_merge has no property '%s'", name);
}

/* sink caps of every pad in, merged source caps out: gst_tensor_merge_get_merged_config,
 * :380-427 (type and every dim but the merge axis must agree) */
extern "C" int
nnsb_merge_set_caps (nnsb_merge * m, uint32_t n, const char *const *caps, char *out, size_t outlen)
{
  if (!m || !caps || n < 1 || n > NNSB_SIZE_LIMIT)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_merge_set_caps: invalid argument");
  if (m->direction < 0)
    return nnsb_set_error (NNSB_EINVAL, "tensor_merge: mode / option are not set");
  nnsb_tensors_config in;
  config_init (&in);
  for (uint32_t i = 0; i < n; i++) {
    Caps c;
    std::string err;
    if (!caps[i] || !caps_parse (caps[i], &c, &err) || c.structures.empty () || !c.is_fixed ())
      return nnsb_set_error (NNSB_EINVAL, "sink_%u: cannot parse fixed caps", i);
    nnsb_tensors_config one;
    if (!config_from_structure (&one, c.structures[0]) || !config_validate (one) || one.num_tensors != 1)
      return nnsb_set_error (NNSB_EINVAL, "sink_%u: not a single-tensor caps", i);
    in.info[i] = one.info[0];
    if (i == 0) {
      in.rate_n = one.rate_n;
      in.rate_d = one.rate_d;
    }
  }
  in.num_tensors = n;
  nnsb_tensor_info merged = in.info[0];
  for (uint32_t i = 1; i < n; i++) {
    if (in.info[i].type != merged.type)
      return nnsb_set_error (NNSB_EINVAL, "tensor_merge: sink_%u has a different type (negotiation error)", i);
    for (int j = 0; j < NNSB_RANK_LIMIT; j++) {
      if (j == m->direction)
        merged.dim[j] += in.info[i].dim[j];
      else if (merged.dim[j] != in.info[i].dim[j])
        return nnsb_set_error (NNSB_EINVAL,
            "tensor_merge: sink_%u differs in dimension %d (negotiation error)", i, j);
    }
  }
  m->in_config = in;
  config_init (&m->out_config);
  m->out_config.num_tensors = 1;
  m->out_config.info[0] = merged;
  m->out_config.rate_n = in.rate_n;
  m->out_config.rate_d = in.rate_d;
  m->configured = true;
  if (out)
    return copy_str (caps_to_string (tensor_caps_from_config (m->out_config)), out, outlen);
  return NNSB_OK;
}

extern "C" int
nnsb_merge_get_config (nnsb_merge * m, nnsb_tensors_config * cfg)
{
  if (!m || !cfg || !m->configured)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_merge_get_config: not configured");
  *cfg = m->out_config;
  return NNSB_OK;
}

/* gst_tensor_merge_generate_mem, :463-600: one memory per sink pad in, the merged tensor out */
extern "C" int
nnsb_merge_chain (nnsb_merge * m, const nnsb_memory * in, uint32_t n, nnsb_memory * out)
{
  if (!m || !in || !out)
    return NNSB_FLOW_ERROR;
  if (!m->configured || n != m->in_config.num_tensors) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_merge: %u memories for %u negotiated pads", n,
        m->configured ? m->in_config.num_tensors : 0u);
    return NNSB_FLOW_ERROR;
  }
  if (ensure_device (m->device_id, &m->ctx, &m->stream) != NNSB_OK)
    return NNSB_FLOW_ERROR;
  const size_t es = nnsb_element_size (m->in_config.info[0].type);
  const uint32_t *dim0 = m->in_config.info[0].dim;
  /* rows = dims above the axis (taken from the first tensor, :512-514) */
  uint64_t rows = 1;
  for (int j = m->direction + 1; j < 4; j++)
    rows *= dim0[j];
  const void *d_in[NNSB_SIZE_LIMIT];
  uint64_t blk[NNSB_SIZE_LIMIT];
  std::vector<void *> staged;
  size_t out_size = 0;
  int r = NNSB_OK;
  for (uint32_t k = 0; k < n && r == NNSB_OK; k++) {
    uint64_t c = es;
    for (int j = 0; j <= m->direction; j++)
      c *= m->in_config.info[k].dim[j];
    blk[k] = c;
    if (in[k].size < c * rows || !in[k].data) {
      nnsb_set_error (NNSB_FLOW_ERROR, "tensor_merge: sink_%u holds %zu bytes, %llu expected", k,
          in[k].size, (unsigned long long) (c * rows));
      r = NNSB_FLOW_ERROR;
      break;
    }
    out_size += in[k].size;             /* outSize += mInfo[i].size, :494 */
    if (in[k].on_device) {
      d_in[k] = in[k].data;
      r = nnsb_memory_wait (&in[k], m->stream);   /* may still be in flight on its producer */
    } else {
      void *d = nullptr;
      r = nnsb_dev_alloc (m->ctx, in[k].size, &d);
      if (r == NNSB_OK) {
        staged.push_back (d);
        r = nnsb_h2d_async (m->stream, d, in[k].data, in[k].size);
        d_in[k] = d;
      }
    }
  }
  void *d_out = nullptr;
  if (r == NNSB_OK)
    r = nnsb_dev_alloc (m->ctx, out_size, &d_out);
  if (r == NNSB_OK)
    r = nnsb_merge_launch (m->stream, n, d_in, blk, d_out, rows);
  if (r == NNSB_OK)
    r = nnsb_stream_sync (m->stream);
  for (void *d : staged)
    nnsb_dev_free (m->ctx, d);
  if (r != NNSB_OK) {
    if (d_out)
      nnsb_dev_free (m->ctx, d_out);
    return NNSB_FLOW_ERROR;
  }
  if (nnsb_memory_adopt (m->ctx, d_out, out_size, 1, out) != NNSB_OK) {
    nnsb_dev_free (m->ctx, d_out);
    return NNSB_FLOW_ERROR;
  }
  return NNSB_FLOW_OK;
}

/* ------------------------------------------------------------------------------------- */
/* tensor_split                                                                            */
/* ------------------------------------------------------------------------------------- */
struct nnsb_split
{
  int device_id = 0;
  nnsb_ctx *ctx = nullptr;
  nnsb_stream *stream = nullptr;
  bool silent = true;
  bool has_seg = false;
  std::vector<std::vector<uint32_t>> seg;       /* tensorseg: one dim[8] per output */
  std::vector<int64_t> pick;                    /* tensorpick */
  bool configured = false;
  nnsb_tensors_config sink_config;
};

extern "C" int
nnsb_split_new (int device_id, nnsb_split ** out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_split_new: null argument");
  nnsb_split *s = new (std::nothrow) nnsb_split ();
  if (!s)
    return nnsb_set_error (NNSB_ENOMEM, "nnsb_split_new: out of memory");
  s->device_id = device_id;
  config_init (&s->sink_config);
  *out = s;
  return NNSB_OK;
}

extern "C" int
nnsb_split_free (nnsb_split * s)
{
  if (!s)
    return NNSB_OK;
  if (s->stream) {
    nnsb_stream_sync (s->stream);
    nnsb_stream_destroy (s->stream);
  }
  if (s->ctx)
    nnsb_ctx_destroy (s->ctx);
  delete s;
  return NNSB_OK;
}

/* gst_tensor_split_set_property, :566-626 */
extern "C" int
nnsb_split_set_property (nnsb_split * s, const char *name, const char *value)
{
  if (!s || !name || !value)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_split_set_property: null argument");
  const std::string n (name);
  if (n == "silent") {
    s->silent = strcmp (value, "false") != 0 && strcmp (value, "0") != 0;
    return NNSB_OK;
  }
  if (n == "tensorpick") {
    /* appended to the list, values through g_ascii_strtoll */
    for (const std::string & tok : split_set (value, ",.;/"))
      s->pick.push_back (strtoll (tok.c_str (), nullptr, 10));
    return NNSB_OK;
  }
  if (n == "tensorseg") {
    s->seg.clear ();
    for (const std::string & tok : split_set (value, ",.;/")) {
      std::vector<uint32_t> d (NNSB_RANK_LIMIT, 1);
      const std::vector<std::string> parts = split_set (tok, ":");
      for (size_t k = 0; k < parts.size () && k < (size_t) NNSB_RANK_LIMIT; k++)
        d[k] = (uint32_t) strtod (parts[k].c_str (), nullptr);        /* g_ascii_strtod, :611 */
      s->seg.push_back (d);
    }
    s->has_seg = true;
    return NNSB_OK;
  }
  return nnsb_set_error (NNSB_EINVAL, "tensor_split has no property '%s'", name);
}

extern "C" int
nnsb_split_get_property (nnsb_split * s, const char *name, char *buf, size_t buflen)
{
  if (!s || !name)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_split_get_property: null argument");
  const std::string n (name);
  if (n == "silent")
    return copy_str (s->silent ? "true" : "false", buf, buflen);
  if (n == "tensorpick") {
    /* :643-658: "a,b,c" */
    std::string v;
    for (size_t i = 0; i < s->pick.size (); i++)
      v += (i ? "," : "") + std::to_string (s->pick[i]);
    return copy_str (v, buf, buflen);
  }
  if (n == "tensorseg") {
    /* :660-690: dims printed with every rank, comma separated */
    if (!s->has_seg)
      return copy_str ("", buf, buflen);
    std::string v;
    for (size_t i = 0; i < s->seg.size (); i++) {
      char d[128];
      nnsb_dim_to_string (s->seg[i].data (), NNSB_RANK_LIMIT, d, sizeof (d));
      v += (i ? "," : "") + std::string (d);
    }
    return copy_str (v, buf, buflen);
  }
  return nnsb_set_error (NNSB_EINVAL, "tensor_split has no property '%s'", name);
}

extern "C" int
nnsb_split_set_caps (nnsb_split * s, const char *caps)
{
  if (!s || !caps)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_split_set_caps: null argument");
  Caps c;
  std::string err;
  if (!caps_parse (caps, &c, &err) || c.structures.empty () || !c.is_fixed ())
    return nnsb_set_error (NNSB_EINVAL, "cannot parse fixed caps '%s'", caps);
  nnsb_tensors_config cfg;
  if (!config_from_structure (&cfg, c.structures[0]) || !config_validate (cfg))
    return nnsb_set_error (NNSB_EINVAL, "Cannot configure tensor info");
  s->sink_config = cfg;
  s->configured = true;
  return NNSB_OK;
}

/* number of source pads a buffer produces (picked segments) */
extern "C" int
nnsb_split_num_outputs (nnsb_split * s)
{
  if (!s || !s->has_seg)
    return 0;
  if (s->pick.empty ())
    return (int) s->seg.size ();
  int n = 0;
  for (size_t i = 0; i < s->seg.size (); i++)
    for (int64_t p : s->pick)
      if (p == (int64_t) i) {
        n++;
        break;
      }
  return n;
}

/* caps of the idx-th source pad in creation order (gst_tensor_split_get_tensor_pad, :345-361) */
extern "C" int
nnsb_split_get_out_caps (nnsb_split * s, uint32_t idx, char *out, size_t outlen)
{
  if (!s || !s->configured || !s->has_seg)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_split_get_out_caps: not configured");
  /* the reference indexes tensorseg with the running pad count, not with the picked index */
  if (idx >= s->seg.size ())
    return nnsb_set_error (NNSB_EINVAL, "nnsb_split_get_out_caps: no such pad");
  nnsb_tensors_config pc;
  config_init (&pc);
  pc.num_tensors = 1;
  for (int i = 0; i < NNSB_RANK_LIMIT; i++)
    pc.info[0].dim[i] = s->seg[idx][i];
  pc.info[0].type = s->sink_config.info[0].type;
  pc.rate_n = s->sink_config.rate_n;
  pc.rate_d = s->sink_config.rate_d;
  return copy_str (caps_to_string (tensor_caps_from_config (pc)), out, outlen);
}

/* gst_tensor_split_chain, :455-530: every picked segment becomes an independent memory holding
 * a copy of its byte range (get_splited, :413-450) */
extern "C" int
nnsb_split_chain (nnsb_split * s, const nnsb_memory * in, nnsb_memory * outs, uint32_t max_out,
    uint32_t * n_out)
{
  if (!s || !in || !outs || !n_out)
    return NNSB_FLOW_ERROR;
  *n_out = 0;
  if (!s->has_seg) {
    nnsb_set_error (NNSB_FLOW_ERROR, "No rule to split incoming buffers.");
    return NNSB_FLOW_ERROR;
  }
  if (!s->configured || !in->data) {
    nnsb_set_error (NNSB_FLOW_ERROR, "tensor_split: caps are not set");
    return NNSB_FLOW_ERROR;
  }
  if ((uint32_t) nnsb_split_num_outputs (s) > max_out) {
    nnsb_set_error (NNSB_ERANGE, "tensor_split: %d outputs, room for %u", nnsb_split_num_outputs (s),
        max_out);
    return NNSB_FLOW_ERROR;
  }
  if (ensure_device (s->device_id, &s->ctx, &s->stream) != NNSB_OK)
    return NNSB_FLOW_ERROR;
  const size_t es = nnsb_element_size (s->sink_config.info[0].type);
  size_t offset = 0;
  int r = in->on_device ? nnsb_memory_wait (in, s->stream) : NNSB_OK;
  for (size_t i = 0; i < s->seg.size () && r == NNSB_OK; i++) {
    const size_t size = (size_t) nnsb_element_count (s->seg[i].data ()) * es;
    bool picked = s->pick.empty ();
    for (int64_t p : s->pick)
      picked = picked || p == (int64_t) i;
    if (picked) {
      if (offset + size > in->size) {
        nnsb_set_error (NNSB_FLOW_ERROR, "tensor_split: segment %zu ends at byte %zu of a %zu-byte buffer",
            i, offset + size, in->size);
        r = NNSB_FLOW_ERROR;
        break;
      }
      void *d = nullptr;
      r = nnsb_dev_alloc (s->ctx, size, &d);
      if (r == NNSB_OK) {
        const char *src = (const char *) in->data + offset;
        r = in->on_device ? nnsb_d2d_async (s->stream, d, src, size)
            : nnsb_h2d_async (s->stream, d, src, size);
        if (nnsb_memory_adopt (s->ctx, d, size, 1, &outs[*n_out]) == NNSB_OK) {
          (*n_out)++;
        } else {
          nnsb_stream_sync (s->stream);
          nnsb_dev_free (s->ctx, d);
          r = NNSB_FLOW_ERROR;
        }
      }
    }
    offset += size;
  }
  if (r == NNSB_OK)
    r = nnsb_stream_sync (s->stream);
  if (r != NNSB_OK) {
    for (uint32_t k = 0; k < *n_out; k++)
      nnsb_memory_release (&outs[k]);
    *n_out = 0;
    return NNSB_FLOW_ERROR;
  }
  return NNSB_FLOW_OK;
}
