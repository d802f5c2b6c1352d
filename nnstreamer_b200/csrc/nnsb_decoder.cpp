/*
 * nnsb_decoder.cpp -- tensor_decoder mode=direct_video on the device
 * (ext/nnstreamer/tensor_decoder/tensordec-directvideo.c): the GstTensorDecoderDef callbacks of
 * the sub-plugin, with dv_decode's row loop (:326-339) done by the pad kernel (device output)
 * or by a pitched D2H copy (host output).  Exact inverse of the converter's stride strip.
 */
#include <cstdio>
#include <cstring>
#include <new>
#include <string>

#include "nnsb_internal.h"

namespace {

/* dv_formats, :80-94 (index = direct_video_formats) */
const char *const k_formats[] = { "UNKNOWN", "GRAY8", "RGB", "BGR", "RGBx", "BGRx", "xRGB", "xBGR",
  "RGBA", "BGRA", "ARGB", "ABGR"
};
constexpr int N_FORMATS = (int) (sizeof (k_formats) / sizeof (k_formats[0]));

/* how many channels a format carries */
int
format_channels (int f)
{
  if (f == 1)
    return 1;
  if (f == 2 || f == 3)
    return 3;
  return f >= 4 ? 4 : 0;
}

/* find_key_strv: case-insensitive match (nnstreamer_plugin_api_util_impl.c:1195-1211) */
int
find_format (const char *param)
{
  for (int i = 0; i < N_FORMATS; i++) {
    const char *a = k_formats[i], *b = param;
    while (*a && *b && ((*a | 0x20) == (*b | 0x20) || *a == *b))
      a++, b++;
    if (!*a && !*b)
      return i;
  }
  return -1;
}

size_t
xraw_bufsize (const uint32_t *dim)
{
  /* dim[0] is bpp and there is zeropadding only when dim[0]%4 > 0, :265-270 */
  return (size_t) (((size_t) dim[0] * dim[1] - 1) / 4 + 1) * 4 * dim[2];
}

}  // namespace

struct nnsb_direct_video
{
  int device_id = 0;
  nnsb_ctx *ctx = nullptr;
  nnsb_stream *stream = nullptr;
  int format = 0;               /* DIRECT_VIDEO_FORMAT_UNKNOWN */
};

extern "C" int
nnsb_direct_video_new (int device_id, nnsb_direct_video ** out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_direct_video_new: null argument");
  nnsb_direct_video *dv = new (std::nothrow) nnsb_direct_video ();
  if (!dv)
    return nnsb_set_error (NNSB_ENOMEM, "Failed to allocate memory for decoder subplugin.");
  dv->device_id = device_id;
  *out = dv;
  return NNSB_OK;
}

extern "C" int
nnsb_direct_video_free (nnsb_direct_video * dv)
{
  if (!dv)
    return NNSB_OK;
  if (dv->stream) {
    nnsb_stream_sync (dv->stream);
    nnsb_stream_destroy (dv->stream);
  }
  if (dv->ctx)
    nnsb_ctx_destroy (dv->ctx);
  delete dv;
  return NNSB_OK;
}

extern "C" int
nnsb_direct_video_set_option (nnsb_direct_video * dv, int op_num, const char *param)
{
  if (!dv) {
    nnsb_set_error (NNSB_EINVAL, "There is no plugin data.");
    return 0;
  }
  if (!param || *param == '\0') {
    nnsb_set_error (NNSB_EINVAL, "Please set the valid value at option.");
    return 0;
  }
  if (op_num == 0) {
    const int f = find_format (param);
    dv->format = f < 0 ? 0 : f;
    if (dv->format == 0)
      return 0;
  }
  return 1;
}

extern "C" int
nnsb_direct_video_get_out_caps (nnsb_direct_video * dv, const nnsb_tensors_config * config,
    char *caps, size_t capslen)
{
  if (!dv || !config || !caps)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_direct_video_get_out_caps: null argument");
  if (config->num_tensors < 1)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_direct_video_get_out_caps: no tensor");
  /* Direct video uses the first tensor only even if it's multi-tensor */
  const uint32_t channel = config->info[0].dim[0];
  int format;
  if (channel == 1 || channel == 3 || channel == 4) {
    if (dv->format == 0)
      format = channel == 1 ? 1 /* GRAY8 */ : channel == 3 ? 2 /* RGB */ : 5 /* BGRx */;
    else if (format_channels (dv->format) == (int) channel)
      format = dv->format;
    else
      return nnsb_set_error (NNSB_EINVAL, "Invalid format. Please check the video format");
  } else {
    return nnsb_set_error (NNSB_EINVAL, "%u channel is not supported", channel);
  }
  /* DECODER_DV_VIDEO_CAPS_STR (:41-43) with format / width / height / framerate fixed */
  std::string s = "video/x-raw, format=(string)";
  s += k_formats[format];
  const uint32_t width = config->info[0].dim[1], height = config->info[0].dim[2];
  s += width > 0 ? ", width=(int)" + std::to_string (width) : ", width=(int)[ 1, 2147483647 ]";
  s += height > 0 ? ", height=(int)" + std::to_string (height) : ", height=(int)[ 1, 2147483647 ]";
  /* setFramerateFromConfig, tensordecutil.c:132-143 */
  if (config->rate_n >= 0 && config->rate_d > 0)
    s += ", framerate=(fraction)" + std::to_string (config->rate_n) + "/" + std::to_string (config->rate_d);
  else
    s += ", framerate=(fraction)[ 0/1, 2147483647/1 ]";
  s += ", views=(int)1, interlace-mode=(string)progressive";
  if (s.size () + 1 > capslen)
    return nnsb_set_error (NNSB_EINVAL, "output buffer too small (%zu needed)", s.size () + 1);
  memcpy (caps, s.c_str (), s.size () + 1);
  return (int) s.size ();
}

extern "C" size_t
nnsb_direct_video_transform_size (const nnsb_tensors_config * config)
{
  if (!config || config->num_tensors < 1)
    return 0;
  return xraw_bufsize (config->info[0].dim);
}

extern "C" int
nnsb_direct_video_decode (nnsb_direct_video * dv, const nnsb_tensors_config * config,
    const nnsb_memory * in, nnsb_memory * out)
{
  if (!dv || !config || !in || !out || !in->data)
    return NNSB_FLOW_ERROR;
  if (config->num_tensors < 1 || config->info[0].type != NNSB_UINT8) {
    nnsb_set_error (NNSB_FLOW_ERROR, "direct_video: the first tensor must be uint8");     /* :304 */
    return NNSB_FLOW_ERROR;
  }
  if (!dv->ctx) {
    if (nnsb_ctx_create (dv->device_id, &dv->ctx) != NNSB_OK
        || nnsb_stream_create (dv->ctx, &dv->stream) != NNSB_OK)
      return NNSB_FLOW_ERROR;
  }
  const uint32_t *dim = config->info[0].dim;
  const size_t row = (size_t) dim[0] * dim[1], rows = dim[2];
  const size_t pitch = ((row - 1) / 4 + 1) * 4;
  const size_t size = xraw_bufsize (dim);
  if (in->size < row * rows) {
    nnsb_set_error (NNSB_FLOW_ERROR, "direct_video: input of %zu bytes, %zu expected", in->size,
        row * rows);
    return NNSB_FLOW_ERROR;
  }
  /* "Don't reallocate. Reuse what's already given", :306-311 */
  const bool reuse = out->data != nullptr && out->size == size;
  void *d_out = nullptr;
  if (!reuse || out->on_device) {
    if (reuse)
      d_out = out->data;
    else if (nnsb_dev_alloc (dv->ctx, size, &d_out) != NNSB_OK)
      return NNSB_FLOW_ERROR;
  }
  int r = NNSB_OK;
  void *staged = nullptr;
  const void *d_in = in->data;
  if (in->on_device)
    r = nnsb_memory_wait (in, dv->stream);      /* may still be in flight on its producer */
  if (r == NNSB_OK && reuse && out->on_device)
    r = nnsb_memory_wait (out, dv->stream);
  if (r == NNSB_OK && !in->on_device) {
    /* host tensor: stage it; the padding is then done on the device or by the copy engine */
    r = nnsb_dev_alloc (dv->ctx, row * rows, &staged);
    if (r == NNSB_OK)
      r = nnsb_h2d_async (dv->stream, staged, in->data, row * rows);
    d_in = staged;
  }
  if (r == NNSB_OK) {
    if (d_out)
      /* No Padding Required (:322-324) degenerates to a flat copy inside the launcher */
      r = nnsb_pad_launch (dv->stream, d_in, d_out, row, pitch, rows);
    else
      r = nnsb_d2h_2d_async (dv->stream, out->data, pitch, d_in, row, row, rows);
  }
  if (r == NNSB_OK)
    r = nnsb_stream_sync (dv->stream);
  if (staged)
    nnsb_dev_free (dv->ctx, staged);
  if (r != NNSB_OK) {
    if (d_out && !reuse)
      nnsb_dev_free (dv->ctx, d_out);
    return NNSB_FLOW_ERROR;
  }
  if (!reuse && nnsb_memory_adopt (dv->ctx, d_out, size, 1, out) != NNSB_OK) {
    nnsb_dev_free (dv->ctx, d_out);
    return NNSB_FLOW_ERROR;
  }
  return NNSB_FLOW_OK;
}
