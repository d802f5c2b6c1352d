/*
 * gsttensor_transform_b200.c -- the `tensor_transform` element on top of libnnstreamer_b200.so.
 *
 * Same element name, klass, pad templates, properties and vfuncs as the reference
 * (gst/nnstreamer/elements/gsttensor_transform.c:128-280); every vfunc forwards to the element
 * mirror of include/nnstreamer_b200.h, which holds the option grammar, the caps logic and the
 * CUDA kernels.  Caps cross the boundary as strings (gst_caps_to_string / _from_string).
 * Not compiled in this repository's image (no GStreamer): `make syntax` in this directory checks
 * it against declaration stubs; see README.md.
 */
#include <stdlib.h>
#include <string.h>
#include <gst/gst.h>
#include <gst/base/gstbasetransform.h>
#include <nnstreamer_b200.h>
#include "gstnnsbcudaallocator.h"

#define CAPS_STRING "other/tensor, framerate = (fraction) [ 0, max ]; " \
    "other/tensors, format = (string) { static, flexible }, framerate = (fraction) [ 0, max ]"
#define CAPS_BUF 8192

typedef struct _GstTensorTransformB200
{
  GstBaseTransform element;
  nnsb_transform *core;
  GstAllocator *cuda_alloc;     /* created at the first buffer, on the GPU `device-id` names */
  int device_id;
} GstTensorTransformB200;

typedef struct _GstTensorTransformB200Class
{
  GstBaseTransformClass parent_class;
} GstTensorTransformB200Class;

enum { PROP_0, PROP_SILENT, PROP_MODE, PROP_OPTION, PROP_ACCELERATION, PROP_APPLY,
  PROP_TRANSPOSE_RANK_LIMIT };

GST_DEBUG_CATEGORY_STATIC (gst_tensor_transform_debug);
#define GST_CAT_DEFAULT gst_tensor_transform_debug

static GstStaticPadTemplate sink_factory = GST_STATIC_PAD_TEMPLATE ("sink", GST_PAD_SINK,
    GST_PAD_ALWAYS, GST_STATIC_CAPS (CAPS_STRING));
static GstStaticPadTemplate src_factory = GST_STATIC_PAD_TEMPLATE ("src", GST_PAD_SRC,
    GST_PAD_ALWAYS, GST_STATIC_CAPS (CAPS_STRING));

G_DEFINE_TYPE (GstTensorTransformB200, gst_tensor_transform_b200, GST_TYPE_BASE_TRANSFORM);

/* gtt_mode_type, gsttensor_transform.c:174-210 */
static GType
mode_get_type (void)
{
  static GType t = 0;
  if (t == 0) {
    static GEnumValue v[] = {
      {NNSB_MODE_DIMCHG, "Mode for changing tensor dimensions", "dimchg"},
      {NNSB_MODE_TYPECAST, "Mode for casting type of tensor", "typecast"},
      {NNSB_MODE_ARITHMETIC, "Mode for arithmetic operations with tensor", "arithmetic"},
      {NNSB_MODE_TRANSPOSE, "Mode for transposing shape of tensor", "transpose"},
      {NNSB_MODE_STAND, "Mode for statistical standardization of tensor", "stand"},
      {NNSB_MODE_CLAMP, "Mode for clamping all elements of tensor into the range", "clamp"},
      {NNSB_MODE_UNKNOWN, "Unknown or not-implemented-yet mode", "unknown"},
      {0, NULL, NULL},
    };
    t = g_enum_register_static ("gtt_mode_type", v);
  }
  return t;
}

static void
set_property (GObject * object, guint prop_id, const GValue * value, GParamSpec * pspec)
{
  GstTensorTransformB200 *self = (GstTensorTransformB200 *) object;
  gchar buf[32];
  switch (prop_id) {
    case PROP_SILENT:
      nnsb_transform_set_property (self->core, "silent", g_value_get_boolean (value) ? "true" : "false");
      break;
    case PROP_MODE:
      g_snprintf (buf, sizeof (buf), "%d", g_value_get_enum (value));
      nnsb_transform_set_property (self->core, "mode", buf);
      break;
    case PROP_OPTION:
      /* an invalid option keeps the previous one and only logs (:985-990) */
      if (nnsb_transform_set_property (self->core, "option", g_value_get_string (value)) != NNSB_OK)
        g_critical ("%s", nnsb_last_error ());
      break;
    case PROP_ACCELERATION:
      nnsb_transform_set_property (self->core, "acceleration", g_value_get_boolean (value) ? "true" : "false");
      break;
    case PROP_APPLY:
      nnsb_transform_set_property (self->core, "apply", g_value_get_string (value));
      break;
    default:
      G_OBJECT_WARN_INVALID_PROPERTY_ID (object, prop_id, pspec);
      break;
  }
}

static void
get_property (GObject * object, guint prop_id, GValue * value, GParamSpec * pspec)
{
  GstTensorTransformB200 *self = (GstTensorTransformB200 *) object;
  gchar buf[CAPS_BUF];
  switch (prop_id) {
    case PROP_SILENT:
      nnsb_transform_get_property (self->core, "silent", buf, sizeof (buf));
      g_value_set_boolean (value, g_strcmp0 (buf, "true") == 0);
      break;
    case PROP_MODE:
      nnsb_transform_get_property (self->core, "mode", buf, sizeof (buf));
      g_value_set_enum (value, nnsb_mode_from_string (buf));
      break;
    case PROP_OPTION:
      if (nnsb_transform_get_property (self->core, "option", buf, sizeof (buf)) == NNSB_ENOENT)
        g_value_set_string (value, NULL);
      else
        g_value_set_string (value, buf);
      break;
    case PROP_ACCELERATION:
      nnsb_transform_get_property (self->core, "acceleration", buf, sizeof (buf));
      g_value_set_boolean (value, g_strcmp0 (buf, "true") == 0);
      break;
    case PROP_APPLY:
      nnsb_transform_get_property (self->core, "apply", buf, sizeof (buf));
      g_value_set_string (value, buf);
      break;
    case PROP_TRANSPOSE_RANK_LIMIT:
      g_value_set_uint (value, NNSB_TRANSPOSE_RANK_LIMIT);
      break;
    default:
      G_OBJECT_WARN_INVALID_PROPERTY_ID (object, prop_id, pspec);
      break;
  }
}

static GstCaps *
transform_caps (GstBaseTransform * trans, GstPadDirection direction, GstCaps * caps, GstCaps * filter)
{
  GstTensorTransformB200 *self = (GstTensorTransformB200 *) trans;
  gchar out[CAPS_BUF];
  gchar *c = gst_caps_to_string (caps);
  gchar *f = (filter && gst_caps_get_size (filter) > 0) ? gst_caps_to_string (filter) : NULL;
  GstCaps *result = NULL;
  if (nnsb_transform_transform_caps (self->core, direction == GST_PAD_SINK ? NNSB_PAD_SINK : NNSB_PAD_SRC,
          c, f, out, sizeof (out)) >= 0)
    result = g_strcmp0 (out, "EMPTY") == 0 ? gst_caps_new_empty () : gst_caps_from_string (out);
  g_free (c);
  g_free (f);
  return result ? result : gst_caps_new_empty ();
}

static GstCaps *
fixate_caps (GstBaseTransform * trans, GstPadDirection direction, GstCaps * caps, GstCaps * othercaps)
{
  GstCaps *result = transform_caps (trans, direction, caps, othercaps);
  gst_caps_unref (othercaps);
  result = gst_caps_make_writable (result);
  return gst_caps_fixate (result);
}

static gboolean
set_caps (GstBaseTransform * trans, GstCaps * incaps, GstCaps * outcaps)
{
  GstTensorTransformB200 *self = (GstTensorTransformB200 *) trans;
  gchar *i = gst_caps_to_string (incaps), *o = gst_caps_to_string (outcaps);
  const gboolean ok = nnsb_transform_set_caps (self->core, i, o) == NNSB_OK;
  if (!ok)
    GST_ERROR_OBJECT (self, "Set Caps Failed! %s", nnsb_last_error ());
  g_free (i);
  g_free (o);
  return ok;
}

static gboolean
transform_size (GstBaseTransform * trans, GstPadDirection direction, GstCaps * caps, gsize size,
    GstCaps * othercaps, gsize * othersize)
{
  (void) trans; (void) direction; (void) caps; (void) size; (void) othercaps;
  *othersize = 0;               /* memories are appended in transform(), :2169-2186 */
  return TRUE;
}

static void
release_host_memory (gpointer data)
{
  nnsb_memory *m = (nnsb_memory *) data;
  nnsb_memory_release (m);
  g_free (m);
}

static GstFlowReturn
transform (GstBaseTransform * trans, GstBuffer * inbuf, GstBuffer * outbuf)
{
  GstTensorTransformB200 *self = (GstTensorTransformB200 *) trans;
  const guint n = gst_buffer_n_memory (inbuf);
  nnsb_memory in[NNSB_SIZE_LIMIT], out[NNSB_SIZE_LIMIT];
  GstMapInfo maps[NNSB_SIZE_LIMIT];
  gboolean mapped[NNSB_SIZE_LIMIT] = { FALSE, };
  uint32_t n_out = 0;
  guint i;
  int flow;

  if (n > NNSB_SIZE_LIMIT)
    return GST_FLOW_ERROR;
  if (!self->cuda_alloc) {
    self->cuda_alloc = gst_nnsb_cuda_allocator_new (self->device_id);
    if (!self->cuda_alloc)
      return GST_FLOW_ERROR;
  }
  for (i = 0; i < n; i++) {
    GstMemory *mem = gst_buffer_peek_memory (inbuf, i);
    if (gst_nnsb_cuda_memory_get (mem, &in[i]))
      continue;                 /* device-resident input from a CUDA-aware peer */
    if (!gst_memory_map (mem, &maps[i], GST_MAP_READ)) {
      flow = NNSB_FLOW_ERROR;
      goto done;
    }
    mapped[i] = TRUE;
    memset (&in[i], 0, sizeof (in[i]));
    in[i].data = maps[i].data;
    in[i].size = maps[i].size;
  }
  flow = nnsb_transform_transform (self->core, in, n, out, &n_out);
  if (flow == NNSB_FLOW_OK) {
    for (i = 0; i < n_out; i++) {
      GstMemory *mem;
      if (out[i].on_device) {
        /* a transformed tensor, or -- for a tensor `apply` skipped -- another reference to the
         * input's device block (:1716); either way the new GstMemory takes the reference over */
        mem = gst_nnsb_cuda_memory_wrap (self->cuda_alloc, &out[i]);
      } else if (!out[i].owned) {
        /* host tensor not selected by `apply`: a view into the mapped input.  With one memory
         * per tensor that is a ref of memory i; when one input blob carried several tensors it
         * is a share of the blob at the tensor's offset (nnstreamer_plugin_api_impl.c:534) */
        if (n == 1 && n_out > 1) {
          const gssize off = (const char *) out[i].data - (const char *) in[0].data;
          mem = gst_memory_share (gst_buffer_peek_memory (inbuf, 0), off, (gssize) out[i].size);
        } else {
          mem = gst_memory_ref (gst_buffer_peek_memory (inbuf, i));
        }
      } else {
        nnsb_memory *keep = g_new (nnsb_memory, 1);
        *keep = out[i];
        mem = gst_memory_new_wrapped (0, out[i].data, out[i].size, 0, out[i].size, keep,
            release_host_memory);
      }
      gst_buffer_append_memory (outbuf, mem);
    }
  } else {
    GST_ERROR_OBJECT (self, "%s", nnsb_last_error ());
  }
done:
  for (i = 0; i < n; i++)
    if (mapped[i])
      gst_memory_unmap (gst_buffer_peek_memory (inbuf, i), &maps[i]);
  return flow == NNSB_FLOW_OK ? GST_FLOW_OK
      : (flow == NNSB_FLOW_NOT_SUPPORTED ? GST_FLOW_NOT_SUPPORTED : GST_FLOW_ERROR);
}

static void
finalize (GObject * object)
{
  GstTensorTransformB200 *self = (GstTensorTransformB200 *) object;
  if (self->cuda_alloc)
    gst_object_unref (self->cuda_alloc);
  nnsb_transform_free (self->core);
  G_OBJECT_CLASS (gst_tensor_transform_b200_parent_class)->finalize (object);
}

static void
gst_tensor_transform_b200_class_init (GstTensorTransformB200Class * klass)
{
  GObjectClass *gobject_class = (GObjectClass *) klass;
  GstElementClass *gstelement_class = (GstElementClass *) klass;
  GstBaseTransformClass *trans_class = (GstBaseTransformClass *) klass;

  GST_DEBUG_CATEGORY_INIT (gst_tensor_transform_debug, "tensor_transform", 0,
      "Element to transforms tensor dimension or type");
  gobject_class->set_property = set_property;
  gobject_class->get_property = get_property;
  gobject_class->finalize = finalize;
  g_object_class_install_property (gobject_class, PROP_SILENT,
      g_param_spec_boolean ("silent", "Silent", "Produce verbose output ?", FALSE, G_PARAM_READWRITE));
  g_object_class_install_property (gobject_class, PROP_MODE,
      g_param_spec_enum ("mode", "Mode", "Mode used for transforming tensor", mode_get_type (),
          NNSB_MODE_UNKNOWN, G_PARAM_READWRITE));
  g_object_class_install_property (gobject_class, PROP_OPTION,
      g_param_spec_string ("option", "Option", "Option for the tensor transform mode ?", "", G_PARAM_READWRITE));
  g_object_class_install_property (gobject_class, PROP_ACCELERATION,
      g_param_spec_boolean ("acceleration", "Acceleration", "ORC semantics (saturating integer ops) ?",
          FALSE, G_PARAM_READWRITE));
  g_object_class_install_property (gobject_class, PROP_APPLY,
      g_param_spec_string ("apply", "Apply", "Select tensors to apply, separated with ',' in case of multiple tensors. Default to apply all tensors.",
          "", G_PARAM_READWRITE));
  g_object_class_install_property (gobject_class, PROP_TRANSPOSE_RANK_LIMIT,
      g_param_spec_uint ("transpose-rank-limit", "Transpose rank limit", "The rank limit of transpose, which varies per version of nnstreamer and may be lower than the global rank limit if it is over 4.",
          0, NNSB_RANK_LIMIT, NNSB_TRANSPOSE_RANK_LIMIT, G_PARAM_READABLE));
  gst_element_class_set_details_simple (gstelement_class, "TensorTransform", "Filter/Tensor",
      "Transforms other/tensor dimensions for different models or frameworks (B200)",
      "nnstreamer_b200");
  gst_element_class_add_pad_template (gstelement_class, gst_static_pad_template_get (&src_factory));
  gst_element_class_add_pad_template (gstelement_class, gst_static_pad_template_get (&sink_factory));
  trans_class->passthrough_on_same_caps = FALSE;        /* :265 */
  trans_class->transform = GST_DEBUG_FUNCPTR (transform);
  trans_class->transform_caps = GST_DEBUG_FUNCPTR (transform_caps);
  trans_class->fixate_caps = GST_DEBUG_FUNCPTR (fixate_caps);
  trans_class->set_caps = GST_DEBUG_FUNCPTR (set_caps);
  trans_class->transform_size = GST_DEBUG_FUNCPTR (transform_size);
}

static void
gst_tensor_transform_b200_init (GstTensorTransformB200 * self)
{
  const char *dev = getenv ("NNSB_DEVICE_ID");  /* one pipeline branch per GPU */
  self->device_id = dev ? atoi (dev) : 0;
  self->cuda_alloc = NULL;
  nnsb_transform_new (self->device_id, &self->core);
}

GType gst_tensor_converter_b200_get_type (void);  /* gsttensor_converter_b200.c */

static gboolean
plugin_init (GstPlugin * plugin)
{
  /* same element names and rank as registerer/nnstreamer.c:80-86,109 */
  return gst_element_register (plugin, "tensor_transform", GST_RANK_NONE,
      gst_tensor_transform_b200_get_type ())
      && gst_element_register (plugin, "tensor_converter", GST_RANK_NONE,
      gst_tensor_converter_b200_get_type ());
}

#ifndef PACKAGE
#define PACKAGE "nnstreamer_b200"
#endif
GST_PLUGIN_DEFINE (GST_VERSION_MAJOR, GST_VERSION_MINOR, nnstreamer_b200,
    "B200-native tensor_transform", plugin_init, "1.0", "LGPL", "nnstreamer_b200", "https://example.invalid")
