/*
 * gsttensor_converter_b200.c -- the `tensor_converter` element (video/x-raw branch) on top of
 * libnnstreamer_b200.so.
 *
 * Same element name, pads and chain / sink_event surface as the reference
 * (gst/nnstreamer/elements/gsttensor_converter.c:153-170 the vmethods, :331-353 the pads,
 * :1046-1109 the _NNS_VIDEO chain case); the work -- caps -> tensors config (parse_video,
 * :1421-1519), the row-stride strip (:1087-1100) and the frames-per-tensor batching
 * (:945-1009) -- is done by the element mirror nnsb_converter_* on the device.  Output buffers
 * carry ONE "nnsb.cuda" GstMemory (gstnnsbcudaallocator.c): a CUDA-aware tensor_transform
 * takes its device view, any other peer maps it (one D2H copy).
 * Only video/x-raw is handled here; the reference's other media types (audio, text, octet,
 * flex, sub-plugins) stay with the reference element.
 *
 * Not compiled in this repository's image (no GStreamer): `make syntax` in this directory checks
 * it against declaration stubs; see README.md.
 */
#include <stdlib.h>
#include <string.h>
#include <gst/gst.h>
#include <nnstreamer_b200.h>
#include "gstnnsbcudaallocator.h"

#define VIDEO_CAPS "video/x-raw, format = (string) { RGB, BGR, RGBx, BGRx, xRGB, xBGR, RGBA, BGRA, ARGB, ABGR, GRAY8 }, " \
    "width = (int) [ 1, max ], height = (int) [ 1, max ], framerate = (fraction) [ 0, max ]"
#define TENSOR_CAPS "other/tensor, framerate = (fraction) [ 0, max ]; " \
    "other/tensors, format = (string) static, framerate = (fraction) [ 0, max ]"
#define CAPS_BUF 8192

typedef struct _GstTensorConverterB200
{
  GstElement element;
  GstPad *sinkpad, *srcpad;
  nnsb_converter *core;
  GstAllocator *cuda_alloc;
  int device_id;
} GstTensorConverterB200;

typedef struct _GstTensorConverterB200Class
{
  GstElementClass parent_class;
} GstTensorConverterB200Class;

enum { PROP_0, PROP_FRAMES_PER_TENSOR, PROP_SILENT };

GST_DEBUG_CATEGORY_STATIC (gst_tensor_converter_debug);
#define GST_CAT_DEFAULT gst_tensor_converter_debug

static GstStaticPadTemplate sink_factory = GST_STATIC_PAD_TEMPLATE ("sink", GST_PAD_SINK,
    GST_PAD_ALWAYS, GST_STATIC_CAPS (VIDEO_CAPS));
static GstStaticPadTemplate src_factory = GST_STATIC_PAD_TEMPLATE ("src", GST_PAD_SRC,
    GST_PAD_ALWAYS, GST_STATIC_CAPS (TENSOR_CAPS));

G_DEFINE_TYPE (GstTensorConverterB200, gst_tensor_converter_b200, GST_TYPE_ELEMENT);

static void
set_property (GObject * object, guint prop_id, const GValue * value, GParamSpec * pspec)
{
  GstTensorConverterB200 *self = (GstTensorConverterB200 *) object;
  gchar buf[32];
  switch (prop_id) {
    case PROP_FRAMES_PER_TENSOR:   /* gsttensor_converter.c:453-456 */
      g_snprintf (buf, sizeof (buf), "%u", g_value_get_uint (value));
      if (nnsb_converter_set_property (self->core, "frames-per-tensor", buf) != NNSB_OK)
        g_critical ("%s", nnsb_last_error ());
      break;
    case PROP_SILENT:
      nnsb_converter_set_property (self->core, "silent", g_value_get_boolean (value) ? "true" : "false");
      break;
    default:
      G_OBJECT_WARN_INVALID_PROPERTY_ID (object, prop_id, pspec);
      break;
  }
}

static void
get_property (GObject * object, guint prop_id, GValue * value, GParamSpec * pspec)
{
  GstTensorConverterB200 *self = (GstTensorConverterB200 *) object;
  gchar buf[64];
  switch (prop_id) {
    case PROP_FRAMES_PER_TENSOR:
      nnsb_converter_get_property (self->core, "frames-per-tensor", buf, sizeof (buf));
      g_value_set_uint (value, (guint) strtoul (buf, NULL, 10));
      break;
    case PROP_SILENT:
      nnsb_converter_get_property (self->core, "silent", buf, sizeof (buf));
      g_value_set_boolean (value, g_strcmp0 (buf, "true") == 0);
      break;
    default:
      G_OBJECT_WARN_INVALID_PROPERTY_ID (object, prop_id, pspec);
      break;
  }
}

/* gst_tensor_converter_sink_event, CAPS case (:600-630): parse the video caps, announce the
 * tensor caps downstream */
static gboolean
sink_event (GstPad * pad, GstObject * parent, GstEvent * event)
{
  GstTensorConverterB200 *self = (GstTensorConverterB200 *) parent;
  if (GST_EVENT_TYPE (event) == GST_EVENT_CAPS) {
    GstCaps *caps = NULL;
    gchar out[CAPS_BUF];
    gchar *text;
    gboolean ok;
    gst_event_parse_caps (event, &caps);
    text = gst_caps_to_string (caps);
    ok = nnsb_converter_set_caps (self->core, text, out, sizeof (out)) >= 0;
    g_free (text);
    gst_event_unref (event);
    if (!ok) {
      GST_ERROR_OBJECT (self, "%s", nnsb_last_error ());
      return FALSE;
    }
    caps = gst_caps_from_string (out);
    ok = gst_pad_push_event (self->srcpad, gst_event_new_caps (caps));
    gst_caps_unref (caps);
    return ok;
  }
  return gst_pad_event_default (pad, parent, event);
}

/* gst_tensor_converter_chain, _NNS_VIDEO (:1046-1109) + chain_chunk (:945-1009) */
static GstFlowReturn
chain (GstPad * pad, GstObject * parent, GstBuffer * buf)
{
  GstTensorConverterB200 *self = (GstTensorConverterB200 *) parent;
  GstMemory *mem;
  GstMapInfo map;
  nnsb_memory frame, tensor;
  gboolean mapped = FALSE;
  int produced = 0, flow;
  GstFlowReturn ret = GST_FLOW_OK;
  (void) pad;

  if (!self->cuda_alloc) {
    self->cuda_alloc = gst_nnsb_cuda_allocator_new (self->device_id);
    if (!self->cuda_alloc) {
      gst_buffer_unref (buf);
      return GST_FLOW_ERROR;
    }
  }
  if (gst_buffer_n_memory (buf) != 1) {
    /* "supposed 1 frame in buffer" (:1060); decoders hand out one memory per frame */
    GST_ERROR_OBJECT (self, "expected one memory per video frame, got %u", gst_buffer_n_memory (buf));
    gst_buffer_unref (buf);
    return GST_FLOW_ERROR;
  }
  mem = gst_buffer_peek_memory (buf, 0);
  if (!gst_nnsb_cuda_memory_get (mem, &frame)) {
    if (!gst_memory_map (mem, &map, GST_MAP_READ)) {
      gst_buffer_unref (buf);
      return GST_FLOW_ERROR;
    }
    mapped = TRUE;
    memset (&frame, 0, sizeof (frame));
    frame.data = map.data;
    frame.size = map.size;
  }
  /* a host frame has been copied (H2D, row padding stripped by the copy engine) when this
   * returns, so the buffer can go back to its pool */
  flow = nnsb_converter_chain (self->core, &frame, &tensor, &produced);
  if (mapped)
    gst_memory_unmap (mem, &map);
  if (flow != NNSB_FLOW_OK) {
    GST_ERROR_OBJECT (self, "%s", nnsb_last_error ());
    ret = GST_FLOW_ERROR;
  } else if (produced) {
    GstBuffer *outbuf = gst_buffer_new ();
    gst_buffer_append_memory (outbuf, gst_nnsb_cuda_memory_wrap (self->cuda_alloc, &tensor));
    /* timestamps of the frame that completed the tensor; the reference interpolates per
     * frames-per-tensor from the adapter (:976-1001) */
    gst_buffer_copy_into (outbuf, buf, GST_BUFFER_COPY_METADATA, 0, (gsize) -1);
    ret = gst_pad_push (self->srcpad, outbuf);
  }
  gst_buffer_unref (buf);
  return ret;
}

static void
finalize (GObject * object)
{
  GstTensorConverterB200 *self = (GstTensorConverterB200 *) object;
  nnsb_converter_free (self->core);
  if (self->cuda_alloc)
    gst_object_unref (self->cuda_alloc);
  G_OBJECT_CLASS (gst_tensor_converter_b200_parent_class)->finalize (object);
}

static void
gst_tensor_converter_b200_class_init (GstTensorConverterB200Class * klass)
{
  GObjectClass *gobject_class = (GObjectClass *) klass;
  GstElementClass *gstelement_class = (GstElementClass *) klass;

  GST_DEBUG_CATEGORY_INIT (gst_tensor_converter_debug, "tensor_converter", 0,
      "Element to convert media stream to tensor stream");
  gobject_class->set_property = set_property;
  gobject_class->get_property = get_property;
  gobject_class->finalize = finalize;
  /* :231-241, :262-270 */
  g_object_class_install_property (gobject_class, PROP_FRAMES_PER_TENSOR,
      g_param_spec_uint ("frames-per-tensor", "Frames per tensor", "The number of frames in output tensor",
          1, 0xffffffffu, 1, G_PARAM_READWRITE));
  g_object_class_install_property (gobject_class, PROP_SILENT,
      g_param_spec_boolean ("silent", "Silent", "Produce verbose output", TRUE, G_PARAM_READWRITE));
  gst_element_class_set_details_simple (gstelement_class, "TensorConverter", "Converter/Tensor",
      "Converts a video stream to a tensor stream of C-Array in B200 HBM", "nnstreamer_b200");
  gst_element_class_add_pad_template (gstelement_class, gst_static_pad_template_get (&src_factory));
  gst_element_class_add_pad_template (gstelement_class, gst_static_pad_template_get (&sink_factory));
}

static void
gst_tensor_converter_b200_init (GstTensorConverterB200 * self)
{
  const char *dev = getenv ("NNSB_DEVICE_ID");
  self->device_id = dev ? atoi (dev) : 0;
  self->cuda_alloc = NULL;
  nnsb_converter_new (self->device_id, &self->core);
  self->sinkpad = gst_pad_new_from_static_template (&sink_factory, "sink");
  gst_pad_set_event_function (self->sinkpad, GST_DEBUG_FUNCPTR (sink_event));
  gst_pad_set_chain_function (self->sinkpad, GST_DEBUG_FUNCPTR (chain));
  gst_element_add_pad (GST_ELEMENT (self), self->sinkpad);
  self->srcpad = gst_pad_new_from_static_template (&src_factory, "src");
  gst_element_add_pad (GST_ELEMENT (self), self->srcpad);
}
