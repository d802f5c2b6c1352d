/* tests/gst_stub/gst/base/gstbasetransform.h -- declaration-only stand-in, see ../gst.h */
#ifndef NNSB_GST_STUB_BASETRANSFORM_H
#define NNSB_GST_STUB_BASETRANSFORM_H
#include <gst/gst.h>

typedef struct _GstBaseTransform { GstElement element; } GstBaseTransform;
typedef struct _GstBaseTransformClass
{
  GstElementClass parent_class;
  gboolean passthrough_on_same_caps;
  GstCaps *(*transform_caps) (GstBaseTransform * trans, GstPadDirection direction, GstCaps * caps, GstCaps * filter);
  GstCaps *(*fixate_caps) (GstBaseTransform * trans, GstPadDirection direction, GstCaps * caps, GstCaps * othercaps);
  gboolean (*set_caps) (GstBaseTransform * trans, GstCaps * incaps, GstCaps * outcaps);
  gboolean (*transform_size) (GstBaseTransform * trans, GstPadDirection direction, GstCaps * caps, gsize size,
      GstCaps * othercaps, gsize * othersize);
  GstFlowReturn (*transform) (GstBaseTransform * trans, GstBuffer * inbuf, GstBuffer * outbuf);
} GstBaseTransformClass;
#define GST_TYPE_BASE_TRANSFORM ((GType) 3)
#endif
