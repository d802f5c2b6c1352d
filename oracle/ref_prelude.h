/*
 * oracle/ref_prelude.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Stub declarations placed in front of the line ranges that oracle/build_ref.py extracts
 * from /root/reference/gst/nnstreamer/elements/gsttensor_transform.{h,c}, so that the
 * reference's six per-mode functions compile without GLib/GStreamer.  Nothing here
 * restates reference arithmetic; it only names the framework types those lines mention.
 */
#ifndef NNS_ORACLE_REF_PRELUDE_H
#define NNS_ORACLE_REF_PRELUDE_H

#include <math.h>
#include <errno.h>
#include <glib.h>
#include <tensor_typedef.h>
#include <nnstreamer_plugin_api.h>   /* the shim: three size helpers */

typedef struct { int unused; } GstBaseTransform;
typedef enum { GST_FLOW_OK = 0, GST_FLOW_ERROR = -5, GST_FLOW_NOT_SUPPORTED = -6 } GstFlowReturn;
typedef enum { GST_PAD_UNKNOWN = 0, GST_PAD_SRC = 1, GST_PAD_SINK = 2 } GstPadDirection;

#define GST_ERROR_OBJECT(...) do { } while (0)
#define GST_WARNING_OBJECT(...) do { } while (0)
#define GST_INFO_OBJECT(...) do { } while (0)
#define ml_logi(...) do { } while (0)
#define ml_logw(...) do { } while (0)
#define ml_loge(...) do { } while (0)
#define ml_logd(...) do { } while (0)
#define ml_logf(...) do { } while (0)
#define UNUSED(expr) do { (void)(expr); } while (0)
#define nns_memcpy memcpy
#define GINT_TO_POINTER(i) ((gpointer) (glong) (i))

static GList *
g_list_find (GList * list, gconstpointer data)
{
  while (list) {
    if (list->data == data)
      break;
    list = list->next;
  }
  return list;
}

/* follows gst_tensor_info_copy (nnstreamer_plugin_api_util_impl.c): type + dimension (name unused here) */
static void
gst_tensor_info_copy (GstTensorInfo * dest, const GstTensorInfo * src)
{
  guint i;
  dest->name = NULL;
  dest->type = src->type;
  for (i = 0; i < NNS_TENSOR_RANK_LIMIT; i++)
    dest->dimension[i] = src->dimension[i];
}

#endif
