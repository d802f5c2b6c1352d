/*
 * oracle/shim/nnstreamer_plugin_api.h -- TEST INFRASTRUCTURE ONLY.
 * Shadows the reference's include/nnstreamer_plugin_api.h (which needs GStreamer) with
 * prototypes for the three size helpers tensor_data.c / the mode functions call.
 * Their bodies live in oracle/ref_wrap.c and follow
 * gst/nnstreamer/nnstreamer_plugin_api_util_impl.c:155-166,1069-1091.
 */
#ifndef NNS_ORACLE_PLUGIN_API_SHIM_H
#define NNS_ORACLE_PLUGIN_API_SHIM_H
#include <glib.h>
#include <tensor_typedef.h>
gulong gst_tensor_get_element_count (const tensor_dim dim);
gsize gst_tensor_get_element_size (tensor_type type);
gsize gst_tensor_info_get_size (const GstTensorInfo * info);
#endif
