/* oracle/shim/gst/gst.h -- TEST INFRASTRUCTURE ONLY. Empty stand-in; GStreamer is not in this image. */
