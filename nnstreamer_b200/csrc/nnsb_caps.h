/*
 * nnsb_caps.h -- a small stand-in for GstCaps, just large enough for the caps negotiation of
 * tensor_transform / tensor_converter (internal; the C-ABI passes caps as strings).
 *
 * GStreamer is not installed in this image, so the BaseTransform vfuncs transform_caps /
 * fixate_caps / set_caps (gsttensor_transform.c:1971-2164) are mirrored on top of this.
 * In a real GStreamer build the glue works on GstCaps and only the config-level helpers
 * (config <-> fields) are shared.  Supported: structures with string / int / fraction
 * values, int and fraction ranges, lists; parse, serialise, intersect, fixate, truncate.
 */
#ifndef NNSB_CAPS_H
#define NNSB_CAPS_H

#include <cstdint>
#include <string>
#include <utility>
#include <vector>

#include "nnsb_internal.h"

namespace nnsb {

struct CapsValue
{
  enum Kind { STRING, INT, FRACTION, INT_RANGE, FRACTION_RANGE, LIST } kind = STRING;
  std::string s;                        /* STRING */
  int64_t i0 = 0, i1 = 0;               /* INT (i0); INT_RANGE [i0, i1]; FRACTION i0/i1 */
  int64_t n0 = 0, d0 = 1, n1 = 0, d1 = 1;       /* FRACTION_RANGE [n0/d0, n1/d1] */
  std::vector<CapsValue> list;          /* LIST */

  static CapsValue str (const std::string & v) { CapsValue c; c.kind = STRING; c.s = v; return c; }
  static CapsValue integer (int64_t v) { CapsValue c; c.kind = INT; c.i0 = v; return c; }
  static CapsValue fraction (int64_t n, int64_t d) { CapsValue c; c.kind = FRACTION; c.i0 = n; c.i1 = d; return c; }
  bool is_fixed () const { return kind == STRING || kind == INT || kind == FRACTION; }
};

struct CapsStructure
{
  std::string name;
  std::vector<std::pair<std::string, CapsValue> > fields;       /* insertion order kept */

  const CapsValue *get (const std::string & field) const;
  CapsValue *get (const std::string & field);
  bool has (const std::string & field) const { return get (field) != nullptr; }
  void set (const std::string & field, const CapsValue & v);
  void remove (const std::string & field);
  const char *get_string (const std::string & field) const;     /* nullptr unless a fixed string */
};

struct Caps
{
  std::vector<CapsStructure> structures;
  bool any = false;

  bool empty () const { return !any && structures.empty (); }
  bool is_fixed () const;
};

/* gst_caps_from_string / gst_caps_to_string */
bool caps_parse (const std::string & text, Caps * out, std::string * err);
std::string caps_to_string (const Caps & caps);
/* gst_caps_intersect_full (a, b, GST_CAPS_INTERSECT_FIRST) */
Caps caps_intersect (const Caps & a, const Caps & b);
/* gst_caps_truncate + gst_caps_fixate */
Caps caps_fixate (const Caps & caps);
bool caps_can_intersect (const Caps & a, const Caps & b);

/* ---- tensors config <-> caps (nnstreamer_plugin_api_impl.c) ------------------------ */
void config_init (nnsb_tensors_config * c);                     /* gst_tensors_config_init */
/* gst_tensors_config_from_structure, :1312-1383 */
bool config_from_structure (nnsb_tensors_config * c, const CapsStructure & st);
/* gst_tensor_caps_from_config, :1296-1304 / gst_tensors_caps_from_config, :1270-1286 */
Caps tensor_caps_from_config (const nnsb_tensors_config & c);
Caps tensors_caps_from_config (const nnsb_tensors_config & c);
/* gst_tensor_caps_update_dimension, :999-1035 */
void caps_update_dimension (Caps * caps, const Caps & peer);
/* gst_tensors_config_validate / _is_equal (util_impl.c) */
bool config_validate (const nnsb_tensors_config & c);
bool config_is_equal (const nnsb_tensors_config & a, const nnsb_tensors_config & b);
/* gst_tensors_info_get_dimensions_string / _types_string */
std::string dims_string (const nnsb_tensors_config & c, unsigned int rank);
std::string types_string (const nnsb_tensors_config & c);

}  // namespace nnsb
#endif
