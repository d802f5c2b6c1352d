/*
 * nnsb_core.cpp -- host core of the B200 tensor_transform path (no GLib, no CUDA).
 *
 * Tensor type / dimension helpers, the 128-byte flexible-tensor header, the per-mode option
 * grammar and the dimension conversion used in caps negotiation.  Written from the behaviour
 * of the reference (file:line cited per function); GRegex/PCRE is not available, so the six
 * grammars of gsttensor_transform.c:70-80 are hand-written deterministic matchers (the
 * transpose pattern uses a look-ahead back-reference POSIX regex cannot express anyway).
 */
#include "nnsb_internal.h"

#include <cerrno>
#include <clocale>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <locale.h>
#include <string>
#include <vector>

/* ------------------------------------------------------------------------------------- */
/* error string                                                                            */
/* ------------------------------------------------------------------------------------- */
static thread_local char g_last_error[512] = "";

int
nnsb_set_error (int code, const char *fmt, ...)
{
  va_list ap;
  va_start (ap, fmt);
  vsnprintf (g_last_error, sizeof (g_last_error), fmt, ap);
  va_end (ap);
  return code;
}

extern "C" const char *
nnsb_last_error (void)
{
  return g_last_error;
}

/* ------------------------------------------------------------------------------------- */
/* small string helpers (ASCII, locale independent like g_ascii_*)                        */
/* ------------------------------------------------------------------------------------- */
static inline char
lower (char c)
{
  return (c >= 'A' && c <= 'Z') ? (char) (c - 'A' + 'a') : c;
}

static inline bool
is_digit (char c)
{
  return c >= '0' && c <= '9';
}

static inline bool
is_space (char c)
{
  return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\f' || c == '\v';
}

static bool
ieq (const std::string & a, const char *b)
{
  size_t n = strlen (b);
  if (a.size () != n)
    return false;
  for (size_t i = 0; i < n; i++)
    if (lower (a[i]) != lower (b[i]))
      return false;
  return true;
}

/* case-insensitive "does s[pos..] start with lit" */
static bool
starts_with_ci (const std::string & s, size_t pos, const char *lit)
{
  size_t n = strlen (lit);
  if (pos + n > s.size ())
    return false;
  for (size_t i = 0; i < n; i++)
    if (lower (s[pos + i]) != lower (lit[i]))
      return false;
  return true;
}

static std::string
strip (const std::string & s)
{
  size_t b = 0, e = s.size ();
  while (b < e && is_space (s[b]))
    b++;
  while (e > b && is_space (s[e - 1]))
    e--;
  return s.substr (b, e - b);
}

/* g_strsplit (s, delim, max_tokens) for a single-character delimiter.
 * max_tokens < 1 means unlimited; an empty input yields an empty vector (GLib behaviour). */
static std::vector < std::string >
split (const std::string & s, char delim, int max_tokens)
{
  std::vector < std::string > out;
  if (s.empty ())
    return out;
  size_t start = 0;
  while (true) {
    if (max_tokens > 0 && (int) out.size () == max_tokens - 1) {
      out.push_back (s.substr (start));
      break;
    }
    size_t p = s.find (delim, start);
    if (p == std::string::npos) {
      out.push_back (s.substr (start));
      break;
    }
    out.push_back (s.substr (start, p - start));
    start = p + 1;
  }
  return out;
}

static locale_t
c_locale ()
{
  static locale_t loc = newlocale (LC_ALL_MASK, "C", (locale_t) 0);
  return loc;
}

/* g_ascii_strtod */
static double
ascii_strtod (const char *s, char **end)
{
  return strtod_l (s, end, c_locale ());
}

/* ------------------------------------------------------------------------------------- */
/* types                                                                                   */
/* ------------------------------------------------------------------------------------- */
/* tensor_element_typename[], nnstreamer_plugin_api_util_impl.c:19-33 */
static const char *const k_type_names[NNSB_TYPE_END + 1] = {
  "int32", "uint32", "int16", "uint16", "int8", "uint8",
  "float64", "float32", "int64", "uint64", "float16", nullptr
};

/* tensor_element_size[], nnstreamer_plugin_api_util_impl.c:38-51 */
static const size_t k_type_sizes[NNSB_TYPE_END + 1] = { 4, 4, 2, 2, 1, 1, 8, 4, 8, 8, 2, 0 };

/* matches (8|16|32|64) style alternatives at s[pos..]; returns chars consumed or 0 */
static size_t
match_alt (const std::string & s, size_t pos, const char *const *alts, int *which)
{
  for (int i = 0; alts[i]; i++) {
    if (starts_with_ci (s, pos, alts[i])) {
      if (which)
        *which = i;
      return strlen (alts[i]);
    }
  }
  return 0;
}

static const char *const k_int_bits[] = { "8", "16", "32", "64", nullptr };
static const char *const k_float_bits[] = { "16", "32", "64", nullptr };

/* Matches `[u]?int(8|16|32|64)|float(16|32|64)` (case-insensitive) at s[pos..].
 * Returns the number of characters consumed (0 = no match) and the type. */
static size_t
match_type_token (const std::string & s, size_t pos, int *type)
{
  size_t p = pos;
  int which = 0;
  bool uns = false;
  if (starts_with_ci (s, p, "u") && starts_with_ci (s, p + 1, "int")) {
    uns = true;
    p += 1;
  }
  if (starts_with_ci (s, p, "int")) {
    size_t n = match_alt (s, p + 3, k_int_bits, &which);
    if (n) {
      static const int st[] = { NNSB_INT8, NNSB_INT16, NNSB_INT32, NNSB_INT64 };
      static const int ut[] = { NNSB_UINT8, NNSB_UINT16, NNSB_UINT32, NNSB_UINT64 };
      if (type)
        *type = uns ? ut[which] : st[which];
      return p + 3 + n - pos;
    }
    return 0;
  }
  if (!uns && starts_with_ci (s, p, "float")) {
    size_t n = match_alt (s, p + 5, k_float_bits, &which);
    if (n) {
      static const int ft[] = { NNSB_FLOAT16, NNSB_FLOAT32, NNSB_FLOAT64 };
      if (type)
        *type = ft[which];
      return 5 + n;
    }
  }
  return 0;
}

/* gst_tensor_get_type, util_impl.c:1098-1170: strip blanks, then an anchored match */
extern "C" int
nnsb_type_from_string (const char *typestr)
{
  if (!typestr)
    return NNSB_TYPE_END;
  std::string s = strip (typestr);
  if (s.empty ())
    return NNSB_TYPE_END;
  int type = NNSB_TYPE_END;
  size_t n = match_type_token (s, 0, &type);
  if (n == 0 || n != s.size ())
    return NNSB_TYPE_END;
  return type;
}

extern "C" const char *
nnsb_type_to_string (int type)
{
  if (type < 0 || type > NNSB_TYPE_END)
    return nullptr;
  return k_type_names[type];
}

extern "C" size_t
nnsb_element_size (int type)
{
  if (type < 0 || type > NNSB_TYPE_END)
    return 0;
  return k_type_sizes[type];
}

extern "C" uint64_t
nnsb_element_count (const uint32_t * dim)
{
  uint64_t count = 1;
  for (int i = 0; i < NNSB_RANK_LIMIT; i++)
    count *= dim[i];
  return count;
}

extern "C" size_t
nnsb_info_size (const nnsb_tensor_info * info)
{
  if (!info)
    return 0;
  return (size_t) nnsb_element_count (info->dim) * nnsb_element_size (info->type);
}

/* gst_tensor_parse_dimension, util_impl.c:921-956 */
extern "C" int
nnsb_dim_parse (const char *dimstr, uint32_t * dim)
{
  if (!dimstr || !dim)
    return 0;
  std::string s = strip (dimstr);
  std::vector < std::string > strv = split (s, ':', NNSB_RANK_LIMIT);
  int rank = 0;
  size_t i = 0;
  for (; i < strv.size (); i++) {
    std::string t = strip (strv[i]);
    if (t.empty ())
      break;
    /* g_ascii_strtoull: leading digits, 0 when none */
    dim[i] = (uint32_t) strtoull (t.c_str (), nullptr, 10);
    rank = (int) i + 1;
  }
  for (; i < NNSB_RANK_LIMIT; i++)
    dim[i] = 1;
  return rank;
}

/* gst_tensor_get_dimension_string / gst_tensor_get_rank_dimension_string, util_impl.c:964-1017 */
extern "C" int
nnsb_dim_to_string (const uint32_t * dim, unsigned int rank, char *buf, size_t buflen)
{
  if (!dim || !buf)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_dim_to_string: null argument");
  unsigned int actual = (rank == 0 || rank > NNSB_RANK_LIMIT) ? NNSB_RANK_LIMIT : rank;
  std::string s;
  for (unsigned int i = 0; i < actual; i++) {
    s += std::to_string (dim[i]);
    if (i + 1 < actual)
      s += ":";
  }
  if (s.size () + 1 > buflen)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_dim_to_string: buffer too small");
  memcpy (buf, s.c_str (), s.size () + 1);
  return (int) s.size ();
}

/* gst_tensor_dimension_is_valid, util_impl.c:893-913 */
extern "C" int
nnsb_dim_is_valid (const uint32_t * dim)
{
  if (!dim)
    return 0;
  for (int i = 0; i < NNSB_RANK_LIMIT; i++)
    if (dim[i] == 0)
      return 0;
  return 1;
}

/* ------------------------------------------------------------------------------------- */
/* flexible tensor header (util_impl.c:1213-1464, 265-286)                                 */
/* ------------------------------------------------------------------------------------- */
#define META_VERSION_VALID(v) (((v) & 0xDE000000u) == 0xDE000000u)
#define META_MAKE_VERSION(major, minor) (((uint32_t) (major) << 12) | (uint32_t) (minor) | 0xDE000000u)
#define META_VERSION META_MAKE_VERSION (1, 0)
#define NNSB_MEDIA_TENSOR 4u    /* _NNS_TENSOR, tensor_typedef.h:178 */

static void
meta_init (nnsb_meta_info * meta)
{
  memset (meta, 0, sizeof (*meta));
  meta->version = META_VERSION;
  meta->type = NNSB_TYPE_END;
  meta->format = NNSB_FORMAT_STATIC;
  meta->media_type = NNSB_MEDIA_TENSOR;
}

/* gst_tensor_meta_info_validate, util_impl.c:1269-1313 */
static bool
meta_validate (const nnsb_meta_info * meta)
{
  if (!meta || !META_VERSION_VALID (meta->version))
    return false;
  if (meta->type >= NNSB_TYPE_END)
    return false;
  if (meta->dimension[0] == 0)
    return false;
  if (meta->format >= NNSB_FORMAT_END)
    return false;
  if (meta->media_type > NNSB_MEDIA_TENSOR)
    return false;
  return true;
}

extern "C" int
nnsb_meta_from_info (const nnsb_tensor_info * info, nnsb_meta_info * meta)
{
  if (!info || !meta)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_meta_from_info: null argument");
  if (info->type < 0 || info->type >= NNSB_TYPE_END || !nnsb_dim_is_valid (info->dim))
    return nnsb_set_error (NNSB_EINVAL, "nnsb_meta_from_info: invalid tensor info");
  meta_init (meta);
  meta->type = (uint32_t) info->type;
  for (int i = 0; i < NNSB_RANK_LIMIT; i++) {
    if (info->dim[i] > 0)
      meta->dimension[i] = info->dim[i];
    else
      break;
  }
  return NNSB_OK;
}

extern "C" int
nnsb_meta_to_info (const nnsb_meta_info * meta, nnsb_tensor_info * info)
{
  if (!info || !meta_validate (meta))
    return nnsb_set_error (NNSB_EINVAL, "nnsb_meta_to_info: invalid meta");
  info->type = (int32_t) meta->type;
  for (int i = 0; i < NNSB_META_RANK_LIMIT; i++) {
    if (i >= NNSB_RANK_LIMIT) {
      if (meta->dimension[i] > 0)
        return nnsb_set_error (NNSB_EINVAL,
            "nnsb_meta_to_info: meta rank exceeds %d (dimension[%d]=%u)", NNSB_RANK_LIMIT, i,
            meta->dimension[i]);
      break;
    }
    info->dim[i] = meta->dimension[i] > 0 ? meta->dimension[i] : 1;
  }
  return NNSB_OK;
}

extern "C" int
nnsb_meta_write_header (const nnsb_meta_info * meta, void *header)
{
  if (!header || !meta_validate (meta))
    return nnsb_set_error (NNSB_EINVAL, "nnsb_meta_write_header: invalid meta or null header");
  uint32_t *val = (uint32_t *) header;
  memset (header, 0, NNSB_META_HEADER_SIZE);
  /* the reference memcpy()s its struct: version, type, dimension[16], format, media_type, nnz */
  val[0] = meta->version;
  val[1] = meta->type;
  memcpy (&val[2], meta->dimension, sizeof (uint32_t) * NNSB_META_RANK_LIMIT);
  val[18] = meta->format;
  val[19] = meta->media_type;
  val[20] = meta->sparse_nnz;
  return NNSB_OK;
}

extern "C" int
nnsb_meta_parse_header (nnsb_meta_info * meta, const void *header)
{
  if (!meta || !header)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_meta_parse_header: null argument");
  const uint32_t *val = (const uint32_t *) header;
  meta_init (meta);
  meta->version = val[0];
  meta->type = val[1];
  memcpy (meta->dimension, &val[2], sizeof (uint32_t) * NNSB_META_RANK_LIMIT);
  meta->format = val[18];
  meta->media_type = val[19];
  if (meta->format == NNSB_FORMAT_SPARSE)
    meta->sparse_nnz = val[20];
  if (!meta_validate (meta))
    return nnsb_set_error (NNSB_EINVAL, "nnsb_meta_parse_header: invalid header");
  return NNSB_OK;
}

extern "C" size_t
nnsb_meta_data_size (const nnsb_meta_info * meta)
{
  if (!meta || !META_VERSION_VALID (meta->version))
    return 0;
  size_t dsize = nnsb_element_size ((int) meta->type);
  if (meta->format == NNSB_FORMAT_SPARSE)
    return meta->sparse_nnz * (dsize + sizeof (uint32_t));
  int i;
  for (i = 0; i < NNSB_META_RANK_LIMIT; i++) {
    if (meta->dimension[i] == 0)
      break;
    dsize *= meta->dimension[i];
  }
  return i > 0 ? dsize : 0;
}

/* ------------------------------------------------------------------------------------- */
/* modes                                                                                   */
/* ------------------------------------------------------------------------------------- */
static const char *const k_mode_nicks[] = {
  "dimchg", "typecast", "arithmetic", "transpose", "stand", "clamp"
};

extern "C" int
nnsb_mode_from_string (const char *nick)
{
  if (!nick)
    return NNSB_MODE_UNKNOWN;
  for (int i = 0; i < 6; i++)
    if (strcmp (nick, k_mode_nicks[i]) == 0)
      return i;
  return NNSB_MODE_UNKNOWN;
}

extern "C" const char *
nnsb_mode_to_string (int mode)
{
  if (mode >= 0 && mode < 6)
    return k_mode_nicks[mode];
  return "unknown";
}

/* ------------------------------------------------------------------------------------- */
/* option grammar                                                                          */
/* ------------------------------------------------------------------------------------- */
/* Matches `[-+]?[0-9]*\.?[0-9]+([eE][-+]?[0-9]+)?` at s[pos..] with the longest extent a
 * backtracking engine could reach; returns chars consumed (0 = no match).  Every token that
 * may follow a number in the grammars (':' '@' ',' an operator name, end) starts with a
 * character outside the number alphabet, so longest-match equals PCRE's result. */
static size_t
match_number (const std::string & s, size_t pos)
{
  size_t p = pos, n = s.size ();
  if (p < n && (s[p] == '-' || s[p] == '+'))
    p++;
  size_t int_start = p;
  while (p < n && is_digit (s[p]))
    p++;
  size_t int_digits = p - int_start;
  size_t mantissa_end = 0;
  if (p < n && s[p] == '.') {
    size_t q = p + 1;
    while (q < n && is_digit (s[q]))
      q++;
    if (q > p + 1)
      mantissa_end = q;         /* [0-9]* '.' [0-9]+ */
  }
  if (!mantissa_end) {
    if (int_digits == 0)
      return 0;
    mantissa_end = p;           /* [0-9]+ with the optional '.' absent */
  }
  p = mantissa_end;
  if (p < n && (s[p] == 'e' || s[p] == 'E')) {
    size_t q = p + 1;
    if (q < n && (s[q] == '-' || s[q] == '+'))
      q++;
    size_t d = q;
    while (q < n && is_digit (s[q]))
      q++;
    if (q > d)
      p = q;
  }
  return p - pos;
}

static size_t
match_digits (const std::string & s, size_t pos)
{
  size_t p = pos;
  while (p < s.size () && is_digit (s[p]))
    p++;
  return p - pos;
}

/* REGEX_ARITH_OPTION, gsttensor_transform.c:76-78 (case-insensitive, anchored) */
static bool
match_arith_option (const std::string & s)
{
  size_t p = 0, n = s.size ();
  /* (typecast:TYPE,)? */
  if (starts_with_ci (s, p, "typecast:")) {
    size_t t = match_type_token (s, p + 9, nullptr);
    if (t && p + 9 + t < n && s[p + 9 + t] == ',')
      p += 9 + t + 1;
    /* else: the optional group does not match; "typecast" is not an operator => fails below */
  }
  /* (per-channel:(false|true@[0-9]+),)? */
  if (starts_with_ci (s, p, "per-channel:")) {
    size_t q = p + 12;
    bool ok = false;
    if (starts_with_ci (s, q, "false")) {
      q += 5;
      ok = true;
    } else if (starts_with_ci (s, q, "true@")) {
      size_t d = match_digits (s, q + 5);
      if (d) {
        q += 5 + d;
        ok = true;
      }
    }
    if (ok && q < n && s[q] == ',')
      p = q + 1;
  }
  /* (((add|mul|div)(:NUM)+(@[0-9]+)?)(,|))+$ */
  int count = 0;
  while (p < n) {
    if (!(starts_with_ci (s, p, "add") || starts_with_ci (s, p, "mul")
            || starts_with_ci (s, p, "div")))
      return false;
    p += 3;
    int nums = 0;
    while (p < n && s[p] == ':') {
      size_t m = match_number (s, p + 1);
      if (!m)
        break;
      p += 1 + m;
      nums++;
    }
    if (nums == 0)
      return false;
    if (p < n && s[p] == '@') {
      size_t d = match_digits (s, p + 1);
      if (d)
        p += 1 + d;
      /* '@' without digits: group is optional, the stray '@' then fails the next iteration */
    }
    if (p < n && s[p] == ',')
      p++;
    count++;
  }
  return count > 0;
}

/* Removes every `typecast:TYPE` that starts at offset >= 1
 * (g_regex_match_full / g_regex_replace with start_position 1, :728-731). */
static std::string
strip_mid_typecast (const std::string & opt, bool * found)
{
  std::string out;
  *found = false;
  if (opt.empty ())
    return out;
  out.push_back (opt[0]);
  size_t p = 1;
  while (p < opt.size ()) {
    if (starts_with_ci (opt, p, "typecast:")) {
      size_t t = match_type_token (opt, p + 9, nullptr);
      if (t) {
        p += 9 + t;
        *found = true;
        continue;
      }
    }
    out.push_back (opt[p++]);
  }
  return out;
}

static int
parse_arith (const std::string & option, nnsb_option * out)
{
  out->arith_out_type = NNSB_TYPE_END;
  out->arith_per_channel = 0;
  out->arith_ch_dim = 0;
  out->n_ops = 0;

  bool had_mid = false;
  std::string str_option = strip_mid_typecast (option, &had_mid);
  if (!match_arith_option (str_option))
    return nnsb_set_error (NNSB_EINVAL,
        "arithmetic: '%s' is not valid option string: it should be in the form of [typecast:TYPE,][per-channel:(false|true@DIM),]add|mul|div:NUMBER[@CH_IDX]..., ...",
        str_option.c_str ());

  std::vector < std::string > str_operators = split (str_option, ',', -1);
  for (size_t i = 0; i < str_operators.size (); i++) {
    std::vector < std::string > str_op = split (str_operators[i], ':', -1);
    if (str_op.empty ())
      continue;                 /* "Invalid option" warning in the reference (:833) */
    size_t num_op = str_op.size ();
    std::vector < std::string > values;
    if (num_op > 1)
      values = split (str_op[1], '@', -1);
    size_t num_values = values.size ();

    if (ieq (str_op[0], "per-channel")) {
      if (num_values > 1 && ieq (values[0], "true")) {
        out->arith_per_channel = 1;
        out->arith_ch_dim = (uint32_t) strtoull (values[1].c_str (), nullptr, 10);
      }
      continue;
    }

    nnsb_arith_op op;
    memset (&op, 0, sizeof (op));
    op.applying_ch = -1;
    op.cast_type = NNSB_TYPE_END;
    if (ieq (str_op[0], "typecast"))
      op.op = NNSB_OP_TYPECAST;
    else if (ieq (str_op[0], "add"))
      op.op = NNSB_OP_ADD;
    else if (ieq (str_op[0], "mul"))
      op.op = NNSB_OP_MUL;
    else if (ieq (str_op[0], "div"))
      op.op = NNSB_OP_DIV;
    else
      op.op = NNSB_OP_UNKNOWN;

    switch (op.op) {
      case NNSB_OP_TYPECAST:
        if (num_op > 1 && num_values > 0) {
          op.cast_type = nnsb_type_from_string (values[0].c_str ());
          out->arith_out_type = op.cast_type;
        } else {
          op.op = NNSB_OP_UNKNOWN;
        }
        break;
      case NNSB_OP_ADD:
      case NNSB_OP_MUL:
      case NNSB_OP_DIV:
        if (num_op > 1 && num_values > 0) {
          const std::string & v = values[0];
          if (v.find ('.') != std::string::npos || v.find ('e') != std::string::npos
              || v.find ('E') != std::string::npos) {
            op.is_float = 1;
            op.dval = ascii_strtod (v.c_str (), nullptr);
          } else {
            op.is_float = 0;
            op.ival = strtoll (v.c_str (), nullptr, 10);
          }
          if (out->arith_per_channel && num_values > 1)
            op.applying_ch = (int32_t) strtoll (values[1].c_str (), nullptr, 10);
        } else {
          op.op = NNSB_OP_UNKNOWN;
        }
        break;
      default:
        break;
    }
    if (op.op != NNSB_OP_UNKNOWN) {
      if (out->n_ops >= NNSB_MAX_OPS)
        return nnsb_set_error (NNSB_EINVAL, "arithmetic: more than %d operators", NNSB_MAX_OPS);
      out->ops[out->n_ops++] = op;
    }
  }
  if (out->n_ops == 0)
    return nnsb_set_error (NNSB_EINVAL, "arithmetic: no operator in '%s'", option.c_str ());
  return NNSB_OK;
}

/* REGEX_STAND_OPTION, gsttensor_transform.c:73 */
static int
parse_stand (const std::string & s, nnsb_option * out)
{
  size_t p = 0, n = s.size ();
  bool ok = false;
  if (starts_with_ci (s, p, "default")) {
    p += 7;
    ok = true;
  } else if (starts_with_ci (s, p, "dc-average")) {
    p += 10;
    ok = true;
  }
  if (ok && p < n && s[p] == ':') {
    size_t t = match_type_token (s, p + 1, nullptr);
    if (t)
      p += 1 + t;
  }
  if (ok && starts_with_ci (s, p, ",per-channel:")) {
    if (starts_with_ci (s, p + 13, "true"))
      p += 13 + 4;
    else if (starts_with_ci (s, p + 13, "false"))
      p += 13 + 5;
  }
  if (!ok || p != n)
    return nnsb_set_error (NNSB_EINVAL,
        "stand: '%s' is not a valid option string: it should be in the form of (default|dc-average)[:TYPE][,per-channel:(false|true)]",
        s.c_str ());

  out->stand_out_type = NNSB_TYPE_END;
  out->stand_per_channel = 0;
  out->stand_mode = NNSB_STAND_END;
  std::vector < std::string > options = split (s, ',', -1);
  for (size_t i = 0; i < options.size (); i++) {
    std::vector < std::string > strv = split (options[i], ':', -1);
    if (strv.empty ())
      continue;
    if (ieq (strv[0], "default") || ieq (strv[0], "dc-average")) {
      out->stand_mode = ieq (strv[0], "default") ? NNSB_STAND_DEFAULT : NNSB_STAND_DC_AVERAGE;
      if (strv.size () > 1)
        out->stand_out_type = nnsb_type_from_string (strv[1].c_str ());
    } else if (ieq (strv[0], "per-channel")) {
      if (strv.size () > 1 && ieq (strv[1], "true"))
        out->stand_per_channel = 1;
    } else {
      out->stand_mode = NNSB_STAND_END;
    }
  }
  return NNSB_OK;
}

/* REGEX_CLAMP_OPTION, gsttensor_transform.c:74-75 and :911-950 */
static int
parse_clamp (const std::string & s, nnsb_option * out)
{
  size_t a = match_number (s, 0);
  if (!a || a >= s.size () || s[a] != ':')
    goto bad;
  {
    size_t b = match_number (s, a + 1);
    if (!b || a + 1 + b != s.size ())
      goto bad;
    std::string smin = s.substr (0, a), smax = s.substr (a + 1);
    errno = 0;
    out->clamp_min = ascii_strtod (smin.c_str (), nullptr);
    if (errno == ERANGE)
      return nnsb_set_error (NNSB_EINVAL, "clamp: CLAMP_MIN value has an invalid range");
    errno = 0;
    out->clamp_max = ascii_strtod (smax.c_str (), nullptr);
    if (errno == ERANGE)
      return nnsb_set_error (NNSB_EINVAL, "clamp: CLAMP_MAX value has an invalid range");
    if (out->clamp_min > out->clamp_max)
      return nnsb_set_error (NNSB_EINVAL, "clamp: CLAMP_MIN is larger than CLAMP_MAX");
    return NNSB_OK;
  }
bad:
  return nnsb_set_error (NNSB_EINVAL,
      "clamp: '%s' is not valid option string: it should be in the form of [CLAMP_MIN:CLAMP_MAX]",
      s.c_str ());
}

extern "C" int
nnsb_option_parse (int mode, const char *option, nnsb_option * out)
{
  if (!out)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_option_parse: null output");
  memset (out, 0, sizeof (*out));
  out->mode = NNSB_MODE_UNKNOWN;
  out->typecast_to = NNSB_TYPE_END;
  out->arith_out_type = NNSB_TYPE_END;
  out->stand_out_type = NNSB_TYPE_END;
  if (!option)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_option_parse: null option");
  std::string s (option);
  int ret = NNSB_EINVAL;

  switch (mode) {
    case NNSB_MODE_DIMCHG:
      /* REGEX_DIMCHG_OPTION "^([0-7]):([0-7])$", :70 */
      if (s.size () == 3 && s[0] >= '0' && s[0] <= '7' && s[1] == ':' && s[2] >= '0'
          && s[2] <= '7') {
        out->dimchg_from = s[0] - '0';
        out->dimchg_to = s[2] - '0';
        ret = NNSB_OK;
      } else {
        nnsb_set_error (NNSB_EINVAL,
            "dimchg: '%s' is not valid option string: it should be in the form of IDX_DIM_FROM:IDX_DIM_TO",
            option);
      }
      break;
    case NNSB_MODE_TYPECAST:
    {
      /* REGEX_TYPECAST_OPTION, :71 (anchored, no blanks allowed) */
      int type = NNSB_TYPE_END;
      size_t n = match_type_token (s, 0, &type);
      if (n && n == s.size ()) {
        out->typecast_to = type;
        ret = NNSB_OK;
      } else {
        nnsb_set_error (NNSB_EINVAL,
            "typecast: '%s' is not valid data type for tensor", option);
      }
      break;
    }
    case NNSB_MODE_ARITHMETIC:
      ret = parse_arith (s, out);
      break;
    case NNSB_MODE_TRANSPOSE:
      /* REGEX_TRANSPOSE_OPTION "^(?:([0-2]):(?!.*\1)){3}3$", :72: a permutation of 0,1,2 then 3 */
      if (s.size () == 7 && s[1] == ':' && s[3] == ':' && s[5] == ':' && s[6] == '3'
          && s[0] >= '0' && s[0] <= '2' && s[2] >= '0' && s[2] <= '2' && s[4] >= '0'
          && s[4] <= '2' && s[0] != s[2] && s[0] != s[4] && s[2] != s[4]) {
        out->trans_order[0] = (uint8_t) (s[0] - '0');
        out->trans_order[1] = (uint8_t) (s[2] - '0');
        out->trans_order[2] = (uint8_t) (s[4] - '0');
        out->trans_order[3] = 3;
        ret = NNSB_OK;
      } else {
        nnsb_set_error (NNSB_EINVAL,
            "transpose: '%s' is not valid option string: it should be in the form of NEW_IDX_DIM0:NEW_IDX_DIM1:NEW_IDX_DIM2:3",
            option);
      }
      break;
    case NNSB_MODE_STAND:
      ret = parse_stand (s, out);
      break;
    case NNSB_MODE_CLAMP:
      ret = parse_clamp (s, out);
      break;
    default:
      nnsb_set_error (NNSB_EINVAL, "Cannot identify mode %d", mode);
      break;
  }
  if (ret == NNSB_OK)
    out->mode = mode;
  return ret;
}

/* ------------------------------------------------------------------------------------- */
/* gst_tensor_transform_convert_dimension, gsttensor_transform.c:1850-1959                 */
/* ------------------------------------------------------------------------------------- */
extern "C" int
nnsb_convert_dimension (const nnsb_option * opt, int direction,
    const nnsb_tensor_info * in_info, nnsb_tensor_info * out_info)
{
  if (!opt || !in_info || !out_info)
    return nnsb_set_error (NNSB_EINVAL, "nnsb_convert_dimension: null argument");
  *out_info = *in_info;
  const bool sink = (direction == NNSB_PAD_SINK);

  switch (opt->mode) {
    case NNSB_MODE_DIMCHG:
    {
      const unsigned from = (unsigned) opt->dimchg_from, to = (unsigned) opt->dimchg_to;
      for (unsigned i = 0; i < NNSB_RANK_LIMIT; i++) {
        if ((i < from && i < to) || (i > from && i > to) || from == to) {
          out_info->dim[i] = in_info->dim[i];
        } else if (sink) {
          if (i == to)
            out_info->dim[i] = in_info->dim[from];
          else if (from > to)
            out_info->dim[i] = in_info->dim[i - 1];
          else
            out_info->dim[i] = in_info->dim[i + 1];
        } else {
          if (i == from)
            out_info->dim[i] = in_info->dim[to];
          else if (from > to)
            out_info->dim[i] = in_info->dim[i + 1];
          else
            out_info->dim[i] = in_info->dim[i - 1];
        }
      }
      break;
    }
    case NNSB_MODE_TYPECAST:
      out_info->type = sink ? opt->typecast_to : NNSB_TYPE_END;
      break;
    case NNSB_MODE_ARITHMETIC:
      if (opt->arith_out_type != NNSB_TYPE_END)
        out_info->type = sink ? opt->arith_out_type : NNSB_TYPE_END;
      break;
    case NNSB_MODE_TRANSPOSE:
      for (int i = 0; i < NNSB_TRANSPOSE_RANK_LIMIT; i++) {
        if (sink)
          out_info->dim[i] = in_info->dim[opt->trans_order[i]];
        else
          out_info->dim[opt->trans_order[i]] = in_info->dim[i];
      }
      break;
    case NNSB_MODE_STAND:
      if (sink) {
        if (opt->stand_out_type != NNSB_TYPE_END)
          out_info->type = opt->stand_out_type;
      } else {
        out_info->type = NNSB_TYPE_END;
      }
      break;
    case NNSB_MODE_CLAMP:
      break;
    default:
      return nnsb_set_error (NNSB_EINVAL, "nnsb_convert_dimension: unknown mode %d", opt->mode);
  }
  return NNSB_OK;
}
