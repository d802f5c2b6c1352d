/*
 * nnsb_caps.cpp -- mini caps (see nnsb_caps.h) and the tensors-config <-> caps helpers of
 * nnstreamer_plugin_api_impl.c:712-830,925-1035,1270-1383, written against it.
 */
#include "nnsb_caps.h"

#include <climits>
#include <cstdlib>
#include <cstring>

namespace nnsb {

/* ------------------------------------------------------------------------------------- */
/* structure helpers                                                                       */
/* ------------------------------------------------------------------------------------- */
const CapsValue *
CapsStructure::get (const std::string & field) const
{
  for (const auto & f : fields)
    if (f.first == field)
      return &f.second;
  return nullptr;
}

CapsValue *
CapsStructure::get (const std::string & field)
{
  for (auto & f : fields)
    if (f.first == field)
      return &f.second;
  return nullptr;
}

void
CapsStructure::set (const std::string & field, const CapsValue & v)
{
  CapsValue *p = get (field);
  if (p)
    *p = v;
  else
    fields.push_back (std::make_pair (field, v));
}

void
CapsStructure::remove (const std::string & field)
{
  for (size_t i = 0; i < fields.size (); i++)
    if (fields[i].first == field) {
      fields.erase (fields.begin () + (long) i);
      return;
    }
}

const char *
CapsStructure::get_string (const std::string & field) const
{
  const CapsValue *v = get (field);
  return (v && v->kind == CapsValue::STRING) ? v->s.c_str () : nullptr;
}

bool
Caps::is_fixed () const
{
  if (any || structures.size () != 1)
    return false;
  for (const auto & f : structures[0].fields)
    if (!f.second.is_fixed ())
      return false;
  return true;
}

/* ------------------------------------------------------------------------------------- */
/* parser                                                                                  */
/* ------------------------------------------------------------------------------------- */
namespace {

struct Parser
{
  const std::string & t;
  size_t p = 0;
  std::string err;
  explicit Parser (const std::string & text) : t (text) {}

  void ws () { while (p < t.size () && (t[p] == ' ' || t[p] == '\t' || t[p] == '\n')) p++; }
  bool eof () { ws (); return p >= t.size (); }
  bool peek (char c) { ws (); return p < t.size () && t[p] == c; }
  bool take (char c) { if (peek (c)) { p++; return true; } return false; }

  /* a bare or quoted token; stops at , ; ] } unless quoted / escaped */
  bool token (std::string * out)
  {
    ws ();
    out->clear ();
    if (p < t.size () && t[p] == '"') {
      p++;
      while (p < t.size () && t[p] != '"') {
        if (t[p] == '\\' && p + 1 < t.size ())
          p++;
        out->push_back (t[p++]);
      }
      if (p >= t.size ()) { err = "unterminated string"; return false; }
      p++;
      return true;
    }
    while (p < t.size ()) {
      const char c = t[p];
      if (c == '\\' && p + 1 < t.size ()) { out->push_back (t[p + 1]); p += 2; continue; }
      if (c == ',' || c == ';' || c == ']' || c == '}' || c == ')' ) break;
      out->push_back (c);
      p++;
    }
    while (!out->empty () && (out->back () == ' ' || out->back () == '\t'))
      out->pop_back ();
    return !out->empty ();
  }

  bool fraction_of (const std::string & s, int64_t * n, int64_t * d)
  {
    if (s == "max") { *n = INT_MAX; *d = 1; return true; }
    if (s == "min") { *n = 0; *d = 1; return true; }
    char *end = nullptr;
    *n = strtoll (s.c_str (), &end, 10);
    if (end == s.c_str ()) return false;
    *d = 1;
    if (*end == '/') {
      const char *q = end + 1;
      *d = strtoll (q, &end, 10);
      if (end == q || *d == 0) return false;
    }
    return *end == '\0';
  }

  bool scalar (const std::string & type, const std::string & tok, CapsValue * v)
  {
    if (type == "int" || type == "i" || type == "uint") {
      char *end = nullptr;
      const long long x = strtoll (tok.c_str (), &end, 10);
      if (end == tok.c_str () || *end) { err = "bad int '" + tok + "'"; return false; }
      *v = CapsValue::integer (x);
      return true;
    }
    if (type == "fraction") {
      int64_t n, d;
      if (!fraction_of (tok, &n, &d)) { err = "bad fraction '" + tok + "'"; return false; }
      *v = CapsValue::fraction (n, d);
      return true;
    }
    *v = CapsValue::str (tok);
    return true;
  }

  bool value (std::string type, CapsValue * v)
  {
    ws ();
    if (take ('(')) {             /* explicit (type) */
      std::string ty;
      while (p < t.size () && t[p] != ')') ty.push_back (t[p++]);
      if (!take (')')) { err = "missing ')'"; return false; }
      while (!ty.empty () && ty.back () == ' ') ty.pop_back ();
      while (!ty.empty () && ty.front () == ' ') ty.erase (ty.begin ());
      type = ty;
      ws ();
    }
    if (take ('[')) {             /* range */
      std::string a, b;
      if (!token (&a) || !take (',') || !token (&b) || !take (']')) { err = "bad range"; return false; }
      if (type.empty ()) type = (a.find ('/') != std::string::npos || b == "max") ? "fraction" : "int";
      if (type == "fraction") {
        v->kind = CapsValue::FRACTION_RANGE;
        if (!fraction_of (a, &v->n0, &v->d0) || !fraction_of (b, &v->n1, &v->d1)) { err = "bad fraction range"; return false; }
      } else {
        v->kind = CapsValue::INT_RANGE;
        v->i0 = strtoll (a.c_str (), nullptr, 10);
        v->i1 = strtoll (b.c_str (), nullptr, 10);
      }
      return true;
    }
    if (take ('{')) {             /* list */
      v->kind = CapsValue::LIST;
      v->list.clear ();
      while (true) {
        CapsValue e;
        if (!value (type, &e)) return false;
        v->list.push_back (e);
        if (take (',')) continue;
        if (take ('}')) break;
        err = "bad list";
        return false;
      }
      if (v->list.size () == 1) { CapsValue only = v->list[0]; *v = only; }
      return true;
    }
    std::string tok;
    if (!token (&tok)) { err = "missing value"; return false; }
    if (type.empty ()) {
      /* untyped: int if all digits, fraction if n/d, else string (what gst_value guesses) */
      bool digits = !tok.empty (), frac = false;
      for (size_t i = 0; i < tok.size (); i++) {
        const char c = tok[i];
        if (c == '/' && i > 0) { frac = true; continue; }
        if (!(c >= '0' && c <= '9') && !(i == 0 && (c == '-' || c == '+'))) digits = false;
      }
      type = digits ? (frac ? "fraction" : "int") : "string";
    }
    return scalar (type, tok, v);
  }

  bool structure (CapsStructure * st)
  {
    ws ();
    st->name.clear ();
    while (p < t.size () && t[p] != ',' && t[p] != ';' && t[p] != ' ') st->name.push_back (t[p++]);
    if (st->name.empty ()) { err = "missing media type"; return false; }
    while (take (',')) {
      ws ();
      if (p >= t.size () || t[p] == ';') break;       /* trailing comma */
      std::string field;
      while (p < t.size () && t[p] != '=' && t[p] != ' ') field.push_back (t[p++]);
      if (!take ('=')) { err = "missing '=' after '" + field + "'"; return false; }
      CapsValue v;
      if (!value ("", &v)) return false;
      st->set (field, v);
    }
    return true;
  }
};

int64_t
gcd64 (int64_t a, int64_t b)
{
  if (a < 0) a = -a;
  if (b < 0) b = -b;
  while (b) { const int64_t t = a % b; a = b; b = t; }
  return a ? a : 1;
}

/* sign of n0/d0 - n1/d1 */
int
frac_cmp (int64_t n0, int64_t d0, int64_t n1, int64_t d1)
{
  const __int128 l = (__int128) n0 * d1, r = (__int128) n1 * d0;
  return l < r ? -1 : (l > r ? 1 : 0);
}

bool
needs_quotes (const std::string & s)
{
  if (s.empty ())
    return true;
  for (char c : s) {
    const bool ok = (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9')
        || c == '_' || c == '-' || c == '+' || c == '/' || c == ':' || c == '.';
    if (!ok)
      return true;
  }
  return false;
}

std::string
value_to_string (const CapsValue & v, bool with_type)
{
  std::string s;
  switch (v.kind) {
    case CapsValue::STRING:
      if (with_type) s += "(string)";
      if (needs_quotes (v.s)) {
        s += "\"";
        for (char c : v.s) {
          if (c == '"' || c == '\\' || c == ',') s += "\\";
          s.push_back (c);
        }
        s += "\"";
      } else {
        s += v.s;
      }
      break;
    case CapsValue::INT:
      if (with_type) s += "(int)";
      s += std::to_string (v.i0);
      break;
    case CapsValue::FRACTION:
      if (with_type) s += "(fraction)";
      s += std::to_string (v.i0) + "/" + std::to_string (v.i1);
      break;
    case CapsValue::INT_RANGE:
      if (with_type) s += "(int)";
      s += "[ " + std::to_string (v.i0) + ", " + std::to_string (v.i1) + " ]";
      break;
    case CapsValue::FRACTION_RANGE:
      if (with_type) s += "(fraction)";
      s += "[ " + std::to_string (v.n0) + "/" + std::to_string (v.d0) + ", "
          + std::to_string (v.n1) + "/" + std::to_string (v.d1) + " ]";
      break;
    case CapsValue::LIST:
    {
      const char *ty = "";
      if (!v.list.empty ()) {
        switch (v.list[0].kind) {
          case CapsValue::STRING: ty = "(string)"; break;
          case CapsValue::INT: case CapsValue::INT_RANGE: ty = "(int)"; break;
          default: ty = "(fraction)"; break;
        }
      }
      if (with_type) s += ty;
      s += "{ ";
      for (size_t i = 0; i < v.list.size (); i++) {
        if (i) s += ", ";
        s += value_to_string (v.list[i], false);
      }
      s += " }";
      break;
    }
  }
  return s;
}

bool value_intersect (const CapsValue & a, const CapsValue & b, CapsValue * out);

bool
scalar_equal (const CapsValue & a, const CapsValue & b)
{
  if (a.kind != b.kind)
    return false;
  switch (a.kind) {
    case CapsValue::STRING: return a.s == b.s;
    case CapsValue::INT: return a.i0 == b.i0;
    case CapsValue::FRACTION: return frac_cmp (a.i0, a.i1, b.i0, b.i1) == 0;
    default: return false;
  }
}

bool
list_intersect (const CapsValue & list, const CapsValue & other, CapsValue * out)
{
  CapsValue r;
  r.kind = CapsValue::LIST;
  for (const auto & e : list.list) {
    CapsValue x;
    if (value_intersect (e, other, &x))
      r.list.push_back (x);
  }
  if (r.list.empty ())
    return false;
  if (r.list.size () == 1) {
    CapsValue only = r.list[0];
    *out = only;
  } else {
    *out = r;
  }
  return true;
}

bool
value_intersect (const CapsValue & a, const CapsValue & b, CapsValue * out)
{
  if (a.kind == CapsValue::LIST)
    return list_intersect (a, b, out);
  if (b.kind == CapsValue::LIST)
    return list_intersect (b, a, out);
  if (a.is_fixed () && b.is_fixed ()) {
    if (!scalar_equal (a, b))
      return false;
    *out = a;
    return true;
  }
  /* ranges */
  if (a.kind == CapsValue::INT_RANGE && b.kind == CapsValue::INT) {
    if (b.i0 < a.i0 || b.i0 > a.i1) return false;
    *out = b;
    return true;
  }
  if (b.kind == CapsValue::INT_RANGE && a.kind == CapsValue::INT)
    return value_intersect (b, a, out);
  if (a.kind == CapsValue::INT_RANGE && b.kind == CapsValue::INT_RANGE) {
    const int64_t lo = a.i0 > b.i0 ? a.i0 : b.i0, hi = a.i1 < b.i1 ? a.i1 : b.i1;
    if (lo > hi) return false;
    if (lo == hi) { *out = CapsValue::integer (lo); return true; }
    out->kind = CapsValue::INT_RANGE; out->i0 = lo; out->i1 = hi;
    return true;
  }
  if (a.kind == CapsValue::FRACTION_RANGE && b.kind == CapsValue::FRACTION) {
    if (frac_cmp (b.i0, b.i1, a.n0, a.d0) < 0 || frac_cmp (b.i0, b.i1, a.n1, a.d1) > 0) return false;
    *out = b;
    return true;
  }
  if (b.kind == CapsValue::FRACTION_RANGE && a.kind == CapsValue::FRACTION)
    return value_intersect (b, a, out);
  if (a.kind == CapsValue::FRACTION_RANGE && b.kind == CapsValue::FRACTION_RANGE) {
    CapsValue r;
    r.kind = CapsValue::FRACTION_RANGE;
    if (frac_cmp (a.n0, a.d0, b.n0, b.d0) >= 0) { r.n0 = a.n0; r.d0 = a.d0; } else { r.n0 = b.n0; r.d0 = b.d0; }
    if (frac_cmp (a.n1, a.d1, b.n1, b.d1) <= 0) { r.n1 = a.n1; r.d1 = a.d1; } else { r.n1 = b.n1; r.d1 = b.d1; }
    const int c = frac_cmp (r.n0, r.d0, r.n1, r.d1);
    if (c > 0) return false;
    if (c == 0) { *out = CapsValue::fraction (r.n0, r.d0); return true; }
    *out = r;
    return true;
  }
  return false;
}

}  // namespace

bool
caps_parse (const std::string & text, Caps * out, std::string * err)
{
  out->structures.clear ();
  out->any = false;
  Parser ps (text);
  if (ps.eof ())
    return true;                /* empty caps */
  {
    std::string trimmed = text;
    while (!trimmed.empty () && trimmed.back () == ' ') trimmed.pop_back ();
    if (trimmed == "ANY") { out->any = true; return true; }
    if (trimmed == "EMPTY" || trimmed == "NONE") return true;
  }
  while (!ps.eof ()) {
    CapsStructure st;
    if (!ps.structure (&st)) {
      if (err) *err = ps.err;
      return false;
    }
    out->structures.push_back (st);
    if (!ps.take (';'))
      break;
  }
  if (!ps.eof ()) {
    if (err) *err = "trailing characters at offset " + std::to_string (ps.p);
    return false;
  }
  return true;
}

std::string
caps_to_string (const Caps & caps)
{
  if (caps.any)
    return "ANY";
  if (caps.structures.empty ())
    return "EMPTY";
  std::string s;
  for (size_t i = 0; i < caps.structures.size (); i++) {
    if (i) s += "; ";
    s += caps.structures[i].name;
    for (const auto & f : caps.structures[i].fields)
      s += ", " + f.first + "=" + value_to_string (f.second, true);
  }
  return s;
}

static bool
structure_intersect (const CapsStructure & a, const CapsStructure & b, CapsStructure * out)
{
  if (a.name != b.name)
    return false;
  out->name = a.name;
  out->fields.clear ();
  for (const auto & f : a.fields) {
    const CapsValue *o = b.get (f.first);
    if (!o) {
      out->fields.push_back (f);
      continue;
    }
    CapsValue v;
    if (!value_intersect (f.second, *o, &v))
      return false;
    out->fields.push_back (std::make_pair (f.first, v));
  }
  for (const auto & f : b.fields)
    if (!a.has (f.first))
      out->fields.push_back (f);
  return true;
}

Caps
caps_intersect (const Caps & a, const Caps & b)
{
  if (a.any)
    return b;
  if (b.any)
    return a;
  Caps r;
  for (const auto & sa : a.structures)
    for (const auto & sb : b.structures) {
      CapsStructure st;
      if (structure_intersect (sa, sb, &st))
        r.structures.push_back (st);
    }
  return r;
}

bool
caps_can_intersect (const Caps & a, const Caps & b)
{
  return !caps_intersect (a, b).empty ();
}

Caps
caps_fixate (const Caps & caps)
{
  Caps r;
  if (caps.any || caps.structures.empty ())
    return r;
  CapsStructure st = caps.structures[0];        /* gst_caps_truncate */
  for (auto & f : st.fields) {
    CapsValue v = f.second;
    while (v.kind == CapsValue::LIST && !v.list.empty ()) {
      CapsValue first = v.list[0];
      v = first;
    }
    if (v.kind == CapsValue::INT_RANGE)
      v = CapsValue::integer (v.i0);
    else if (v.kind == CapsValue::FRACTION_RANGE)
      v = CapsValue::fraction (v.n0, v.d0);
    f.second = v;
  }
  r.structures.push_back (st);
  return r;
}

/* ------------------------------------------------------------------------------------- */
/* tensors config                                                                          */
/* ------------------------------------------------------------------------------------- */
void
config_init (nnsb_tensors_config * c)
{
  memset (c, 0, sizeof (*c));
  c->num_tensors = 0;
  c->format = NNSB_FORMAT_STATIC;
  for (int i = 0; i < NNSB_SIZE_LIMIT; i++)
    c->info[i].type = NNSB_TYPE_END;    /* dimension 0: not configured */
  c->rate_n = -1;
  c->rate_d = -1;
}

static std::vector<std::string>
split_multi (const std::string & s)
{
  /* ',' and '.' both separate tensors (util_impl.c:502,540) */
  std::vector<std::string> out;
  std::string cur;
  for (char ch : s) {
    if (ch == ',' || ch == '.') {
      out.push_back (cur);
      cur.clear ();
    } else {
      cur.push_back (ch);
    }
  }
  out.push_back (cur);
  return out;
}

bool
config_from_structure (nnsb_tensors_config * c, const CapsStructure & st)
{
  config_init (c);
  if (st.name == "other/tensor") {
    c->num_tensors = 1;
    if (const char *d = st.get_string ("dimension"))
      nnsb_dim_parse (d, c->info[0].dim);
    if (const char *t = st.get_string ("type"))
      c->info[0].type = nnsb_type_from_string (t);
  } else if (st.name == "other/tensors") {
    if (const char *f = st.get_string ("format")) {
      const std::string fs (f);
      if (fs == "static") c->format = NNSB_FORMAT_STATIC;
      else if (fs == "flexible") c->format = NNSB_FORMAT_FLEXIBLE;
      else if (fs == "sparse") c->format = NNSB_FORMAT_SPARSE;
    }
    if (c->format == NNSB_FORMAT_STATIC) {
      const CapsValue *n = st.get ("num_tensors");
      if (n && n->kind == CapsValue::INT)
        c->num_tensors = (uint32_t) n->i0;
      if (c->num_tensors > NNSB_SIZE_LIMIT)
        c->num_tensors = NNSB_SIZE_LIMIT;
      if (const char *d = st.get_string ("dimensions")) {
        std::vector<std::string> v = split_multi (d);
        for (size_t i = 0; i < v.size () && i < NNSB_SIZE_LIMIT; i++)
          nnsb_dim_parse (v[i].c_str (), c->info[i].dim);
      }
      if (const char *t = st.get_string ("types")) {
        std::vector<std::string> v = split_multi (t);
        for (size_t i = 0; i < v.size () && i < NNSB_SIZE_LIMIT; i++)
          c->info[i].type = nnsb_type_from_string (v[i].c_str ());
      }
    }
  } else {
    return false;
  }
  const CapsValue *fr = st.get ("framerate");
  if (fr && fr->kind == CapsValue::FRACTION) {
    c->rate_n = (int32_t) fr->i0;
    c->rate_d = (int32_t) fr->i1;
  }
  return true;
}

std::string
dims_string (const nnsb_tensors_config & c, unsigned int rank)
{
  std::string s;
  char buf[128];
  for (uint32_t i = 0; i < c.num_tensors; i++) {
    nnsb_dim_to_string (c.info[i].dim, rank, buf, sizeof (buf));
    if (i) s += ",";
    s += buf;
  }
  return s;
}

std::string
types_string (const nnsb_tensors_config & c)
{
  std::string s;
  for (uint32_t i = 0; i < c.num_tensors; i++) {
    const char *t = nnsb_type_to_string (c.info[i].type);
    if (i) s += ",";
    s += t ? t : "";
  }
  return s;
}

static CapsValue
rate_range ()
{
  CapsValue v;
  v.kind = CapsValue::FRACTION_RANGE;
  v.n0 = 0; v.d0 = 1; v.n1 = INT_MAX; v.d1 = 1;        /* GST_TENSOR_RATE_RANGE */
  return v;
}

Caps
tensor_caps_from_config (const nnsb_tensors_config & c)
{
  Caps caps;
  if (c.num_tensors > 1)
    return caps;
  CapsStructure st;
  st.name = "other/tensor";
  st.set ("framerate", rate_range ());                  /* GST_TENSOR_CAP_DEFAULT */
  if (nnsb_dim_is_valid (c.info[0].dim)) {
    char buf[128];
    nnsb_dim_to_string (c.info[0].dim, 0, buf, sizeof (buf));
    st.set ("dimension", CapsValue::str (buf));
  }
  if (c.info[0].type != NNSB_TYPE_END && nnsb_type_to_string (c.info[0].type))
    st.set ("type", CapsValue::str (nnsb_type_to_string (c.info[0].type)));
  if (c.rate_n >= 0 && c.rate_d > 0)
    st.set ("framerate", CapsValue::fraction (c.rate_n, c.rate_d));
  caps.structures.push_back (st);       /* gst_caps_truncate drops the rank-4 compat structure */
  return caps;
}

Caps
tensors_caps_from_config (const nnsb_tensors_config & c)
{
  Caps caps;
  CapsStructure st;
  st.name = "other/tensors";
  if (c.format == NNSB_FORMAT_FLEXIBLE) {
    st.set ("format", CapsValue::str ("flexible"));     /* GST_TENSORS_FLEX_CAP_DEFAULT */
    st.set ("framerate", rate_range ());
  } else {
    st.set ("format", CapsValue::str ("static"));       /* GST_TENSORS_CAP_DEFAULT */
    CapsValue nr;
    nr.kind = CapsValue::INT_RANGE;
    nr.i0 = 1;
    nr.i1 = NNSB_SIZE_LIMIT;
    st.set ("num_tensors", nr);
    st.set ("framerate", rate_range ());
    if (c.num_tensors > 0) {
      st.set ("num_tensors", CapsValue::integer (c.num_tensors));
      st.set ("dimensions", CapsValue::str (dims_string (c, 0)));
      st.set ("types", CapsValue::str (types_string (c)));
    }
  }
  if (c.rate_n >= 0 && c.rate_d > 0)
    st.set ("framerate", CapsValue::fraction (c.rate_n, c.rate_d));
  caps.structures.push_back (st);
  return caps;
}

static bool
dims_same (const std::string & a, const std::string & b, bool multi)
{
  std::vector<std::string> va = multi ? split_multi (a) : std::vector<std::string> (1, a);
  std::vector<std::string> vb = multi ? split_multi (b) : std::vector<std::string> (1, b);
  if (va.size () != vb.size ())
    return false;
  for (size_t i = 0; i < va.size (); i++) {
    uint32_t d1[NNSB_RANK_LIMIT], d2[NNSB_RANK_LIMIT];
    nnsb_dim_parse (va[i].c_str (), d1);
    nnsb_dim_parse (vb[i].c_str (), d2);
    if (memcmp (d1, d2, sizeof (d1)) != 0)
      return false;
  }
  return true;
}

void
caps_update_dimension (Caps * caps, const Caps & peer)
{
  for (auto & st : caps->structures) {
    for (const auto & sp : peer.structures) {
      const char *a, *b;
      if ((a = st.get_string ("dimension")) && (b = sp.get_string ("dimension"))) {
        if (dims_same (a, b, false))
          st.set ("dimension", CapsValue::str (b));
      } else if ((a = st.get_string ("dimensions")) && (b = sp.get_string ("dimensions"))) {
        if (dims_same (a, b, true))
          st.set ("dimensions", CapsValue::str (b));
      }
    }
  }
}

bool
config_validate (const nnsb_tensors_config & c)
{
  if (c.rate_n < 0 || c.rate_d <= 0)
    return false;
  if (c.format < 0 || c.format >= NNSB_FORMAT_END)
    return false;
  if (c.format != NNSB_FORMAT_STATIC)
    return true;
  if (c.num_tensors < 1 || c.num_tensors > NNSB_SIZE_LIMIT)
    return false;
  for (uint32_t i = 0; i < c.num_tensors; i++) {
    if (c.info[i].type < 0 || c.info[i].type >= NNSB_TYPE_END)
      return false;
    if (!nnsb_dim_is_valid (c.info[i].dim))
      return false;
  }
  return true;
}

bool
config_is_equal (const nnsb_tensors_config & a, const nnsb_tensors_config & b)
{
  if (!config_validate (a) || !config_validate (b))
    return false;
  /* _compare_rate: reduce, then compare */
  const int64_t ga = gcd64 (a.rate_n, a.rate_d), gb = gcd64 (b.rate_n, b.rate_d);
  if (a.rate_n / ga != b.rate_n / gb || a.rate_d / ga != b.rate_d / gb)
    return false;
  if (a.format != b.format)
    return false;
  if (a.format != NNSB_FORMAT_STATIC)
    return true;
  if (a.num_tensors != b.num_tensors)
    return false;
  for (uint32_t i = 0; i < a.num_tensors; i++) {
    if (a.info[i].type != b.info[i].type)
      return false;
    if (memcmp (a.info[i].dim, b.info[i].dim, sizeof (a.info[i].dim)) != 0)
      return false;
  }
  return true;
}

}  // namespace nnsb
