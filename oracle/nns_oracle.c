/*
 * oracle/nns_oracle.c -- TEST INFRASTRUCTURE ONLY.  Never linked into, imported by or
 * executed from the product (nnstreamer_b200/); only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference leg may load it.
 *
 * A plain-C restatement of the reference's CPU algorithm for the tensor_transform hot path
 * and tensor_converter's video stride strip, written from the behaviour of
 *   gst/nnstreamer/elements/gsttensor_transform.c   (NNStreamer 2.3.0)
 *   gst/nnstreamer/tensor_data.c
 *   gst/nnstreamer/elements/nnstreamer-orc.orc      (acceleration=true semantics)
 *   gst/nnstreamer/elements/gsttensor_converter.c
 * Each function cites the lines it follows.
 *
 * Parity pinning:
 *   - acceleration=false paths are pinned against the reference's OWN code compiled here
 *     (oracle/_ref/libnns_ref.so, see oracle/build_ref.py) by tests/test_oracle_vs_ref.py
 *     and against the golden vectors of the reference's unit tests (tests/golden/).
 *   - acceleration=true (ORC) is restated from the .orc programs and ORC's published opcode
 *     semantics (liborc >= 0.4.17, un-vendored, meson.build:174-175).  In-range results are
 *     pinned by the reference's own orc* unit-test vectors; overflow / saturation behaviour
 *     is "parity unpinned" -- the reference asserts nothing there.
 *   - dimchg with from > 0: the reference over-runs its buffers (copyblocklimit, :1162-1163);
 *     this oracle implements the intended permutation of convert_dimension (:1869-1882).
 *
 * float -> integer casts are undefined behaviour in C when out of range; the reference, as
 * compiled by gcc for x86-64, executes cvttss2si/cvttsd2si, whose out-of-range result is the
 * "integer indefinite" value.  That behaviour is written out explicitly here.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { T_I32 = 0, T_U32, T_I16, T_U16, T_I8, T_U8, T_F64, T_F32, T_I64, T_U64, T_F16, T_END };
enum { OP_TYPECAST = 0, OP_ADD = 1, OP_MUL = 2, OP_DIV = 3 };
#define RANK 8

typedef _Float16 f16;

typedef struct
{
  int op;                       /* OP_* */
  int applying_ch;              /* -1: every channel */
  int is_float;                 /* literal class, gsttensor_transform.c:796-807 */
  int cast_type;
  double dval;
  int64_t ival;
} oracle_op_t;

/* tagged scalar, the role tensor_data_s plays in the reference (tensor_data.h:23-27) */
typedef struct
{
  int type;
  union
  {
    int32_t i32; uint32_t u32; int16_t i16; uint16_t u16; int8_t i8; uint8_t u8;
    double f64; float f32; int64_t i64; uint64_t u64; f16 h;
  } v;
} val_t;

static const size_t k_size[T_END + 1] = { 4, 4, 2, 2, 1, 1, 8, 4, 8, 8, 2, 0 };

size_t
oracle_element_size (int type)
{
  return (type >= 0 && type <= T_END) ? k_size[type] : 0;
}

static uint64_t
count_of (const uint32_t * dim)
{
  uint64_t n = 1;
  int i;
  for (i = 0; i < RANK; i++)
    n *= dim[i];
  return n;
}

/* x86 cvttsd2si r32 / r64 */
static int32_t
cvtt32 (double d)
{
  if (d > -2147483649.0 && d < 2147483648.0)
    return (int32_t) d;
  return INT32_MIN;
}

static int64_t
cvtt64 (double d)
{
  if (d >= -9223372036854775808.0 && d < 9223372036854775808.0)
    return (int64_t) d;
  return INT64_MIN;
}

static int
is_float_type (int t)
{
  return t == T_F32 || t == T_F64 || t == T_F16;
}

static void
val_load (val_t * x, int type, const void *p)
{
  x->type = type;
  x->v.u64 = 0;
  memcpy (&x->v, p, k_size[type]);
}

static void
val_store (const val_t * x, void *p)
{
  memcpy (p, &x->v, k_size[x->type]);
}

/* value of x as double / int64 / uint64 (exact for the respective source classes) */
static double
val_as_double (const val_t * x)
{
  switch (x->type) {
    case T_I32: return (double) x->v.i32;
    case T_U32: return (double) x->v.u32;
    case T_I16: return (double) x->v.i16;
    case T_U16: return (double) x->v.u16;
    case T_I8: return (double) x->v.i8;
    case T_U8: return (double) x->v.u8;
    case T_F64: return x->v.f64;
    case T_F32: return (double) x->v.f32;
    case T_I64: return (double) x->v.i64;
    case T_U64: return (double) x->v.u64;
    case T_F16: return (double) x->v.h;
    default: return 0.0;
  }
}

/* C cast, gst_tensor_data_typecast (tensor_data.c:202-279):
 * a floating source reaches an unsigned target through the signed type of the same width */
static void
c_cast (val_t * x, int to)
{
  val_t r;
  r.type = to;
  r.v.u64 = 0;
  if (x->type == to)
    return;
  if (is_float_type (x->type)) {
    /* float16/float32 widen exactly to double; double -> float rounds once */
    const double d = val_as_double (x);
    switch (to) {
      case T_I32: r.v.i32 = cvtt32 (d); break;
      case T_U32: r.v.u32 = (uint32_t) cvtt32 (d); break;
      case T_I16: r.v.i16 = (int16_t) cvtt32 (d); break;
      case T_U16: r.v.u16 = (uint16_t) (int16_t) cvtt32 (d); break;
      case T_I8: r.v.i8 = (int8_t) cvtt32 (d); break;
      case T_U8: r.v.u8 = (uint8_t) (int8_t) cvtt32 (d); break;
      case T_F64: r.v.f64 = d; break;
      case T_F32: r.v.f32 = (x->type == T_F64) ? (float) x->v.f64 : (float) d; break;
      case T_I64: r.v.i64 = cvtt64 (d); break;
      case T_U64: r.v.u64 = (uint64_t) cvtt64 (d); break;
      case T_F16: r.v.h = (x->type == T_F64) ? (f16) x->v.f64 : (f16) x->v.f32; break;
      default: break;
    }
  } else {
    const int is_u64 = (x->type == T_U64);
    int64_t s = 0;
    uint64_t u = 0;
    switch (x->type) {
      case T_I32: s = x->v.i32; break;
      case T_U32: s = x->v.u32; break;
      case T_I16: s = x->v.i16; break;
      case T_U16: s = x->v.u16; break;
      case T_I8: s = x->v.i8; break;
      case T_U8: s = x->v.u8; break;
      case T_I64: s = x->v.i64; break;
      case T_U64: u = x->v.u64; s = (int64_t) u; break;
      default: break;
    }
    if (!is_u64)
      u = (uint64_t) s;
    switch (to) {
      case T_I32: r.v.i32 = (int32_t) u; break;
      case T_U32: r.v.u32 = (uint32_t) u; break;
      case T_I16: r.v.i16 = (int16_t) u; break;
      case T_U16: r.v.u16 = (uint16_t) u; break;
      case T_I8: r.v.i8 = (int8_t) u; break;
      case T_U8: r.v.u8 = (uint8_t) u; break;
      case T_F64: r.v.f64 = is_u64 ? (double) u : (double) s; break;
      case T_F32: r.v.f32 = is_u64 ? (float) u : (float) s; break;
      case T_I64: r.v.i64 = (int64_t) u; break;
      case T_U64: r.v.u64 = u; break;
      case T_F16: r.v.h = is_u64 ? (f16) u : (f16) s; break;
      default: break;
    }
  }
  *x = r;
}

/* ---- ORC opcode restatements (only those nnstreamer-orc.orc uses) ------------------- */
static int16_t orc_convssslw (int32_t x) { return (int16_t) (x < -32768 ? -32768 : (x > 32767 ? 32767 : x)); }
static int8_t orc_convssswb (int16_t x) { return (int8_t) (x < -128 ? -128 : (x > 127 ? 127 : x)); }

static int32_t
orc_convfl (float f)
{
  uint32_t bits;
  int32_t t = cvtt32 ((double) f);
  memcpy (&bits, &f, 4);
  if (t == INT32_MIN && !(bits & 0x80000000u))
    t = INT32_MAX;
  return t;
}

static int32_t
orc_convdl (double d)
{
  uint64_t bits;
  int32_t t = cvtt32 (d);
  memcpy (&bits, &d, 8);
  if (t == INT32_MIN && !(bits >> 63))
    t = INT32_MAX;
  return t;
}

static int
orc_type_ok (int t)
{
  return t != T_I64 && t != T_U64 && t != T_F16;
}

/* nns_orc_conv_<from>_to_<to>; both types must satisfy orc_type_ok */
static void
orc_cast (val_t * x, int to)
{
  val_t r;
  const int from = x->type;
  r.type = to;
  r.v.u64 = 0;
  if (from == T_F32 || from == T_F64) {
    if (to == T_F32 || to == T_F64) {
      c_cast (x, to);           /* copyl / convfd / convdf / copyq */
      return;
    } else {
      const int32_t l = (from == T_F32) ? orc_convfl (x->v.f32) : orc_convdl (x->v.f64);
      switch (to) {
        case T_I32: r.v.i32 = l; break;
        case T_U32: r.v.u32 = (uint32_t) l; break;
        case T_I16: r.v.i16 = orc_convssslw (l); break;
        case T_U16: r.v.u16 = (uint16_t) orc_convssslw (l); break;
        case T_I8: r.v.i8 = orc_convssswb (orc_convssslw (l)); break;
        case T_U8: r.v.u8 = (uint8_t) l; break;               /* convlw, convwb */
        default: break;
      }
    }
  } else if (to == T_F32 || to == T_F64) {
    /* widen by the source signedness, then convlf / convld on a SIGNED 32-bit value */
    int32_t l = 0;
    switch (from) {
      case T_I32: l = x->v.i32; break;
      case T_U32: l = (int32_t) x->v.u32; break;
      case T_I16: l = x->v.i16; break;
      case T_U16: l = x->v.u16; break;
      case T_I8: l = x->v.i8; break;
      case T_U8: l = x->v.u8; break;
      default: break;
    }
    if (to == T_F32)
      r.v.f32 = (float) l;
    else
      r.v.f64 = (double) l;
  } else {
    const size_t fs = k_size[from], ts = k_size[to];
    if (ts >= fs) {
      c_cast (x, to);           /* copy / conv[su]bw / conv[su]wl: plain widening */
      return;
    }
    if (fs == 4) {
      const int32_t l = (from == T_I32) ? x->v.i32 : (int32_t) x->v.u32;
      switch (to) {
        case T_I16: r.v.i16 = orc_convssslw (l); break;
        case T_U16: r.v.u16 = (uint16_t) orc_convssslw (l); break;
        case T_I8: r.v.i8 = orc_convssswb (orc_convssslw (l)); break;
        case T_U8: r.v.u8 = (uint8_t) l; break;
        default: break;
      }
    } else {
      const int16_t w = (from == T_I16) ? x->v.i16 : (int16_t) x->v.u16;
      if (to == T_I8)
        r.v.i8 = orc_convssswb (w);
      else
        r.v.u8 = (uint8_t) w;
    }
  }
  *x = r;
}

/* one operator on a value of the working type; b is already cast to that type.
 * C flavour: handle_operator (gsttensor_transform.c:563-582).
 * ORC flavour: nns_orc_{add,mul}_c_* (saturating), division as the C loop (:529-535). */
static void
apply_op (val_t * a, const val_t * b, int op, int orc)
{
  switch (a->type) {
#define INT_CASE(T, f, W, lo, hi, U) \
    case T: \
      if (op == OP_ADD) { \
        if (orc) { W s = (W) a->v.f + (W) b->v.f; a->v.f = s < (W) (lo) ? (lo) : (s > (W) (hi) ? (hi) : s); } \
        else a->v.f = (U) ((U) a->v.f + (U) b->v.f); \
      } else if (op == OP_MUL) { \
        if (orc) { W s = (W) a->v.f * (W) b->v.f; a->v.f = s < (W) (lo) ? (lo) : (s > (W) (hi) ? (hi) : s); } \
        else a->v.f = (U) ((U) a->v.f * (U) b->v.f); \
      } else if (op == OP_DIV) { \
        if (b->v.f == 0) break;      /* C: skipped; ORC: SIGFPE in the reference, skipped here */ \
        if ((lo) != 0 && b->v.f == (__typeof__ (a->v.f)) -1) a->v.f = (U) (0 - (U) a->v.f); \
        else a->v.f = a->v.f / b->v.f; \
      } \
      break;
      INT_CASE (T_I8, i8, int32_t, -128, 127, uint8_t)
      INT_CASE (T_U8, u8, int32_t, 0, 255, uint8_t)
      INT_CASE (T_I16, i16, int32_t, -32768, 32767, uint16_t)
      INT_CASE (T_U16, u16, int64_t, 0, 65535, uint16_t)
      INT_CASE (T_I32, i32, int64_t, INT32_MIN, INT32_MAX, uint32_t)
#undef INT_CASE
    case T_U32:
      if (op == OP_ADD) {
        if (orc) { uint64_t s = (uint64_t) a->v.u32 + b->v.u32; a->v.u32 = s > 0xffffffffull ? 0xffffffffu : (uint32_t) s; }
        else a->v.u32 += b->v.u32;
      } else if (op == OP_MUL) {
        if (orc) { uint64_t s = (uint64_t) a->v.u32 * b->v.u32; a->v.u32 = s > 0xffffffffull ? 0xffffffffu : (uint32_t) s; }
        else a->v.u32 *= b->v.u32;
      } else if (op == OP_DIV) {
        if (b->v.u32 != 0) a->v.u32 /= b->v.u32;
      }
      break;
    case T_I64:                 /* never ORC (orc_supported, :462-463) */
      if (op == OP_ADD) a->v.i64 = (int64_t) ((uint64_t) a->v.i64 + (uint64_t) b->v.i64);
      else if (op == OP_MUL) a->v.i64 = (int64_t) ((uint64_t) a->v.i64 * (uint64_t) b->v.i64);
      else if (op == OP_DIV) {
        if (b->v.i64 == 0) break;
        if (b->v.i64 == -1) a->v.i64 = (int64_t) (0 - (uint64_t) a->v.i64);
        else a->v.i64 /= b->v.i64;
      }
      break;
    case T_U64:
      if (op == OP_ADD) a->v.u64 += b->v.u64;
      else if (op == OP_MUL) a->v.u64 *= b->v.u64;
      else if (op == OP_DIV) { if (b->v.u64 != 0) a->v.u64 /= b->v.u64; }
      break;
#define FLT_CASE(T, f, CT) \
    case T: \
      if (op == OP_ADD) a->v.f = (CT) (a->v.f + b->v.f); \
      else if (op == OP_MUL) a->v.f = (CT) (a->v.f * b->v.f); \
      else if (op == OP_DIV) { \
        if (!orc && b->v.f == 0) break;       /* C path only (:572-575) */ \
        a->v.f = (CT) (a->v.f / b->v.f); \
      } \
      break;
      FLT_CASE (T_F32, f32, float)
      FLT_CASE (T_F64, f64, double)
      FLT_CASE (T_F16, h, f16)
#undef FLT_CASE
    default:
      break;
  }
}

/* operand literal -> working type, once (gst_tensor_data_typecast on op_s->value,
 * gsttensor_transform.c:1268,1349,1394) */
static val_t
operand_of (const oracle_op_t * op, int type)
{
  val_t v;
  if (op->is_float) {
    v.type = T_F64;
    v.v.u64 = 0;
    v.v.f64 = op->dval;
  } else {
    v.type = T_I64;
    v.v.i64 = op->ival;
  }
  c_cast (&v, type);
  return v;
}

static int
orc_supported (int accel, int it, int ot)
{
  /* gsttensor_transform.c:462-463 */
  return accel && !(it == T_I64 || it == T_U64 || ot == T_I64 || ot == T_U64);
}

/* conversion used on the element path for the given policy.  With acceleration the float16
 * conversions run the C loops _conv_to_f16 / _conv_from_f16 (:366-428). */
static void
convert (val_t * x, int to, int orc)
{
  if (orc && orc_type_ok (x->type) && orc_type_ok (to))
    orc_cast (x, to);
  else
    c_cast (x, to);
}

/* ------------------------------------------------------------------------------------- */
/* typecast: gst_tensor_transform_typecast, gsttensor_transform.c:1202-1229                */
/* ------------------------------------------------------------------------------------- */
int
oracle_typecast (const void *in, void *out, uint64_t n, int itype, int otype, int accel)
{
  const int orc = orc_supported (accel, itype, otype);
  const size_t is = k_size[itype], os = k_size[otype];
  uint64_t i;
  for (i = 0; i < n; i++) {
    val_t x;
    val_load (&x, itype, (const char *) in + i * is);
    convert (&x, otype, orc);
    val_store (&x, (char *) out + i * os);
  }
  return 0;
}

/* ------------------------------------------------------------------------------------- */
/* arithmetic: gst_tensor_transform_arithmetic, gsttensor_transform.c:1240-1412            */
/* ------------------------------------------------------------------------------------- */
int
oracle_arithmetic (const void *in, void *out, int itype, int otype, const uint32_t * dim,
    int per_channel, unsigned int ch_dim, int n_ops, const oracle_op_t * ops, int accel)
{
  const int orc = orc_supported (accel, itype, otype);
  const uint64_t n = count_of (dim);
  const size_t is = k_size[itype], os = k_size[otype];
  uint64_t ch_size = 1, nch = 1, i;
  unsigned int d;
  int k;
  val_t *operands;

  if (per_channel) {
    if (ch_dim >= RANK)
      return -1;
    for (d = 0; d < ch_dim; d++)
      ch_size *= dim[d];        /* :1317-1322 */
    nch = dim[ch_dim];
  }
  operands = (val_t *) calloc (n_ops > 0 ? n_ops : 1, sizeof (val_t));
  for (k = 0; k < n_ops; k++)
    if (ops[k].op != OP_TYPECAST)
      operands[k] = operand_of (&ops[k], otype);

  for (i = 0; i < n; i++) {
    const int64_t ch = per_channel ? (int64_t) ((i / ch_size) % nch) : -1;
    val_t x;
    val_load (&x, itype, (const char *) in + i * is);
    /* the (only) typecast sits at the head of the list (:728-737); without one the working
     * type is the input type, which then equals otype */
    convert (&x, otype, orc);
    for (k = 0; k < n_ops; k++) {
      if (ops[k].op == OP_TYPECAST)
        continue;
      if (per_channel && ops[k].applying_ch != -1 && ops[k].applying_ch != ch)
        continue;               /* :1351 */
      apply_op (&x, &operands[k], ops[k].op, orc);
    }
    val_store (&x, (char *) out + i * os);
  }
  free (operands);
  return 0;
}

/* ------------------------------------------------------------------------------------- */
/* transpose: gst_tensor_transform_transpose, gsttensor_transform.c:1417-1500              */
/* out.dim[i] = in.dim[order[i]] for i < 4 (:1924-1928); order[3] == 3                     */
/* ------------------------------------------------------------------------------------- */
int
oracle_transpose (const void *in, void *out, int type, const uint32_t * in_dim,
    const uint8_t * order)
{
  const size_t es = k_size[type];
  uint64_t istride[4], odim[4], hi = 1, o[4], blk, outidx = 0;
  int i;
  istride[0] = 1;
  for (i = 1; i < 4; i++)
    istride[i] = istride[i - 1] * in_dim[i - 1];
  for (i = 0; i < 4; i++)
    odim[i] = in_dim[order[i]];
  for (i = 4; i < RANK; i++)
    hi *= in_dim[i];
  /* the reference loop only walks the first four dims; higher dims are 1 in its tests.
   * Blocks of the higher dims are handled one after the other here. */
  for (blk = 0; blk < hi; blk++) {
    const uint64_t base = blk * istride[3] * in_dim[3];
    for (o[3] = 0; o[3] < odim[3]; o[3]++)
      for (o[2] = 0; o[2] < odim[2]; o[2]++)
        for (o[1] = 0; o[1] < odim[1]; o[1]++)
          for (o[0] = 0; o[0] < odim[0]; o[0]++) {
            uint64_t inidx = base;
            for (i = 0; i < 4; i++)
              inidx += o[i] * istride[order[i]];
            memcpy ((char *) out + outidx * es, (const char *) in + inidx * es, es);
            outidx++;
          }
  }
  return 0;
}

/* ------------------------------------------------------------------------------------- */
/* dimchg: gst_tensor_transform_dimchg, gsttensor_transform.c:1120-1191, dims :1869-1882   */
/* ------------------------------------------------------------------------------------- */
int
oracle_dimchg (const void *in, void *out, int type, const uint32_t * in_dim, int from, int to)
{
  const size_t es = k_size[type];
  uint64_t istride[RANK], odim[RANK], src_stride[RANK], o[RANK], n, outidx;
  int i;
  if (from == to) {
    memcpy (out, in, count_of (in_dim) * es);   /* :1133-1139 */
    return 0;
  }
  if (from > to)
    return -5;                  /* GST_FLOW_ERROR, :1179-1188 */
  istride[0] = 1;
  for (i = 1; i < RANK; i++)
    istride[i] = istride[i - 1] * in_dim[i - 1];
  for (i = 0; i < RANK; i++) {
    int src;
    if (i < from || i > to)
      src = i;
    else if (i == to)
      src = from;
    else
      src = i + 1;
    odim[i] = in_dim[src];
    src_stride[i] = istride[src];
  }
  n = count_of (in_dim);
  memset (o, 0, sizeof (o));
  for (outidx = 0; outidx < n; outidx++) {
    uint64_t inidx = 0;
    for (i = 0; i < RANK; i++)
      inidx += o[i] * src_stride[i];
    memcpy ((char *) out + outidx * es, (const char *) in + inidx * es, es);
    for (i = 0; i < RANK; i++) {
      if (++o[i] < odim[i])
        break;
      o[i] = 0;
    }
  }
  return 0;
}

/* ------------------------------------------------------------------------------------- */
/* stand: gst_tensor_transform_stand (:1512-1619) with the statistics of                   */
/* tensor_data.c:314-493 (running mean; sum of pow (x - mean, 2) / N; 1e-10 if zero)       */
/* ------------------------------------------------------------------------------------- */
static double
elem_as_double (const void *base, int type, uint64_t idx)
{
  val_t x;
  val_load (&x, type, (const char *) base + idx * k_size[type]);
  return val_as_double (&x);
}

int
oracle_stand_stats (const void *in, int itype, const uint32_t * dim, int per_channel,
    double *mean, double *std)
{
  const uint64_t total = count_of (dim);
  const uint64_t nch = per_channel ? dim[0] : 1;
  const uint64_t num = total / nch;
  uint64_t ch, i;
  for (ch = 0; ch < nch; ch++) {
    double average = 0.0, s = 0.0;
    for (i = 0; i < num; i++) {
      const double value = elem_as_double (in, itype, i * nch + ch);
      average = (value - average) / (double) (i + 1) + average;      /* :342,:390 */
    }
    mean[ch] = average;
    if (std) {
      for (i = 0; i < num; i++) {
        const double value = elem_as_double (in, itype, i * nch + ch);
        s += pow (value - average, 2) / (double) num;                 /* :435,:485 */
      }
      std[ch] = (s != 0.0) ? sqrt (s) : 1e-10;                       /* :438,:488 */
    }
  }
  return 0;
}

int
oracle_stand (const void *in, void *out, int itype, int otype, const uint32_t * dim, int mode,
    int per_channel)
{
  const uint64_t total = count_of (dim);
  const uint64_t nch = per_channel ? dim[0] : 1;
  double *mean = (double *) calloc (nch, sizeof (double));
  double *std = (double *) calloc (nch, sizeof (double));
  uint64_t i;
  if (mode != 0 && mode != 1) {
    free (mean);
    free (std);
    return -5;
  }
  oracle_stand_stats (in, itype, dim, per_channel, mean, mode == 0 ? std : NULL);
  for (i = 0; i < total; i++) {
    const uint64_t ch = i % nch;
    val_t y;
    double tmp = elem_as_double (in, itype, i);
    if (mode == 0)
      tmp = fabs ((tmp - mean[ch]) / std[ch]);    /* :1556,:1569 */
    else
      tmp -= mean[ch];                            /* :1587,:1600 */
    y.type = T_F64;
    y.v.u64 = 0;
    y.v.f64 = tmp;
    c_cast (&y, otype);
    val_store (&y, (char *) out + i * k_size[otype]);
  }
  free (mean);
  free (std);
  return 0;
}

/* ------------------------------------------------------------------------------------- */
/* clamp: gst_tensor_transform_clamp, gsttensor_transform.c:1632-1658                      */
/* ------------------------------------------------------------------------------------- */
int
oracle_clamp (const void *in, void *out, int type, uint64_t n, double lo, double hi)
{
  uint64_t i;
  for (i = 0; i < n; i++) {
    val_t y;
    double tmp = elem_as_double (in, type, i);
    tmp = (tmp > hi) ? hi : ((tmp < lo) ? lo : tmp);    /* glib CLAMP */
    y.type = T_F64;
    y.v.u64 = 0;
    y.v.f64 = tmp;
    c_cast (&y, type);
    val_store (&y, (char *) out + i * k_size[type]);
  }
  return 0;
}

/* ------------------------------------------------------------------------------------- */
/* tensor_converter video stride strip, gsttensor_converter.c:1075-1100:                   */
/* size = C*W bytes per row, source pitch = size rounded up to 4                           */
/* ------------------------------------------------------------------------------------- */
int
oracle_stride_strip (const void *src, void *dst, uint32_t color, uint32_t width, uint32_t height)
{
  const size_t size = (size_t) color * width;
  size_t offset = size;
  uint32_t h;
  if (offset % 4)
    offset += 4 - (offset % 4);
  for (h = 0; h < height; h++)
    memcpy ((char *) dst + h * size, (const char *) src + h * offset, size);
  return 0;
}

/* ------------------------------------------------------------------------------------- */
/* tensor_aggregator: the block interleave of gst_tensor_aggregator_concat               */
/* (gsttensor_aggregator.c:768-787).  src holds frames_out frames of frame_size bytes     */
/* back to back; block_size = element size * prod (frame dim[0..frames_dim]).             */
/* ------------------------------------------------------------------------------------- */
int
oracle_aggregator_concat (const void *src, void *dst, uint64_t frame_size, uint32_t frames_out,
    uint64_t block_size)
{
  uint64_t src_idx = 0, dest_idx = 0;
  uint32_t f;
  if (block_size == 0 || frame_size == 0 || frame_size % block_size)
    return -1;
  do {
    for (f = 0; f < frames_out; f++) {
      memcpy ((char *) dst + dest_idx, (const char *) src + src_idx + frame_size * f, block_size);
      dest_idx += block_size;
    }
    src_idx += block_size;
  } while (src_idx < frame_size);
  return 0;
}

/* ------------------------------------------------------------------------------------- */
/* tensor_decoder direct_video, dv_decode (tensordec-directvideo.c:322-339): dense rows of  */
/* dim[0]*dim[1] bytes -> rows padded to a multiple of 4.  The reference leaves the padding */
/* bytes uninitialised; dst is expected to be pre-filled by the caller (tests use zero).    */
/* ------------------------------------------------------------------------------------- */
int
oracle_direct_video (const void *src, void *dst, uint32_t color, uint32_t width, uint32_t height)
{
  const size_t row = (size_t) color * width;
  const size_t pitch = ((row - 1) / 4 + 1) * 4;
  uint32_t h;
  if (row % 4 == 0) {
    memcpy (dst, src, row * height);
    return 0;
  }
  for (h = 0; h < height; h++)
    memcpy ((char *) dst + h * pitch, (const char *) src + h * row, row);
  return 0;
}

/* ------------------------------------------------------------------------------------- */
/* tensor_merge mode=linear, gst_tensor_merge_generate_mem (gsttensor_merge.c:506-586).   */
/* dims: n x 4 (only rank 4 takes part, as in the reference); direction 0..3.             */
/* ------------------------------------------------------------------------------------- */
int
oracle_merge_linear (int n, const void *const *in, const uint32_t *dims, uint64_t element_size,
    int direction, void *out)
{
  const uint32_t *dim = dims;           /* outer loops run over the first tensor's dims */
  char *outptr = (char *) out;
  uint32_t i, j, l;
  int k, jj;
  uint64_t c, s;
  switch (direction) {
    case 0:
      for (l = 0; l < dim[3]; l++)
        for (i = 0; i < dim[2]; i++)
          for (j = 0; j < dim[1]; j++)
            for (k = 0; k < n; k++) {
              c = dims[k * 4 + 0];
              s = element_size * c;
              memcpy (outptr, (const char *) in[k] + ((uint64_t) l * dim[2] * dim[1] + (uint64_t) i * dim[1] + j) * s, s);
              outptr += s;
            }
      return 0;
    case 1:
      for (l = 0; l < dim[3]; l++)
        for (i = 0; i < dim[2]; i++)
          for (k = 0; k < n; k++) {
            c = 1;
            for (jj = 0; jj < 2; jj++)
              c *= dims[k * 4 + jj];
            s = element_size * c;
            memcpy (outptr, (const char *) in[k] + ((uint64_t) l * dim[2] + i) * s, s);
            outptr += s;
          }
      return 0;
    case 2:
      for (l = 0; l < dim[3]; l++)
        for (k = 0; k < n; k++) {
          c = 1;
          for (jj = 0; jj < 3; jj++)
            c *= dims[k * 4 + jj];
          s = element_size * c;
          memcpy (outptr, (const char *) in[k] + (uint64_t) l * s, s);
          outptr += s;
        }
      return 0;
    case 3:
      for (k = 0; k < n; k++) {
        c = 1;
        for (jj = 0; jj < 4; jj++)
          c *= dims[k * 4 + jj];
        s = element_size * c;
        memcpy (outptr, in[k], s);
        outptr += s;
      }
      return 0;
    default:
      return -1;
  }
}
