/*
 * nnsb_plan_host.cpp -- host-side compilation of a parsed arithmetic option into the dense
 * "chain" the CUDA kernel executes (compiled with g++, not nvcc: needs _Float16).
 *
 * Reference behaviour being compiled (gsttensor_transform.c):
 *   - each operand literal is int64 or double (:796-807) and is C-cast ONCE to the working
 *     type by gst_tensor_data_typecast (&op_s->value, type) (:1349,:1394, ORC path :1268);
 *   - operators run in list order, each rounding to the working type (:563-582);
 *   - per-channel: operator k touches an element iff applying_ch is -1 or equals the
 *     element's channel (:1351);
 *   - C path: a zero divisor is logged and the element left unchanged (:572-575, the return
 *     value is ignored at :1352,:1395).
 *
 * Compilation: walk the list once and merge consecutive per-channel operators with the same
 * opcode on different channels into one POSITION holding one operand per channel (identity
 * where nothing applies: add -0.0 / 0, mul 1, div 1 -- all exact).  "add@0,add@1,add@2,
 * div@0,div@1,div@2" becomes 2 positions instead of 6 predicated passes.
 */
#include "nnsb_internal.h"

#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

/* x86 cvttsd2si with a 32-bit destination: out-of-range and NaN give the
 * "integer indefinite" 0x80000000 -- what the reference's (int32_t) cast compiles to. */
static inline int32_t
x86_cvtt32 (double d)
{
  if (d > -2147483649.0 && d < 2147483648.0)
    return (int32_t) d;
  return INT32_MIN;
}

/* x86 cvttsd2si with a 64-bit destination */
static inline int64_t
x86_cvtt64 (double d)
{
  if (d >= -9223372036854775808.0 && d < 9223372036854775808.0)
    return (int64_t) d;
  return INT64_MIN;
}

static inline uint16_t
f16_bits (_Float16 h)
{
  uint16_t b;
  memcpy (&b, &h, sizeof (b));
  return b;
}

extern "C" nnsb_cell
nnsb_cast_operand (const nnsb_arith_op * op, int type)
{
  nnsb_cell c;
  c.raw = 0;
  if (op->is_float) {
    const double d = op->dval;
    switch (type) {
      case NNSB_INT32: c.i32 = x86_cvtt32 (d); break;
      case NNSB_UINT32: c.u32 = (uint32_t) x86_cvtt32 (d); break;       /* via int32, tensor_data.c:219-222 */
      case NNSB_INT16: c.i16 = (int16_t) x86_cvtt32 (d); break;
      case NNSB_UINT16: c.u16 = (uint16_t) (int16_t) x86_cvtt32 (d); break;   /* :229-232 */
      case NNSB_INT8: c.i8 = (int8_t) x86_cvtt32 (d); break;
      case NNSB_UINT8: c.u8 = (uint8_t) (int8_t) x86_cvtt32 (d); break;       /* :239-242 */
      case NNSB_FLOAT64: c.f64 = d; break;
      case NNSB_FLOAT32: c.f32 = (float) d; break;
      case NNSB_INT64: c.i64 = x86_cvtt64 (d); break;
      case NNSB_UINT64: c.u64 = (uint64_t) x86_cvtt64 (d); break;       /* :264-267 */
      case NNSB_FLOAT16: c.f16_bits = f16_bits ((_Float16) d); break;
      default: break;
    }
  } else {
    const int64_t v = op->ival;
    switch (type) {
      case NNSB_INT32: c.i32 = (int32_t) v; break;
      case NNSB_UINT32: c.u32 = (uint32_t) v; break;
      case NNSB_INT16: c.i16 = (int16_t) v; break;
      case NNSB_UINT16: c.u16 = (uint16_t) v; break;
      case NNSB_INT8: c.i8 = (int8_t) v; break;
      case NNSB_UINT8: c.u8 = (uint8_t) v; break;
      case NNSB_FLOAT64: c.f64 = (double) v; break;
      case NNSB_FLOAT32: c.f32 = (float) v; break;
      case NNSB_INT64: c.i64 = v; break;
      case NNSB_UINT64: c.u64 = (uint64_t) v; break;
      case NNSB_FLOAT16: c.f16_bits = f16_bits ((_Float16) v); break;
      default: break;
    }
  }
  return c;
}

static bool
cell_is_zero (const nnsb_cell & c, int type)
{
  switch (type) {
    case NNSB_INT32: case NNSB_UINT32: return c.u32 == 0;
    case NNSB_INT16: case NNSB_UINT16: return c.u16 == 0;
    case NNSB_INT8: case NNSB_UINT8: return c.u8 == 0;
    case NNSB_FLOAT64: return c.f64 == 0.0;
    case NNSB_FLOAT32: return c.f32 == 0.0f;
    case NNSB_INT64: case NNSB_UINT64: return c.u64 == 0;
    case NNSB_FLOAT16: return (c.f16_bits & 0x7fff) == 0;
    default: return true;
  }
}

static bool
type_is_float (int type)
{
  return type == NNSB_FLOAT64 || type == NNSB_FLOAT32 || type == NNSB_FLOAT16;
}

static nnsb_cell
identity_cell (int opcode, int type)
{
  nnsb_arith_op lit;
  memset (&lit, 0, sizeof (lit));
  if (opcode == NNSB_OP_ADD) {
    if (type_is_float (type)) {
      lit.is_float = 1;
      lit.dval = -0.0;          /* x + (-0.0) == x for every x, including -0.0 */
    } else {
      lit.ival = 0;
    }
  } else {
    lit.ival = 1;               /* x * 1 == x, x / 1 == x */
  }
  return nnsb_cast_operand (&lit, type);
}

struct Position
{
  int opcode;
  bool partial;                 /* created by a per-channel operator: unset slots may be filled */
  std::vector < nnsb_cell > val;
  std::vector < char >set;
};

extern "C" int
nnsb_chain_compile (const nnsb_option * opt, int work_type, int acceleration,
    const uint32_t * dim, nnsb_chain_desc * desc, nnsb_cell ** table)
{
  memset (desc, 0, sizeof (*desc));
  *table = nullptr;
  if (work_type < 0 || work_type >= NNSB_TYPE_END)
    return nnsb_set_error (NNSB_EINVAL, "arithmetic: invalid working type %d", work_type);

  uint32_t nch = 1;
  uint64_t ch_size = 1;
  const bool per_channel = opt->arith_per_channel != 0;
  if (per_channel) {
    if (opt->arith_ch_dim >= NNSB_RANK_LIMIT)
      return nnsb_set_error (NNSB_EINVAL,
          "arithmetic: per-channel dimension index %u exceeds the rank limit %d",
          opt->arith_ch_dim, NNSB_RANK_LIMIT);
    /* ch_size / ch_offset, gsttensor_transform.c:1317-1322 */
    for (uint32_t i = 0; i < opt->arith_ch_dim; i++)
      ch_size *= dim[i];
    nch = dim[opt->arith_ch_dim];
    if (nch == 0)
      return nnsb_set_error (NNSB_EINVAL, "arithmetic: channel dimension is 0");
  }

  std::vector < Position > pos;
  for (int k = 0; k < opt->n_ops; k++) {
    const nnsb_arith_op & op = opt->ops[k];
    if (op.op == NNSB_OP_TYPECAST)
      continue;                 /* handled as the load-side conversion */
    if (op.op != NNSB_OP_ADD && op.op != NNSB_OP_MUL && op.op != NNSB_OP_DIV)
      continue;
    nnsb_cell v = nnsb_cast_operand (&op, work_type);
    if (op.op == NNSB_OP_DIV && cell_is_zero (v, work_type)) {
      /* C path: element unchanged.  ORC path: divf/divd (and the float16 loop) really divide
       * -> inf/NaN; the ORC integer loop would raise SIGFPE, which we do not reproduce. */
      if (!(acceleration && type_is_float (work_type)))
        continue;
    }
    if (!per_channel || op.applying_ch == -1) {
      Position p;
      p.opcode = op.op;
      p.partial = false;
      p.val.assign (nch, v);
      p.set.assign (nch, 1);
      pos.push_back (p);
    } else {
      if (op.applying_ch < 0 || (uint32_t) op.applying_ch >= nch)
        continue;               /* never equals a channel index (:1351) */
      const uint32_t c = (uint32_t) op.applying_ch;
      if (!pos.empty () && pos.back ().partial && pos.back ().opcode == op.op
          && !pos.back ().set[c]) {
        pos.back ().val[c] = v;
        pos.back ().set[c] = 1;
      } else {
        Position p;
        p.opcode = op.op;
        p.partial = true;
        p.val.assign (nch, identity_cell (op.op, work_type));
        p.set.assign (nch, 0);
        p.val[c] = v;
        p.set[c] = 1;
        pos.push_back (p);
      }
    }
  }
  if (pos.size () > NNSB_MAX_OPS)
    return nnsb_set_error (NNSB_EINVAL, "arithmetic: chain too long");

  if (work_type == NNSB_FLOAT32) {
    /* float32 chains: the upper half of every divisor cell carries RN (1 / d) for the
     * division-by-constant of the fast kernel (k_chain.cuh, fchain_kernel); 0 = not eligible
     * (then the launch takes the generic kernel).  Other cells keep a zero upper half. */
    for (auto & ps : pos)
      for (auto & cell : ps.val) {
        const uint32_t lo = (uint32_t) (cell.raw & 0xffffffffu);
        uint32_t hi = 0;
        if (ps.opcode == NNSB_OP_DIV) {
          float d;
          memcpy (&d, &lo, sizeof (d));
          const uint32_t expo = (lo >> 23) & 0xffu, frac = lo & 0x7fffffu;
          /* normal, 2^-30 <= |d| <= 2^30, significand not all ones (Markstein's exception) */
          if (expo >= 97 && expo <= 157 && frac != 0x7fffffu) {
            const float r = 1.0f / d;   /* correctly rounded (no contraction, SSE) */
            memcpy (&hi, &r, sizeof (hi));
          }
        }
        cell.raw = ((uint64_t) hi << 32) | lo;
      }
  }

  desc->n_pos = (int32_t) pos.size ();
  desc->nch = (int32_t) nch;
  desc->ch_size = ch_size;
  for (size_t i = 0; i < pos.size (); i++) {
    desc->opcode[i] = pos[i].opcode;
    desc->scalar[i] = pos[i].val[0];
  }
  if (nch > 1 && !pos.empty ()) {
    nnsb_cell *t = (nnsb_cell *) malloc (sizeof (nnsb_cell) * pos.size () * nch);
    if (!t)
      return nnsb_set_error (NNSB_ENOMEM, "arithmetic: out of host memory");
    for (size_t i = 0; i < pos.size (); i++)
      for (uint32_t c = 0; c < nch; c++)
        t[i * nch + c] = pos[i].val[c];
    *table = t;
  }
  return NNSB_OK;
}
