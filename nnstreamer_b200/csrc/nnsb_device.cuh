/*
 * nnsb_device.cuh -- device-side building blocks shared by the sm_100a kernels:
 *   - streaming 4/8/16/32-byte global loads and stores (256-bit LDG/STG exist on sm_100),
 *   - the reference's element conversion rules in both flavours
 *       CCast   : plain C casts as compiled by gcc for x86-64 (tensor_data.c:35-39,202-279)
 *       OrcCast : the ORC programs nns_orc_conv_* (nnstreamer-orc.orc), opcode by opcode
 *   - the arithmetic operators in both flavours (gsttensor_transform.c:563-582 and the
 *     nns_orc_{add,mul,div}_c_* programs).
 * No FMA contraction anywhere: every float operation is an explicit *_rn intrinsic.
 */
#ifndef NNSB_DEVICE_CUH
#define NNSB_DEVICE_CUH

#include <cuda_fp16.h>
#include <stdint.h>

#include "nnsb_internal.h"

namespace nnsb {

/* ------------------------------------------------------------------------------------- */
/* streaming vector access                                                                 */
/* ------------------------------------------------------------------------------------- */
template <int BYTES> struct Bytes;
template <> struct Bytes<1> { uint8_t v; };
template <> struct Bytes<2> { uint16_t v; };
template <> struct Bytes<4> { uint32_t v; };
template <> struct Bytes<8> { uint32_t v[2]; };
template <> struct Bytes<16> { uint32_t v[4]; };
template <> struct Bytes<32> { uint32_t v[8]; };

/* read-once global load: non-coherent path, do not allocate in L1 */
template <int BYTES> __device__ __forceinline__ void
ld_stream (void *dst, const void *src)
{
  if constexpr (BYTES == 32) {
    uint32_t *d = (uint32_t *) dst;
    asm volatile ("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r" (d[0]), "=r" (d[1]), "=r" (d[2]), "=r" (d[3]), "=r" (d[4]), "=r" (d[5]),
          "=r" (d[6]), "=r" (d[7]) : "l" (src));
  } else if constexpr (BYTES == 16) {
    uint32_t *d = (uint32_t *) dst;
    asm volatile ("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
        : "=r" (d[0]), "=r" (d[1]), "=r" (d[2]), "=r" (d[3]) : "l" (src));
  } else if constexpr (BYTES == 8) {
    uint32_t *d = (uint32_t *) dst;
    asm volatile ("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];"
        : "=r" (d[0]), "=r" (d[1]) : "l" (src));
  } else if constexpr (BYTES == 4) {
    uint32_t *d = (uint32_t *) dst;
    asm volatile ("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r" (d[0]) : "l" (src));
  } else if constexpr (BYTES == 2) {
    uint16_t *d = (uint16_t *) dst;
    asm volatile ("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h" (d[0]) : "l" (src));
  } else {
    uint32_t t;
    asm volatile ("ld.global.nc.L1::no_allocate.u8 %0, [%1];" : "=r" (t) : "l" (src));
    *(uint8_t *) dst = (uint8_t) t;
  }
}

/* write-once global store: streaming (evict-first) */
template <int BYTES> __device__ __forceinline__ void
st_stream (void *dst, const void *src)
{
  if constexpr (BYTES == 32) {
    const uint32_t *s = (const uint32_t *) src;
    asm volatile ("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
        :: "l" (dst), "r" (s[0]), "r" (s[1]), "r" (s[2]), "r" (s[3]), "r" (s[4]), "r" (s[5]),
           "r" (s[6]), "r" (s[7]) : "memory");
  } else if constexpr (BYTES == 16) {
    const uint32_t *s = (const uint32_t *) src;
    asm volatile ("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
        :: "l" (dst), "r" (s[0]), "r" (s[1]), "r" (s[2]), "r" (s[3]) : "memory");
  } else if constexpr (BYTES == 8) {
    const uint32_t *s = (const uint32_t *) src;
    asm volatile ("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};"
        :: "l" (dst), "r" (s[0]), "r" (s[1]) : "memory");
  } else if constexpr (BYTES == 4) {
    const uint32_t *s = (const uint32_t *) src;
    asm volatile ("st.global.L1::no_allocate.u32 [%0], %1;" :: "l" (dst), "r" (s[0]) : "memory");
  } else if constexpr (BYTES == 2) {
    const uint16_t *s = (const uint16_t *) src;
    asm volatile ("st.global.L1::no_allocate.u16 [%0], %1;" :: "l" (dst), "h" (s[0]) : "memory");
  } else {
    uint32_t t = *(const uint8_t *) src;
    asm volatile ("st.global.L1::no_allocate.u8 [%0], %1;" :: "l" (dst), "r" (t) : "memory");
  }
}

/* V elements of T as a register array that can be moved with one access */
template <typename T, int V> struct alignas (sizeof (T) * V <= 32 ? sizeof (T) * V : 32) Vec
{
  T e[V];
};

template <typename T, int V> __device__ __forceinline__ void
load_vec (Vec<T, V> &v, const T *p)
{
  constexpr int B = sizeof (T) * V;
  if constexpr (B <= 32) {
    ld_stream<B> (&v, p);
  } else {
#pragma unroll
    for (int i = 0; i < B / 32; i++)
      ld_stream<32> ((char *) &v + 32 * i, (const char *) p + 32 * i);
  }
}

template <typename T, int V> __device__ __forceinline__ void
store_vec (T *p, const Vec<T, V> &v)
{
  constexpr int B = sizeof (T) * V;
  if constexpr (B <= 32) {
    st_stream<B> (p, &v);
  } else {
#pragma unroll
    for (int i = 0; i < B / 32; i++)
      st_stream<32> ((char *) p + 32 * i, (const char *) &v + 32 * i);
  }
}

/* ------------------------------------------------------------------------------------- */
/* type traits                                                                             */
/* ------------------------------------------------------------------------------------- */
template <typename T> struct TypeId;
template <> struct TypeId<int32_t> { static constexpr int id = NNSB_INT32; };
template <> struct TypeId<uint32_t> { static constexpr int id = NNSB_UINT32; };
template <> struct TypeId<int16_t> { static constexpr int id = NNSB_INT16; };
template <> struct TypeId<uint16_t> { static constexpr int id = NNSB_UINT16; };
template <> struct TypeId<int8_t> { static constexpr int id = NNSB_INT8; };
template <> struct TypeId<uint8_t> { static constexpr int id = NNSB_UINT8; };
template <> struct TypeId<double> { static constexpr int id = NNSB_FLOAT64; };
template <> struct TypeId<float> { static constexpr int id = NNSB_FLOAT32; };
template <> struct TypeId<int64_t> { static constexpr int id = NNSB_INT64; };
template <> struct TypeId<uint64_t> { static constexpr int id = NNSB_UINT64; };
template <> struct TypeId<__half> { static constexpr int id = NNSB_FLOAT16; };

template <typename T> struct IsFloat { static constexpr bool v = false; };
template <> struct IsFloat<float> { static constexpr bool v = true; };
template <> struct IsFloat<double> { static constexpr bool v = true; };
template <> struct IsFloat<__half> { static constexpr bool v = true; };

template <typename T> struct IsSigned { static constexpr bool v = (T (-1) < T (0)); };
template <> struct IsSigned<__half> { static constexpr bool v = true; };

template <typename T> struct Is64Int { static constexpr bool v = false; };
template <> struct Is64Int<int64_t> { static constexpr bool v = true; };
template <> struct Is64Int<uint64_t> { static constexpr bool v = true; };

template <int BYTES> struct SignedOf;
template <> struct SignedOf<1> { typedef int8_t type; };
template <> struct SignedOf<2> { typedef int16_t type; };
template <> struct SignedOf<4> { typedef int32_t type; };
template <> struct SignedOf<8> { typedef int64_t type; };

/* ------------------------------------------------------------------------------------- */
/* float -> integer the way x86 does it (cvttss2si / cvttsd2si):                           */
/* truncation; NaN and out-of-range give the "integer indefinite" 0x80..0                  */
/* ------------------------------------------------------------------------------------- */
__device__ __forceinline__ int32_t x86_cvtt32 (float f)
{
  /* cvt.rzi saturates: negative overflow already yields INT_MIN; positive overflow and
   * NaN (comparison false) must be turned into INT_MIN too */
  return (f < 2147483648.0f) ? __float2int_rz (f) : (int32_t) 0x80000000;
}
__device__ __forceinline__ int32_t x86_cvtt32 (double d)
{
  return (d < 2147483648.0) ? __double2int_rz (d) : (int32_t) 0x80000000;
}
__device__ __forceinline__ int64_t x86_cvtt64 (float f)
{
  return (f < 9223372036854775808.0f) ? __float2ll_rz (f) : (int64_t) 0x8000000000000000LL;
}
__device__ __forceinline__ int64_t x86_cvtt64 (double d)
{
  return (d < 9223372036854775808.0) ? __double2ll_rz (d) : (int64_t) 0x8000000000000000LL;
}

/* exact u8 -> f32: place the byte in the mantissa of 2^23 and subtract 2^23 */
__device__ __forceinline__ float u8_to_f32 (uint32_t byte)
{
  return __fsub_rn (__uint_as_float (0x4B000000u | byte), 8388608.0f);
}

/* float <-> float conversions the way x86 does them (cvtss2sd, cvtsd2ss, vcvtph2ps,
 * vcvtps2ph / libgcc): round to nearest even; a NaN keeps its sign and the top payload bits
 * and becomes quiet.  CUDA's cvt returns a canonical NaN instead, hence the explicit paths. */
__device__ __forceinline__ float f16_to_f32 (__half h)
{
  const unsigned short b = __half_as_ushort (h);
  if ((b & 0x7fffu) > 0x7c00u)
    return __uint_as_float (((uint32_t) (b & 0x8000u) << 16) | 0x7fc00000u | ((uint32_t) (b & 0x3ffu) << 13));
  return __half2float (h);
}
__device__ __forceinline__ __half f32_to_f16 (float f)
{
  if (f != f) {
    const uint32_t u = __float_as_uint (f);
    return __ushort_as_half ((unsigned short) (((u >> 16) & 0x8000u) | 0x7e00u | ((u >> 13) & 0x1ffu)));
  }
  return __float2half_rn (f);
}
__device__ __forceinline__ __half f64_to_f16 (double d)
{
  if (d != d) {
    const unsigned long long u = (unsigned long long) __double_as_longlong (d);
    return __ushort_as_half ((unsigned short) (((u >> 48) & 0x8000u) | 0x7e00u | ((u >> 42) & 0x1ffu)));
  }
  return __double2half (d);
}
__device__ __forceinline__ double f32_to_f64 (float f)
{
  if (f != f) {
    const unsigned long long u = __float_as_uint (f);
    return __longlong_as_double ((long long) (((u & 0x80000000ull) << 32) | 0x7ff8000000000000ull
            | ((u & 0x7fffffull) << 29)));
  }
  return (double) f;
}
__device__ __forceinline__ float f64_to_f32 (double d)
{
  if (d != d) {
    const unsigned long long u = (unsigned long long) __double_as_longlong (d);
    return __uint_as_float ((uint32_t) ((u >> 32) & 0x80000000u) | 0x7fc00000u
        | (uint32_t) ((u >> 29) & 0x3fffffu));
  }
  return __double2float_rn (d);
}

/* ------------------------------------------------------------------------------------- */
/* CCast: `(To) from` as the reference's C path performs it                                */
/* gst_tensor_data_typecast, tensor_data.c:202-279: float sources reach an unsigned target */
/* through the signed type of the same width (:219-222,229-232,239-242,264-267)            */
/* ------------------------------------------------------------------------------------- */
template <typename To, typename From> struct CCast
{
  static __device__ __forceinline__ To run (From x)
  {
    if constexpr (IsFloat<From>::v && !IsFloat<To>::v) {
      /* float/double/half -> integer */
      typedef typename SignedOf<sizeof (To)>::type S;
      S s;
      if constexpr (sizeof (To) == 8) {
        if constexpr (sizeof (From) == 8) s = (S) x86_cvtt64 ((double) x);
        else if constexpr (sizeof (From) == 4) s = (S) x86_cvtt64 ((float) x);
        else s = (S) x86_cvtt64 (__half2float (x));
      } else {
        int32_t t;
        if constexpr (sizeof (From) == 8) t = x86_cvtt32 ((double) x);
        else if constexpr (sizeof (From) == 4) t = x86_cvtt32 ((float) x);
        else t = x86_cvtt32 (__half2float (x));
        s = (S) t;              /* low bits, like the movsx / byte store after cvtt */
      }
      return (To) s;
    } else if constexpr (IsFloat<From>::v && IsFloat<To>::v) {
      if constexpr (sizeof (From) == sizeof (To)) return x;
      else if constexpr (sizeof (From) == 2 && sizeof (To) == 4) return f16_to_f32 (x);
      else if constexpr (sizeof (From) == 2 && sizeof (To) == 8) return f32_to_f64 (f16_to_f32 (x));
      else if constexpr (sizeof (To) == 2 && sizeof (From) == 4) return f32_to_f16 (x);
      else if constexpr (sizeof (To) == 2 && sizeof (From) == 8) return f64_to_f16 (x);
      else if constexpr (sizeof (To) == 4 && sizeof (From) == 8) return f64_to_f32 (x);
      else return f32_to_f64 (x);
    } else if constexpr (!IsFloat<From>::v && IsFloat<To>::v) {
      /* integer -> float: round to nearest even, one rounding */
      if constexpr (sizeof (To) == 2) {
        if constexpr (sizeof (From) == 8) {
          if constexpr (IsSigned<From>::v) return __ll2half_rn ((long long) x);
          else return __ull2half_rn ((unsigned long long) x);
        } else if constexpr (IsSigned<From>::v) return __int2half_rn ((int) x);
        else return __uint2half_rn ((unsigned int) x);
      } else if constexpr (sizeof (To) == 4) {
        if constexpr (sizeof (From) == 1 && !IsSigned<From>::v) return u8_to_f32 (x);
        else if constexpr (sizeof (From) == 8) {
          if constexpr (IsSigned<From>::v) return __ll2float_rn ((long long) x);
          else return __ull2float_rn ((unsigned long long) x);
        } else if constexpr (IsSigned<From>::v) return __int2float_rn ((int) x);
        else return __uint2float_rn ((unsigned int) x);
      } else {
        if constexpr (sizeof (From) == 8) {
          if constexpr (IsSigned<From>::v) return __ll2double_rn ((long long) x);
          else return __ull2double_rn ((unsigned long long) x);
        } else if constexpr (IsSigned<From>::v) return __int2double_rn ((int) x);
        else return __uint2double_rn ((unsigned int) x);
      }
    } else {
      return (To) x;            /* integer -> integer: modular */
    }
  }
};

/* ------------------------------------------------------------------------------------- */
/* OrcCast: the nns_orc_conv_<A>_to_<B> programs (nnstreamer-orc.orc).  Only the eight     */
/* types ORC handles; 64-bit integers and float16 always take the C path                   */
/* (orc_supported, gsttensor_transform.c:462-463; _conv_*_f16, :366-428).                  */
/* ------------------------------------------------------------------------------------- */
__device__ __forceinline__ int16_t orc_convssslw (int32_t x)  /* signed saturate 32 -> 16 */
{
  return (int16_t) max (-32768, min (32767, x));
}
__device__ __forceinline__ int8_t orc_convssswb (int16_t x)   /* signed saturate 16 -> 8 */
{
  return (int8_t) max (-128, min (127, (int) x));
}
/* convfl / convdl: C cast, then positive overflow (and positive NaN) -> 0x7fffffff */
__device__ __forceinline__ int32_t orc_convfl (float f)
{
  int32_t t = x86_cvtt32 (f);
  if (t == (int32_t) 0x80000000 && !(__float_as_uint (f) & 0x80000000u)) t = 0x7fffffff;
  return t;
}
__device__ __forceinline__ int32_t orc_convdl (double d)
{
  int32_t t = x86_cvtt32 (d);
  if (t == (int32_t) 0x80000000 && !((unsigned long long) __double_as_longlong (d) >> 63))
    t = 0x7fffffff;
  return t;
}

template <typename To, typename From> struct OrcCast
{
  static __device__ __forceinline__ To run (From x)
  {
    if constexpr (IsFloat<From>::v) {
      if constexpr (IsFloat<To>::v) {
        return CCast<To, From>::run (x);                /* copyl / convfd / convdf / copyq */
      } else {
        int32_t l;
        if constexpr (sizeof (From) == 4) l = orc_convfl (x); else l = orc_convdl (x);
        if constexpr (sizeof (To) == 4) return (To) l;                       /* convfl */
        else if constexpr (sizeof (To) == 2) return (To) orc_convssslw (l);   /* s16 and u16 */
        else if constexpr (IsSigned<To>::v) return (To) orc_convssswb (orc_convssslw (l));
        else return (To) (uint8_t) l;                                        /* convlw, convwb */
      }
    } else if constexpr (IsFloat<To>::v) {
      if constexpr (sizeof (From) == 4) {
        /* convlf / convld read the 32-bit source as SIGNED, also for uint32 */
        return CCast<To, int32_t>::run ((int32_t) x);
      } else {
        return CCast<To, From>::run (x);                /* widen (sign/zero), then convlf/ld */
      }
    } else {
      /* integer -> integer */
      if constexpr (sizeof (To) >= sizeof (From)) {
        return (To) x;                                  /* copy / conv[su]bw / conv[su]wl */
      } else if constexpr (sizeof (From) == 4 && sizeof (To) == 2) {
        return (To) orc_convssslw ((int32_t) x);        /* s16 and u16 targets, src as signed */
      } else if constexpr (sizeof (From) == 4 && sizeof (To) == 1) {
        if constexpr (IsSigned<To>::v) return (To) orc_convssswb (orc_convssslw ((int32_t) x));
        else return (To) (uint8_t) (uint32_t) x;        /* convlw, convwb: truncate */
      } else {
        /* 16 -> 8 */
        if constexpr (IsSigned<To>::v) return (To) orc_convssswb ((int16_t) x);
        else return (To) (uint8_t) (uint16_t) x;        /* convwb: truncate */
      }
    }
  }
};

template <typename To, typename From, bool ORC> struct Cast
{
  static __device__ __forceinline__ To run (From x)
  {
    if constexpr (ORC) return OrcCast<To, From>::run (x);
    else return CCast<To, From>::run (x);
  }
};

/* ------------------------------------------------------------------------------------- */
/* operators                                                                               */
/* ------------------------------------------------------------------------------------- */
template <typename T> __device__ __forceinline__ T cell_get (const nnsb_cell &c)
{
  T v;
  memcpy (&v, &c, sizeof (T));
  return v;
}

/* x86 SSE NaN rules for a binary operation r = a op b: a NaN operand is returned quieted
 * (the first operand wins), an invalid operation (inf-inf, 0*inf, 0/0, inf/inf) returns the
 * default NaN, which has the SIGN BIT SET.  CUDA returns canonical 0x7fffffff instead. */
__device__ __forceinline__ float nan_fix (float r, float a, float b)
{
  if (r == r) return r;
  if (a != a) return __uint_as_float (__float_as_uint (a) | 0x00400000u);
  if (b != b) return __uint_as_float (__float_as_uint (b) | 0x00400000u);
  return __uint_as_float (0xffc00000u);
}
__device__ __forceinline__ double nan_fix (double r, double a, double b)
{
  if (r == r) return r;
  if (a != a) return __longlong_as_double (__double_as_longlong (a) | 0x0008000000000000ll);
  if (b != b) return __longlong_as_double (__double_as_longlong (b) | 0x0008000000000000ll);
  return __longlong_as_double ((long long) 0xfff8000000000000ull);
}

/* C path: handle_operator, gsttensor_transform.c:563-582 (+=, *=, /= in the working type;
 * narrow integers are promoted to int and truncated back, i.e. modular) */
template <typename T, bool ORC> struct Arith
{
  static __device__ __forceinline__ T add (T a, T b)
  {
    if constexpr (sizeof (T) == 2 && IsFloat<T>::v) {
      const float fa = f16_to_f32 (a), fb = f16_to_f32 (b);
      return f32_to_f16 (nan_fix (__fadd_rn (fa, fb), fa, fb));
    } else if constexpr (sizeof (T) == 4 && IsFloat<T>::v) {
      return nan_fix (__fadd_rn (a, b), a, b);
    } else if constexpr (sizeof (T) == 8 && IsFloat<T>::v) {
      return nan_fix (__dadd_rn (a, b), a, b);
    } else if constexpr (ORC) {
      /* addssb/addusb/addssw/addusw/addssl/addusl: saturating */
      if constexpr (sizeof (T) == 4 && IsSigned<T>::v) {
        long long s = (long long) a + (long long) b;
        return (T) max (-2147483648LL, min (2147483647LL, s));
      } else if constexpr (sizeof (T) == 4) {
        unsigned int s = (unsigned int) a + (unsigned int) b;
        return (T) (s < (unsigned int) a ? 0xffffffffu : s);
      } else if constexpr (IsSigned<T>::v) {
        int s = (int) a + (int) b;
        constexpr int lo = sizeof (T) == 1 ? -128 : -32768, hi = sizeof (T) == 1 ? 127 : 32767;
        return (T) max (lo, min (hi, s));
      } else {
        int s = (int) a + (int) b;
        constexpr int hi = sizeof (T) == 1 ? 255 : 65535;
        return (T) min (hi, s);
      }
    } else {
      typedef typename SignedOf<sizeof (T)>::type S;
      if constexpr (sizeof (T) == 8) return (T) ((unsigned long long) a + (unsigned long long) b);
      else { (void) sizeof (S); return (T) ((unsigned int) a + (unsigned int) b); }
    }
  }

  static __device__ __forceinline__ T mul (T a, T b)
  {
    if constexpr (sizeof (T) == 2 && IsFloat<T>::v) {
      const float fa = f16_to_f32 (a), fb = f16_to_f32 (b);
      return f32_to_f16 (nan_fix (__fmul_rn (fa, fb), fa, fb));
    } else if constexpr (sizeof (T) == 4 && IsFloat<T>::v) {
      return nan_fix (__fmul_rn (a, b), a, b);
    } else if constexpr (sizeof (T) == 8 && IsFloat<T>::v) {
      return nan_fix (__dmul_rn (a, b), a, b);
    } else if constexpr (ORC) {
      /* mul[su]bw+conv, mul[su]wl+conv, mul[su]lq+conv: widening multiply, saturating narrow */
      if constexpr (sizeof (T) == 4 && IsSigned<T>::v) {
        long long p = (long long) a * (long long) b;
        return (T) max (-2147483648LL, min (2147483647LL, p));
      } else if constexpr (sizeof (T) == 4) {
        unsigned long long p = (unsigned long long) a * (unsigned long long) b;
        return (T) (p > 0xffffffffULL ? 0xffffffffULL : p);
      } else if constexpr (IsSigned<T>::v) {
        int p = (int) a * (int) b;
        constexpr int lo = sizeof (T) == 1 ? -128 : -32768, hi = sizeof (T) == 1 ? 127 : 32767;
        return (T) max (lo, min (hi, p));
      } else {
        unsigned int p = (unsigned int) a * (unsigned int) b;
        constexpr unsigned int hi = sizeof (T) == 1 ? 255u : 65535u;
        return (T) min (hi, p);
      }
    } else {
      if constexpr (sizeof (T) == 8) return (T) ((unsigned long long) a * (unsigned long long) b);
      else return (T) ((unsigned int) a * (unsigned int) b);
    }
  }

  /* the divisor is never zero here for the C path (filtered when the chain is compiled) */
  static __device__ __forceinline__ T div (T a, T b)
  {
    if constexpr (sizeof (T) == 2 && IsFloat<T>::v) {
      const float fa = f16_to_f32 (a), fb = f16_to_f32 (b);
      return f32_to_f16 (nan_fix (__fdiv_rn (fa, fb), fa, fb));
    } else if constexpr (sizeof (T) == 4 && IsFloat<T>::v) {
      return nan_fix (__fdiv_rn (a, b), a, b);
    } else if constexpr (sizeof (T) == 8 && IsFloat<T>::v) {
      return nan_fix (__ddiv_rn (a, b), a, b);
    } else if constexpr (IsSigned<T>::v) {
      /* C division truncates toward zero; INT_MIN / -1 (a trap on x86) is defined here as
       * the wrapped value so the kernel cannot fault */
      if constexpr (sizeof (T) == 8) {
        if (b == T (-1)) return (T) (0ULL - (unsigned long long) a);
        return (T) ((long long) a / (long long) b);
      } else {
        if (b == T (-1)) return (T) (0u - (unsigned int) (int) a);
        return (T) ((int) a / (int) b);
      }
    } else {
      if constexpr (sizeof (T) == 8) return (T) ((unsigned long long) a / (unsigned long long) b);
      else return (T) ((unsigned int) a / (unsigned int) b);
    }
  }
};

}  // namespace nnsb
#endif
