/*
 * k_chain.cuh -- K1/K2: fused typecast + arithmetic chain, one pass over HBM.
 *
 * Replaces gst_tensor_transform_typecast (gsttensor_transform.c:1202-1229) and
 * gst_tensor_transform_arithmetic (:1240-1412): the reference's ORC path makes one full pass
 * over the output per operator, its C path walks a GSList through a typed union per element.
 * Here every element is loaded once (In), converted (C or ORC rules), run through the
 * compiled chain in the working type Out with per-operation rounding (no FMA contraction,
 * IEEE division) and stored once.
 *
 * Data movement: each thread owns V consecutive elements per step where the wider of the
 * two element types fills a 256-bit access (u8->f32: one 8-byte load, one 32-byte store);
 * consecutive threads own consecutive vectors, so every warp access is a dense 256/512/1024
 * byte run.  U independent loads are issued before the first use.  Loads are
 * ld.global.nc.L1::no_allocate, stores st.global.L1::no_allocate (read/write-once streams).
 *
 * Algorithmic bytes per element: sizeof(In) + sizeof(Out)  (5 for u8->f32).
 */
#ifndef NNSB_K_CHAIN_CUH
#define NNSB_K_CHAIN_CUH

#include "nnsb_device.cuh"
#include "nnsb_kernels.h"

namespace nnsb {

struct ChainDev
{
  int32_t n_pos;
  int32_t nch;                  /* 1 = not per-channel */
  uint64_t ch_size;
  uint64_t step_r;              /* (grid threads * V) % ch_size */
  uint32_t step_c;              /* ((grid threads * V) / ch_size) % nch */
  const nnsb_cell *table;       /* [n_pos][nch] when nch > 1 */
  int32_t opcode[NNSB_MAX_OPS];
  nnsb_cell scalar[NNSB_MAX_OPS];
};

constexpr int CHAIN_THREADS = 256;

template <typename In, typename Out> struct ChainVec
{
  static constexpr int W = sizeof (In) > sizeof (Out) ? sizeof (In) : sizeof (Out);
  /* at most 16 elements per thread keeps the 1-byte pairs under ~100 registers */
  static constexpr int V = (32 / W) > 16 ? 16 : (32 / W);
};

template <typename Out> __device__ __forceinline__ Out
chain_operand (const ChainDev &p, int pos, uint32_t ch, bool perch)
{
  if (perch) {
    unsigned long long raw = __ldg ((const unsigned long long *) (p.table + (size_t) pos * p.nch + ch));
    Out v;
    memcpy (&v, &raw, sizeof (Out));
    return v;
  }
  return cell_get<Out> (p.scalar[pos]);
}

template <typename Out, bool ORC, int V> __device__ __forceinline__ void
chain_apply (Vec<Out, V> &y, const uint32_t (&ch)[V], const ChainDev &p, bool perch)
{
  for (int pos = 0; pos < p.n_pos; pos++) {
    const int opc = p.opcode[pos];
    if (opc == NNSB_OP_ADD) {
#pragma unroll
      for (int k = 0; k < V; k++)
        y.e[k] = Arith<Out, ORC>::add (y.e[k], chain_operand<Out> (p, pos, ch[k], perch));
    } else if (opc == NNSB_OP_MUL) {
#pragma unroll
      for (int k = 0; k < V; k++)
        y.e[k] = Arith<Out, ORC>::mul (y.e[k], chain_operand<Out> (p, pos, ch[k], perch));
    } else {
#pragma unroll
      for (int k = 0; k < V; k++)
        y.e[k] = Arith<Out, ORC>::div (y.e[k], chain_operand<Out> (p, pos, ch[k], perch));
    }
  }
}

#ifndef NNSB_CHAIN_MIN_BLOCKS
#define NNSB_CHAIN_MIN_BLOCKS 3
#endif
template <typename In, typename Out, bool ORC, int U>
__global__ void __launch_bounds__ (CHAIN_THREADS, NNSB_CHAIN_MIN_BLOCKS)
chain_kernel (const In *__restrict__ in, Out *__restrict__ out, size_t n, size_t n_vec,
    const __grid_constant__ ChainDev p)
{
  constexpr int V = ChainVec<In, Out>::V;
  const size_t tid = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
  const size_t nthreads = (size_t) gridDim.x * blockDim.x;
  const bool perch = p.nch > 1;

  /* channel walker: (r, c) = position inside the channel run and channel index of the
   * first element of this thread's current vector */
  uint64_t r = 0;
  uint32_t c = 0;
  if (perch) {
    const uint64_t e0 = (uint64_t) tid * V;
    r = e0 % p.ch_size;
    c = (uint32_t) ((e0 / p.ch_size) % (uint64_t) p.nch);
  }

  for (size_t v = tid; v < n_vec; v += nthreads * U) {
    Vec<In, V> x[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t vu = v + (size_t) u * nthreads;
      if (vu < n_vec)
        load_vec (x[u], in + vu * V);
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t vu = v + (size_t) u * nthreads;
      if (vu < n_vec) {
        Vec<Out, V> y;
        uint32_t ch[V];
#pragma unroll
        for (int k = 0; k < V; k++)
          y.e[k] = Cast<Out, In, ORC>::run (x[u].e[k]);
        if (perch) {
          uint64_t rr = r;
          uint32_t cc = c;
#pragma unroll
          for (int k = 0; k < V; k++) {
            ch[k] = cc;
            if (++rr == p.ch_size) {
              rr = 0;
              cc = (cc + 1 == (uint32_t) p.nch) ? 0 : cc + 1;
            }
          }
        } else {
#pragma unroll
          for (int k = 0; k < V; k++)
            ch[k] = 0;
        }
        chain_apply<Out, ORC, V> (y, ch, p, perch);
        store_vec (out + vu * V, y);
      }
      if (perch) {
        /* advance the walker by one grid stride (nthreads * V elements) */
        r += p.step_r;
        uint32_t carry = 0;
        if (r >= p.ch_size) {
          r -= p.ch_size;
          carry = 1;
        }
        c += p.step_c + carry;
        if (c >= (uint32_t) p.nch)
          c -= (uint32_t) p.nch;
      }
    }
  }

  /* scalar tail (and the whole tensor when the pointers are not vector-aligned) */
  for (size_t e = n_vec * V + tid; e < n; e += nthreads) {
    Vec<Out, 1> y;
    uint32_t ch[1] = { 0 };
    In xv;
    ld_stream<sizeof (In)> (&xv, in + e);
    y.e[0] = Cast<Out, In, ORC>::run (xv);
    if (perch)
      ch[0] = (uint32_t) (((uint64_t) e / p.ch_size) % (uint64_t) p.nch);
    chain_apply<Out, ORC, 1> (y, ch, p, perch);
    st_stream<sizeof (Out)> (out + e, &y.e[0]);
  }
}

/* ------------------------------------------------------------------------------------- */
/* 1-byte inputs: the whole chain is a function of (channel, byte) -> 256 entries per        */
/* channel.  Each block evaluates the table once with the SAME code path as the generic      */
/* kernel (so results are identical by construction), then streams: one shared-memory        */
/* lookup per element instead of a convert + N rounded operations (IEEE div is ~10 issue     */
/* slots; at 5 B/element the generic u8->f32 kernel is issue-bound, not HBM-bound).           */
/* ------------------------------------------------------------------------------------- */
template <typename Out> struct LutVec
{
  static constexpr int V = (32 / sizeof (Out)) > 16 ? 16 : (32 / sizeof (Out));
};

template <typename In, typename Out, bool INVARIANT, int U>
__device__ __forceinline__ void
lut_stream (const In *__restrict__ in, Out *__restrict__ out, size_t n_vec, const Out *lut,
    const ChainDev &p, size_t tid, size_t nthreads, bool perch)
{
  constexpr int V = LutVec<Out>::V;
  uint64_t r = 0;
  uint32_t c = 0;
  if (perch) {
    const uint64_t e0 = (uint64_t) tid * V;
    r = e0 % p.ch_size;
    c = (uint32_t) ((e0 / p.ch_size) % (uint64_t) p.nch);
  }
  /* per-lane table rows; loop-invariant when the grid stride is a multiple of the channel
   * period (the launcher sizes the grid for that whenever it can) */
  const Out *row[V];
#pragma unroll
  for (int k = 0; k < V; k++)
    row[k] = lut;
  if (INVARIANT && perch) {
    uint64_t rr = r;
    uint32_t cc = c;
#pragma unroll
    for (int k = 0; k < V; k++) {
      row[k] = lut + (size_t) cc * 256;
      if (++rr == p.ch_size) {
        rr = 0;
        cc = (cc + 1 == (uint32_t) p.nch) ? 0 : cc + 1;
      }
    }
  }
  for (size_t v = tid; v < n_vec; v += nthreads * U) {
    Vec<In, V> x[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t vu = v + (size_t) u * nthreads;
      if (vu < n_vec)
        load_vec (x[u], in + vu * V);
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t vu = v + (size_t) u * nthreads;
      if (!INVARIANT && perch) {
        uint64_t rr = r;
        uint32_t cc = c;
#pragma unroll
        for (int k = 0; k < V; k++) {
          row[k] = lut + (size_t) cc * 256;
          if (++rr == p.ch_size) {
            rr = 0;
            cc = (cc + 1 == (uint32_t) p.nch) ? 0 : cc + 1;
          }
        }
        r += p.step_r;
        uint32_t carry = 0;
        if (r >= p.ch_size) {
          r -= p.ch_size;
          carry = 1;
        }
        c += p.step_c + carry;
        if (c >= (uint32_t) p.nch)
          c -= (uint32_t) p.nch;
      }
      if (vu < n_vec) {
        Vec<Out, V> y;
#pragma unroll
        for (int k = 0; k < V; k++)
          y.e[k] = row[k][(uint8_t) x[u].e[k]];
        store_vec (out + vu * V, y);
      }
    }
  }
}

template <typename In, typename Out, bool ORC, int U>
__global__ void __launch_bounds__ (CHAIN_THREADS)
lut_chain_kernel (const In *__restrict__ in, Out *__restrict__ out, size_t n, size_t n_vec,
    const __grid_constant__ ChainDev p)
{
  static_assert (sizeof (In) == 1, "table lookup needs a 1-byte input type");
  constexpr int V = LutVec<Out>::V;
  extern __shared__ __align__ (16) unsigned char lut_raw[];
  Out *lut = reinterpret_cast<Out *> (lut_raw);
  const bool perch = p.nch > 1;
  const int entries = p.nch * 256;
  for (int i = threadIdx.x; i < entries; i += CHAIN_THREADS) {
    const uint8_t byte = (uint8_t) (i & 255);
    In xv;
    memcpy (&xv, &byte, 1);
    Vec<Out, 1> y;
    uint32_t ch[1] = { (uint32_t) (i >> 8) };
    y.e[0] = Cast<Out, In, ORC>::run (xv);
    chain_apply<Out, ORC, 1> (y, ch, p, perch);
    lut[i] = y.e[0];
  }
  __syncthreads ();

  const size_t tid = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
  const size_t nthreads = (size_t) gridDim.x * blockDim.x;
  if (p.step_r == 0 && p.step_c == 0)
    lut_stream<In, Out, true, U> (in, out, n_vec, lut, p, tid, nthreads, perch);
  else
    lut_stream<In, Out, false, U> (in, out, n_vec, lut, p, tid, nthreads, perch);

  for (size_t e = n_vec * V + tid; e < n; e += nthreads) {
    In xv;
    ld_stream<1> (&xv, in + e);
    uint8_t byte;
    memcpy (&byte, &xv, 1);
    const uint32_t c = perch ? (uint32_t) (((uint64_t) e / p.ch_size) % (uint64_t) p.nch) : 0;
    const Out y = lut[(size_t) c * 256 + byte];
    st_stream<sizeof (Out)> (out + e, &y);
  }
}

static inline uint64_t
gcd_u64 (uint64_t a, uint64_t b)
{
  while (b) {
    const uint64_t t = a % b;
    a = b;
    b = t;
  }
  return a;
}

/* Shrinks `blocks` so that one grid stride (blocks * threads * V elements) is a multiple of
 * the channel period ch_size * nch; returns 0 if that would cost more than half the grid. */
static inline size_t
blocks_for_invariant_channels (size_t blocks, uint64_t elems_per_block, uint64_t period)
{
  if (period <= 1)
    return blocks;
  const uint64_t need = period / gcd_u64 (elems_per_block, period);
  if (need == 0 || need > blocks / 2 + 1)
    return 0;
  return blocks / need * need;
}

/* ------------------------------------------------------------------------------------- */
/* float32 working type, inputs wider than a byte: the generic kernel above is issue-bound     */
/* (operand fetch, NaN fix-up branches and a ~12-instruction IEEE division per element and     */
/* position: 2.1-4.5 TB/s).  This kernel runs a FAST chain per vector and falls back to the   */
/* careful one (chain_apply, identical by construction) for the vectors that need it:         */
/*   - operands come from registers (scalar chains) or a tiny shared-memory table indexed by   */
/*     a per-lane channel offset that is loop-invariant (the launcher sizes the grid so that   */
/*     one grid stride is a multiple of the channel period);                                   */
/*   - add / mul are the plain IEEE operations; they differ from the careful path only when    */
/*     the result is NaN (x86 payload rules) -- checked once per vector at the end;            */
/*   - x / d with the host-provided r = RN (1 / d) is Markstein's correctly rounded            */
/*     q0 = RN (x r), e = x - d q0 (exact, FMA), q = RN (q0 + e r); valid without over- or       */
/*     underflow, which the kernel guarantees by taking the fast result only for               */
/*     2^-63 <= |x| <= 2^63 (the host admits 2^-30 <= |d| <= 2^30, significand not all ones)   */
/*     and +-0 (q0 has the right sign); anything else marks the vector for the careful path.   */
/* ------------------------------------------------------------------------------------- */
constexpr int FCHAIN_MAX_POS = 4;
constexpr int FCHAIN_MAX_NCH = 64;

/* the careful chain for one vector, out of line: it runs for a handful of vectors at most and
 * must not cost the fast path its registers */
template <int V> struct ChanIdx
{
  uint32_t c[V];
};

template <typename In, bool ORC, int V>
__device__ __noinline__ Vec<float, V>
fchain_careful (Vec<In, V> x, ChanIdx<V> ch, const ChainDev &p, bool perch)
{
  Vec<float, V> y;
#pragma unroll
  for (int k = 0; k < V; k++)
    y.e[k] = Cast<float, In, ORC>::run (x.e[k]);
  chain_apply<float, ORC, V> (y, ch.c, p, perch);
  return y;
}

#ifndef NNSB_FCHAIN_MIN_BLOCKS
#define NNSB_FCHAIN_MIN_BLOCKS 3
#endif
template <typename In, bool ORC, int U>
__global__ void __launch_bounds__ (CHAIN_THREADS, NNSB_FCHAIN_MIN_BLOCKS)
fchain_kernel (const In *__restrict__ in, float *__restrict__ out, size_t n, size_t n_vec,
    const __grid_constant__ ChainDev p)
{
  constexpr int V = ChainVec<In, float>::V;
  __shared__ unsigned long long s_tab[FCHAIN_MAX_POS * FCHAIN_MAX_NCH];
  const size_t tid = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
  const size_t nthreads = (size_t) gridDim.x * blockDim.x;
  const bool perch = p.nch > 1;
  if (perch) {
    /* [channel][position]: a lane's operands sit at (its channel offset) + constant */
    for (int i = threadIdx.x; i < p.n_pos * p.nch; i += CHAIN_THREADS)
      s_tab[(i % p.nch) * FCHAIN_MAX_POS + i / p.nch] = p.table[i].raw;
    __syncthreads ();
  }
  /* channel of lane k: the same for every vector this thread touches */
  uint32_t ch[V];
  {
    uint64_t rr = 0;
    uint32_t cc = 0;
    if (perch) {
      const uint64_t e0 = (uint64_t) tid * V;
      rr = e0 % p.ch_size;
      cc = (uint32_t) ((e0 / p.ch_size) % (uint64_t) p.nch);
    }
#pragma unroll
    for (int k = 0; k < V; k++) {
      ch[k] = cc;
      if (perch && ++rr == p.ch_size) {
        rr = 0;
        cc = (cc + 1 == (uint32_t) p.nch) ? 0 : cc + 1;
      }
    }
  }

  for (size_t v = tid; v < n_vec; v += nthreads * U) {
    Vec<In, V> x[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t vu = v + (size_t) u * nthreads;
      if (vu < n_vec)
        load_vec (x[u], in + vu * V);
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t vu = v + (size_t) u * nthreads;
      if (vu >= n_vec)
        continue;
      Vec<float, V> y;
#pragma unroll
      for (int k = 0; k < V; k++)
        y.e[k] = Cast<float, In, ORC>::run (x[u].e[k]);
      bool bad = false;
#pragma unroll
      for (int pos = 0; pos < FCHAIN_MAX_POS; pos++) {
        if (pos >= p.n_pos)
          break;
        const int opc = p.opcode[pos];
        const unsigned long long sc = p.scalar[pos].raw;
        float c[V], r[V];
#pragma unroll
        for (int k = 0; k < V; k++) {
          const unsigned long long cell = perch ? s_tab[ch[k] * FCHAIN_MAX_POS + pos] : sc;
          c[k] = __uint_as_float ((uint32_t) cell);
          r[k] = __uint_as_float ((uint32_t) (cell >> 32));
        }
        if (opc == NNSB_OP_ADD) {
#pragma unroll
          for (int k = 0; k < V; k++)
            y.e[k] = __fadd_rn (y.e[k], c[k]);
        } else if (opc == NNSB_OP_MUL) {
#pragma unroll
          for (int k = 0; k < V; k++)
            y.e[k] = __fmul_rn (y.e[k], c[k]);
        } else {
#pragma unroll
          for (int k = 0; k < V; k++) {
            const float a = y.e[k];
            const uint32_t ea = (__float_as_uint (a) >> 23) & 0xffu;
            const float q0 = __fmul_rn (a, r[k]);
            const float e = __fmaf_rn (-c[k], q0, a);
            const float q1 = __fmaf_rn (e, r[k], q0);
            const bool zero = a == 0.0f;
            bad = bad || (!zero && (ea - 64u) > 126u);
            y.e[k] = zero ? q0 : q1;
          }
        }
      }
#pragma unroll
      for (int k = 0; k < V; k++)
        bad = bad || !(y.e[k] == y.e[k]);
      if (bad) {
        /* NaN, infinity, denormal or out-of-range dividend somewhere in this vector */
        ChanIdx<V> ci;
#pragma unroll
        for (int k = 0; k < V; k++)
          ci.c[k] = ch[k];
        y = fchain_careful<In, ORC, V> (x[u], ci, p, perch);
      }
      store_vec (out + vu * V, y);
    }
  }

  /* scalar tail (and the whole tensor when the pointers are not vector-aligned) */
  for (size_t e = n_vec * V + tid; e < n; e += nthreads) {
    Vec<float, 1> y;
    uint32_t c1[1] = { 0 };
    In xv;
    ld_stream<sizeof (In)> (&xv, in + e);
    y.e[0] = Cast<float, In, ORC>::run (xv);
    if (perch)
      c1[0] = (uint32_t) (((uint64_t) e / p.ch_size) % (uint64_t) p.nch);
    chain_apply<float, ORC, 1> (y, c1, p, perch);
    st_stream<sizeof (float)> (out + e, &y.e[0]);
  }
}

/* all divisors carry a reciprocal (nnsb_chain_compile) and the chain fits the kernel */
static inline bool
fchain_eligible (const ChainArgs &args)
{
  const int n_pos = args.desc.n_pos, nch = args.desc.nch > 0 ? args.desc.nch : 1;
  if (n_pos < 1 || n_pos > FCHAIN_MAX_POS || nch > FCHAIN_MAX_NCH)
    return false;
  if (nch > 1 && !args.table)
    return false;
  return true;
}

template <typename In, typename Out, bool ORC>
int launch_chain_t (const void *in, void *out, size_t n, const ChainArgs &args,
    const LaunchCfg &cfg, cudaStream_t stream)
{
  constexpr int V = ChainVec<In, Out>::V;
  constexpr int U = 4;
  if (n == 0)
    return 0;
  const bool aligned = ((uintptr_t) in % (V * sizeof (In)) == 0)
      && ((uintptr_t) out % (V * sizeof (Out)) == 0);
  const size_t n_vec = aligned ? n / V : 0;
  const size_t work = n_vec ? n_vec : n;
  /* Many more blocks than fit at once: measured on B200 (profiles/r1_blocks_per_sm_sweep.md),
   * throughput rises monotonically from 8 to 64+ blocks per SM because the hardware block
   * scheduler then balances the SMs dynamically; a persistent one-wave grid is ~10 % slower. */
  const int bps = cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : 256;
  size_t blocks = (work + (size_t) CHAIN_THREADS * U - 1) / ((size_t) CHAIN_THREADS * U);
  const size_t cap = (size_t) cfg.sm_count * bps;
  if (blocks > cap)
    blocks = cap;
  if (blocks == 0)
    blocks = 1;

  ChainDev p;
  p.n_pos = args.desc.n_pos;
  p.nch = args.desc.nch > 0 ? args.desc.nch : 1;
  p.ch_size = args.desc.ch_size ? args.desc.ch_size : 1;
  p.table = args.table;
  for (int i = 0; i < NNSB_MAX_OPS; i++) {
    p.opcode[i] = args.desc.opcode[i];
    p.scalar[i] = args.desc.scalar[i];
  }

  if constexpr (sizeof (In) == 1) {
    /* table path: a table that fits and enough elements to amortise building it (also used
     * for a bare typecast: one lookup is no slower than the conversion and needs fewer
     * registers, i.e. more warps in flight) */
    constexpr int LV = LutVec<Out>::V;
    const size_t lut_bytes = (size_t) p.nch * 256 * sizeof (Out);
    const bool lut_aligned = ((uintptr_t) in % (LV * sizeof (In)) == 0)
        && ((uintptr_t) out % (LV * sizeof (Out)) == 0);
    if (lut_bytes <= 48 * 1024 && n >= 65536 && lut_aligned) {
      const size_t l_vec = n / LV;
      /* every block rebuilds the table, so keep ~20 K elements per block */
      const int lres = cfg.blocks_per_sm > 0 ? cfg.blocks_per_sm : 64;
      const size_t lcap = (size_t) cfg.sm_count * lres;
      size_t lb = (l_vec + (size_t) CHAIN_THREADS * U - 1) / ((size_t) CHAIN_THREADS * U);
      if (lb > lcap)
        lb = lcap;
      if (lb == 0)
        lb = 1;
      const uint64_t period = p.ch_size * (uint64_t) p.nch;
      const size_t li = blocks_for_invariant_channels (lb, (uint64_t) CHAIN_THREADS * LV, period);
      if (li)
        lb = li;
      const uint64_t stride = (uint64_t) lb * CHAIN_THREADS * LV;
      p.step_r = stride % p.ch_size;
      p.step_c = (uint32_t) ((stride / p.ch_size) % (uint64_t) p.nch);
      lut_chain_kernel<In, Out, ORC, U><<<(unsigned int) lb, CHAIN_THREADS, lut_bytes, stream>>> (
          (const In *) in, (Out *) out, n, l_vec, p);
      return (int) cudaGetLastError ();
    }
  }

  if constexpr (sizeof (Out) == 4 && IsFloat<Out>::v && sizeof (In) > 1) {
    if (args.fast_f32 && fchain_eligible (args)) {
      const uint64_t period = p.ch_size * (uint64_t) p.nch;
      const size_t fb = blocks_for_invariant_channels (blocks, (uint64_t) CHAIN_THREADS * V, period);
      if (fb) {
        fchain_kernel<In, ORC, U><<<(unsigned int) fb, CHAIN_THREADS, 0, stream>>> ((const In *) in,
            (float *) out, n, n_vec, p);
        return (int) cudaGetLastError ();
      }
    }
  }

  const uint64_t stride_elems = (uint64_t) blocks * CHAIN_THREADS * V;
  p.step_r = stride_elems % p.ch_size;
  p.step_c = (uint32_t) ((stride_elems / p.ch_size) % (uint64_t) p.nch);
  chain_kernel<In, Out, ORC, U><<<(unsigned int) blocks, CHAIN_THREADS, 0, stream>>> (
      (const In *) in, (Out *) out, n, n_vec, p);
  return (int) cudaGetLastError ();
}

typedef int (*chain_fn) (const void *, void *, size_t, const ChainArgs &, const LaunchCfg &,
    cudaStream_t);

/* one row (fixed In) of the dispatch table, both policies */
template <typename In> struct ChainRow
{
  static chain_fn get (int out_type, bool orc)
  {
    constexpr bool in_orc_ok = !Is64Int<In>::v && TypeId<In>::id != NNSB_FLOAT16;
    switch (out_type) {
#define NNSB_CASE(T) \
      case TypeId<T>::id: \
        if constexpr (in_orc_ok && !Is64Int<T>::v && TypeId<T>::id != NNSB_FLOAT16) { \
          if (orc) return &launch_chain_t<In, T, true>; \
        } \
        return &launch_chain_t<In, T, false>;
      NNSB_CASE (int32_t) NNSB_CASE (uint32_t) NNSB_CASE (int16_t) NNSB_CASE (uint16_t)
      NNSB_CASE (int8_t) NNSB_CASE (uint8_t) NNSB_CASE (double) NNSB_CASE (float)
      NNSB_CASE (int64_t) NNSB_CASE (uint64_t) NNSB_CASE (__half)
#undef NNSB_CASE
      default:
        return nullptr;
    }
  }
};

}  // namespace nnsb
#endif
