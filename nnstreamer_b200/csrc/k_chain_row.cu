/*
 * k_chain_row.cu -- instantiates the chain kernels for ONE input type (compile with
 * -DNNSB_ROW=<nnsb_tensor_type>), so the 11 rows build in parallel.
 */
#include "k_chain.cuh"

#ifndef NNSB_ROW
#error "compile with -DNNSB_ROW=<0..10>"
#endif

namespace nnsb {

#if NNSB_ROW == 0
typedef int32_t RowT;
#elif NNSB_ROW == 1
typedef uint32_t RowT;
#elif NNSB_ROW == 2
typedef int16_t RowT;
#elif NNSB_ROW == 3
typedef uint16_t RowT;
#elif NNSB_ROW == 4
typedef int8_t RowT;
#elif NNSB_ROW == 5
typedef uint8_t RowT;
#elif NNSB_ROW == 6
typedef double RowT;
#elif NNSB_ROW == 7
typedef float RowT;
#elif NNSB_ROW == 8
typedef int64_t RowT;
#elif NNSB_ROW == 9
typedef uint64_t RowT;
#elif NNSB_ROW == 10
typedef __half RowT;
#endif

#define NNSB_CAT2(a, b) a##b
#define NNSB_CAT(a, b) NNSB_CAT2 (a, b)

chain_fn NNSB_CAT (chain_row_, NNSB_ROW) (int out_type, bool orc)
{
  return ChainRow<RowT>::get (out_type, orc);
}

}  // namespace nnsb
