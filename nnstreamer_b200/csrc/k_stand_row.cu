/*
 * k_stand_row.cu -- instantiates the stand kernels for ONE input type
 * (compile with -DNNSB_ROW=<nnsb_tensor_type>).
 */
#include "k_stand.cuh"

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

stand_stats_fn NNSB_CAT (stand_stats_row_, NNSB_ROW) (void)
{
  return StandRow<RowT>::stats ();
}

stand_apply_fn NNSB_CAT (stand_apply_row_, NNSB_ROW) (int out_type)
{
  return StandRow<RowT>::apply (out_type);
}

}  // namespace nnsb
