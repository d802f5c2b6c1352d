/*
 * nnsb_internal.h -- declarations shared by the translation units of libnnstreamer_b200.so.
 * Not part of the public ABI (see include/nnstreamer_b200.h).
 */
#ifndef NNSB_INTERNAL_H
#define NNSB_INTERNAL_H

#include "nnstreamer_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* records the thread-local error message and returns `code` */
int nnsb_set_error (int code, const char *fmt, ...) __attribute__((format (printf, 2, 3)));

/* 8-byte cell holding one operand already cast to the working type (tensor_element,
 * tensor_typedef.h:198-212) */
typedef union {
  int32_t i32; uint32_t u32; int16_t i16; uint16_t u16; int8_t i8; uint8_t u8;
  double f64; float f32; int64_t i64; uint64_t u64; uint16_t f16_bits;
  uint64_t raw;
} nnsb_cell;

/* One position of a compiled arithmetic chain: an opcode shared by all channels and one
 * operand per channel (identity where the operator does not apply). */
#define NNSB_CHAIN_NOP 0
typedef struct {
  int32_t n_pos;                        /* chain length after compilation */
  int32_t nch;                          /* 1 = not per-channel */
  uint64_t ch_size;                     /* product of dims below ch_dim */
  int32_t opcode[NNSB_MAX_OPS];         /* nnsb_operator per position */
  nnsb_cell scalar[NNSB_MAX_OPS];       /* operand when nch == 1 */
} nnsb_chain_desc;

/* Host-side compilation of the operator list into positions (nnsb_plan_host.cpp).
 * table (n_pos * nch cells, row-major [pos][ch]) is malloc()ed when nch > 1. */
int nnsb_chain_compile (const nnsb_option * opt, int work_type, int acceleration,
    const uint32_t *dim, nnsb_chain_desc * desc, nnsb_cell ** table);

/* C-cast of an operand literal to `type` with the reference's rules
 * (gst_tensor_data_typecast, tensor_data.c:202-279) */
nnsb_cell nnsb_cast_operand (const nnsb_arith_op * op, int type);

/* clamp bounds / stand: nothing to precompile */

/* wraps a raw allocation of `ctx` (nnsb_dev_alloc / nnsb_pinned_alloc) into an owned,
 * reference-counted memory (nnsb_runtime.cu); on failure the allocation stays raw */
int nnsb_memory_adopt (nnsb_ctx * ctx, void *ptr, size_t size, int on_device, nnsb_memory * out);
/* 1 when `mem` holds the only reference on its block */
int nnsb_memory_is_unique (const nnsb_memory * mem);
/* 1 = page-locked / managed host memory (the copy engine reads it directly), 0 = pageable */
int nnsb_host_is_pinned (const void *p);

#ifdef __cplusplus
}
#endif
#endif
