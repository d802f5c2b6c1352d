/*
 * nnstreamer_b200.h -- C-ABI of the B200-native tensor_transform / tensor_converter hot path.
 *
 * This is the drop-in boundary: plain C types, plain pointers and
 * sizes, no C++/GLib/torch types.  Everything above it (the GStreamer glue in
 * nnstreamer_b200/gst/, the Python mirror in nnstreamer_b200/element.py, bench.py, tests/)
 * talks to the GPU only through these entry points.  Each declaration cites the reference
 * interface (file:line under the NNStreamer 2.3.0 tree) it replaces.
 *
 * Conventions
 *   - every call returns 0 on success or a negative errno-style code (NNSB_E*);
 *     nnsb_last_error() returns the thread-local message (CUDA error string included).
 *   - there is NO CPU fallback: if no CUDA device / kernel image is usable the call fails.
 *   - dimension order is innermost-first (dim[0] = channel for video), rank limit 8,
 *     exactly as tensor_typedef.h:34,212.
 */
#ifndef NNSTREAMER_B200_H
#define NNSTREAMER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NNSB_API __attribute__((visibility("default")))

/* ---------------------------------------------------------------------------------------
 * Types (tensor_typedef.h:34-36,133-148,163-171,232-260)
 * ------------------------------------------------------------------------------------- */
#define NNSB_RANK_LIMIT 8            /* NNS_TENSOR_RANK_LIMIT, tensor_typedef.h:34 */
#define NNSB_SIZE_LIMIT 16           /* NNS_TENSOR_SIZE_LIMIT, tensor_typedef.h:36 */
#define NNSB_META_RANK_LIMIT 16      /* NNS_TENSOR_META_RANK_LIMIT, tensor_typedef.h:45 */
#define NNSB_TRANSPOSE_RANK_LIMIT 4  /* gsttensor_transform.c:86 */
#define NNSB_MAX_OPS 64              /* arithmetic chain length cap (reference: unbounded GSList) */

/* tensor_type, tensor_typedef.h:133-148 -- the numeric values are wire format (flex header) */
typedef enum {
  NNSB_INT32 = 0, NNSB_UINT32, NNSB_INT16, NNSB_UINT16, NNSB_INT8, NNSB_UINT8,
  NNSB_FLOAT64, NNSB_FLOAT32, NNSB_INT64, NNSB_UINT64, NNSB_FLOAT16,
  NNSB_TYPE_END
} nnsb_tensor_type;

/* tensor_format, tensor_typedef.h:186-193 */
typedef enum {
  NNSB_FORMAT_STATIC = 0, NNSB_FORMAT_FLEXIBLE, NNSB_FORMAT_SPARSE, NNSB_FORMAT_END
} nnsb_tensor_format;

/* tensor_transform_mode, gsttensor_transform.h:57-67 */
typedef enum {
  NNSB_MODE_DIMCHG = 0, NNSB_MODE_TYPECAST, NNSB_MODE_ARITHMETIC, NNSB_MODE_TRANSPOSE,
  NNSB_MODE_STAND, NNSB_MODE_CLAMP, NNSB_MODE_UNKNOWN = -1
} nnsb_mode;

/* tensor_transform_operator, gsttensor_transform.h:69-77 */
typedef enum {
  NNSB_OP_TYPECAST = 0, NNSB_OP_ADD = 1, NNSB_OP_MUL = 2, NNSB_OP_DIV = 3, NNSB_OP_UNKNOWN
} nnsb_operator;

/* tensor_transform_stand_mode, gsttensor_transform.h:79-84 */
typedef enum { NNSB_STAND_DEFAULT = 0, NNSB_STAND_DC_AVERAGE = 1, NNSB_STAND_END } nnsb_stand_mode;

/* GstPadDirection as used by convert_dimension, gsttensor_transform.c:1850-1853 */
typedef enum { NNSB_PAD_SRC = 1, NNSB_PAD_SINK = 2 } nnsb_pad_direction;

/* GstTensorInfo (type + dimension), tensor_typedef.h:232-239 */
typedef struct {
  int32_t type;                         /* nnsb_tensor_type */
  uint32_t dim[NNSB_RANK_LIMIT];
} nnsb_tensor_info;

/* GstTensorsConfig / GstTensorsInfo, tensor_typedef.h:245-260 */
typedef struct {
  uint32_t num_tensors;
  nnsb_tensor_info info[NNSB_SIZE_LIMIT];
  int32_t format;                       /* nnsb_tensor_format */
  int32_t rate_n, rate_d;
} nnsb_tensors_config;

/* GstTensorMetaInfo, tensor_typedef.h:281-296 (the 128-byte flexible-tensor header) */
typedef struct {
  uint32_t version;
  uint32_t type;
  uint32_t dimension[NNSB_META_RANK_LIMIT];
  uint32_t format;
  uint32_t media_type;
  uint32_t sparse_nnz;
} nnsb_meta_info;

/* tensor_transform_operator_s, gsttensor_transform.h:104-109.  The operand keeps the
 * literal's parse class (double if it contains '.', 'e' or 'E', else int64 --
 * gsttensor_transform.c:796-807); the cast to the working type happens at plan time. */
typedef struct {
  int32_t op;                           /* nnsb_operator */
  int32_t applying_ch;                  /* -1 = every channel (:779) */
  int32_t is_float;                     /* operand literal class */
  int32_t cast_type;                    /* target of a leading typecast op */
  double dval;
  int64_t ival;
} nnsb_arith_op;

/* The data_* union of GstTensorTransform, gsttensor_transform.h:153-160, flattened. */
typedef struct {
  int32_t mode;                         /* nnsb_mode */
  int32_t dimchg_from, dimchg_to;       /* tensor_transform_dimchg  (.h:89-92) */
  int32_t typecast_to;                  /* tensor_transform_typecast (.h:97-99) */
  int32_t arith_out_type;               /* tensor_transform_arithmetic (.h:114-118) */
  int32_t arith_per_channel;
  uint32_t arith_ch_dim;
  int32_t n_ops;
  nnsb_arith_op ops[NNSB_MAX_OPS];      /* the `operators` GSList (.h:163) */
  uint8_t trans_order[NNSB_RANK_LIMIT]; /* tensor_transform_transpose (.h:123-125) */
  int32_t stand_mode;                   /* tensor_transform_stand (.h:130-134) */
  int32_t stand_out_type;
  int32_t stand_per_channel;
  double clamp_min, clamp_max;          /* tensor_transform_clamp (.h:139-141) */
} nnsb_option;

/* Error codes (negative errno values) */
#define NNSB_OK 0
#define NNSB_EINVAL (-22)
#define NNSB_ENOMEM (-12)
#define NNSB_ENOTSUP (-95)
#define NNSB_ECUDA (-5)                 /* maps to GST_FLOW_ERROR in the glue */
#define NNSB_ERANGE (-34)

/* Thread-local message of the last failing call on this thread. */
NNSB_API const char *nnsb_last_error (void);

/* ---------------------------------------------------------------------------------------
 * Host core: type / dimension helpers (nnstreamer_plugin_api_util_impl.c)
 * ------------------------------------------------------------------------------------- */
/* gst_tensor_get_type, util_impl.c:1098-1170 (strips blanks, case-insensitive) */
NNSB_API int nnsb_type_from_string (const char *typestr);
/* gst_tensor_get_type_string, util_impl.c:1175-1181; NULL for NNSB_TYPE_END */
NNSB_API const char *nnsb_type_to_string (int type);
/* gst_tensor_get_element_size, util_impl.c:1085-1091 */
NNSB_API size_t nnsb_element_size (int type);
/* gst_tensor_get_element_count, util_impl.c:1069-1080 */
NNSB_API uint64_t nnsb_element_count (const uint32_t *dim);
/* gst_tensor_info_get_size, util_impl.c:155-166 */
NNSB_API size_t nnsb_info_size (const nnsb_tensor_info * info);
/* gst_tensor_parse_dimension, util_impl.c:921-956; returns the rank (0 on error) */
NNSB_API int nnsb_dim_parse (const char *dimstr, uint32_t *dim);
/* gst_tensor_get_dimension_string / _rank_dimension_string, util_impl.c:964-1017.
 * rank 0 prints all 8. Returns the string length or NNSB_EINVAL if buf is too small. */
NNSB_API int nnsb_dim_to_string (const uint32_t *dim, unsigned int rank, char *buf, size_t buflen);
/* gst_tensor_dimension_is_valid, util_impl.c:893-913 */
NNSB_API int nnsb_dim_is_valid (const uint32_t *dim);

/* Flexible-tensor header (util_impl.c:1213-1424).  Header size is fixed at 128 bytes (:1330). */
#define NNSB_META_HEADER_SIZE 128
/* gst_tensor_info_convert_to_meta, util_impl.c:1471-1495 */
NNSB_API int nnsb_meta_from_info (const nnsb_tensor_info * info, nnsb_meta_info * meta);
/* gst_tensor_meta_info_convert, util_impl.c:1432-1464 */
NNSB_API int nnsb_meta_to_info (const nnsb_meta_info * meta, nnsb_tensor_info * info);
/* gst_tensor_meta_info_update_header, util_impl.c:1374-1389 */
NNSB_API int nnsb_meta_write_header (const nnsb_meta_info * meta, void *header);
/* gst_tensor_meta_info_parse_header, util_impl.c:1397-1424 */
NNSB_API int nnsb_meta_parse_header (nnsb_meta_info * meta, const void *header);
/* gst_tensor_meta_info_get_data_size, util_impl.c:1340-1364 */
NNSB_API size_t nnsb_meta_data_size (const nnsb_meta_info * meta);

/* ---------------------------------------------------------------------------------------
 * Host core: mode / option grammar and dimension conversion (gsttensor_transform.c)
 * ------------------------------------------------------------------------------------- */
/* GEnum nicks of gtt_mode_type, gsttensor_transform.c:180-204 */
NNSB_API int nnsb_mode_from_string (const char *nick);
NNSB_API const char *nnsb_mode_to_string (int mode);

/* gst_tensor_transform_set_option_data, gsttensor_transform.c:652-958: validates `option`
 * against the per-mode grammar (REGEX_* :70-80) and fills `out`.  Returns 0 when the
 * reference would set `loaded = TRUE`, NNSB_EINVAL otherwise (out->mode is then UNKNOWN). */
NNSB_API int nnsb_option_parse (int mode, const char *option, nnsb_option * out);

/* gst_tensor_transform_convert_dimension, gsttensor_transform.c:1850-1959.
 * direction SINK: in -> out; SRC: out -> in (type becomes NNSB_TYPE_END where unknowable). */
NNSB_API int nnsb_convert_dimension (const nnsb_option * opt, int direction,
    const nnsb_tensor_info * in_info, nnsb_tensor_info * out_info);

/* ---------------------------------------------------------------------------------------
 * Device runtime: replaces gst_allocator_alloc(NULL,...) + gst_memory_map of
 * gsttensor_transform.c:1725,1758-1766 and tensor_allocator.c:101-128.
 * ------------------------------------------------------------------------------------- */
typedef struct nnsb_ctx nnsb_ctx;        /* one per (element instance, device) */
typedef struct nnsb_stream nnsb_stream;  /* one CUDA stream; per element instance */
typedef struct nnsb_event nnsb_event;

NNSB_API int nnsb_device_count (int *count);
NNSB_API int nnsb_ctx_create (int device_id, nnsb_ctx ** out);
NNSB_API int nnsb_ctx_destroy (nnsb_ctx * ctx);
NNSB_API int nnsb_ctx_device (const nnsb_ctx * ctx);
/* number of SMs of the bound device (148 on B200) */
NNSB_API int nnsb_ctx_sm_count (const nnsb_ctx * ctx);

NNSB_API int nnsb_stream_create (nnsb_ctx * ctx, nnsb_stream ** out);
/* wrap an existing cudaStream_t (e.g. torch's current stream); not owned */
NNSB_API int nnsb_stream_wrap (nnsb_ctx * ctx, void *cuda_stream, nnsb_stream ** out);
NNSB_API int nnsb_stream_sync (nnsb_stream * s);
NNSB_API int nnsb_stream_destroy (nnsb_stream * s);

NNSB_API int nnsb_event_create (nnsb_ctx * ctx, nnsb_event ** out);
NNSB_API int nnsb_event_record (nnsb_event * e, nnsb_stream * s);
NNSB_API int nnsb_event_sync (nnsb_event * e);
/* make stream `s` wait for event `e` (cross-stream ordering inside one element) */
NNSB_API int nnsb_stream_wait_event (nnsb_stream * s, nnsb_event * e);
NNSB_API int nnsb_event_elapsed_ms (nnsb_event * start, nnsb_event * stop, float *ms);
NNSB_API int nnsb_event_destroy (nnsb_event * e);

/* Size-bucketed device pool: freed blocks are kept and reused, so steady-state streaming
 * performs no cudaMalloc.  Blocks are 256-byte aligned. */
NNSB_API int nnsb_dev_alloc (nnsb_ctx * ctx, size_t size, void **out);
NNSB_API int nnsb_dev_free (nnsb_ctx * ctx, void *ptr);
/* nnsb_dev_free returns the block to the pool at once: the memory must be idle (no queued GPU
 * work still touches it).  The _async pair is stream-ordered like cudaMallocAsync: a block
 * freed on `s` may be handed out again immediately to work on `s`, and to other streams only
 * after the work queued on `s` at the time of the free has completed (the pool inserts the
 * wait; the host never blocks). */
NNSB_API int nnsb_dev_alloc_async (nnsb_stream * s, size_t size, void **out);
NNSB_API int nnsb_dev_free_async (nnsb_stream * s, void *ptr);
/* number of cudaMalloc calls this context has made so far (pool growth; tests) */
NNSB_API uint64_t nnsb_ctx_device_mallocs (const nnsb_ctx * ctx);
/* release cached blocks back to the driver */
NNSB_API int nnsb_dev_trim (nnsb_ctx * ctx);
NNSB_API int nnsb_dev_memset_async (nnsb_stream * s, void *dst, int value, size_t n);
/* pooled page-locked host staging memory */
NNSB_API int nnsb_pinned_alloc (nnsb_ctx * ctx, size_t size, void **out);
NNSB_API int nnsb_pinned_free (nnsb_ctx * ctx, void *ptr);

NNSB_API int nnsb_h2d_async (nnsb_stream * s, void *dst_dev, const void *src_host, size_t n);
NNSB_API int nnsb_d2h_async (nnsb_stream * s, void *dst_host, const void *src_dev, size_t n);
NNSB_API int nnsb_d2d_async (nnsb_stream * s, void *dst_dev, const void *src_dev, size_t n);
/* pitched H2D: strips the video row padding for free on the copy engine
 * (the copy gsttensor_converter.c:1087-1100 does with one memcpy per row) */
NNSB_API int nnsb_h2d_2d_async (nnsb_stream * s, void *dst_dev, size_t dst_pitch,
    const void *src_host, size_t src_pitch, size_t row_bytes, size_t rows);

/* ---------------------------------------------------------------------------------------
 * Plans: one parsed (mode, option) bound to one input tensor info -> the kernels.
 * Replaces the switch of gst_tensor_transform_transform, gsttensor_transform.c:1773-1802,
 * and the six functions it calls (:1120 dimchg, :1202 typecast, :1240 arithmetic,
 * :1443 transpose, :1512 stand, :1632 clamp).
 * ------------------------------------------------------------------------------------- */
typedef struct nnsb_plan nnsb_plan;

/* acceleration: the element's `acceleration` property (gsttensor_transform.c:241-243).
 * 0 = plain C semantics (integer ops wrap); 1 = ORC semantics for the types ORC supports
 * (saturating integer add/mul, signed-saturating narrowing; orc_supported, :462-463). */
NNSB_API int nnsb_plan_create (nnsb_ctx * ctx, const nnsb_option * opt,
    const nnsb_tensor_info * in_info, int acceleration, nnsb_plan ** out);
NNSB_API int nnsb_plan_destroy (nnsb_plan * plan);
NNSB_API int nnsb_plan_out_info (const nnsb_plan * plan, nnsb_tensor_info * out_info);
NNSB_API size_t nnsb_plan_in_size (const nnsb_plan * plan);   /* bytes of one input tensor */
NNSB_API size_t nnsb_plan_out_size (const nnsb_plan * plan);  /* bytes of one output tensor */

/* Transform `batch` tensors laid out back to back (frame f at d_in + f*in_size ->
 * d_out + f*out_size) with ONE set of launches.  Per-tensor semantics are preserved
 * (stand computes its statistics per tensor), i.e. the result equals `batch` calls of the
 * reference function.  d_in/d_out are device pointers; asynchronous on `s`. */
NNSB_API int nnsb_plan_launch (nnsb_plan * plan, const void *d_in, void *d_out,
    uint32_t batch, nnsb_stream * s);
/* Element-range launch for the element-wise modes (typecast, arithmetic, clamp): transforms
 * elements [first_elem, first_elem + n_elems) of the buffers d_in / d_out point to.  Lets the
 * host-buffer path pipeline H2D copy, kernel and D2H copy chunk by chunk on two streams.
 * first_elem must be a multiple of nnsb_plan_chunk_granule() (keeps the 32-byte vector
 * alignment and, for per-channel arithmetic, the channel phase).  No reference counterpart:
 * the reference's loops (gsttensor_transform.c:1222-1226,1375-1409) run on host memory. */
NNSB_API int nnsb_plan_is_elementwise (const nnsb_plan * plan);
NNSB_API uint64_t nnsb_plan_chunk_granule (const nnsb_plan * plan);
NNSB_API int nnsb_plan_launch_elems (nnsb_plan * plan, const void *d_in, void *d_out,
    uint64_t first_elem, uint64_t n_elems, nnsb_stream * s);
/* number of kernels the last nnsb_plan_launch on this plan enqueued */
NNSB_API int nnsb_plan_last_launch_count (const nnsb_plan * plan);
/* One line naming the kernel path the plan compiled to (e.g. "permute: vector table, box
 * 64x64, swizzle 6"), for logs and tests.  Returns the string length or NNSB_EINVAL.  No
 * reference counterpart (the reference has a single code path per mode). */
NNSB_API int nnsb_plan_describe (const nnsb_plan * plan, char *buf, size_t buflen);
/* stand only: copy the per-tensor statistics of the last launch (mean[, std]) to host */
NNSB_API int nnsb_plan_stand_stats (nnsb_plan * plan, nnsb_stream * s, uint32_t batch,
    double *mean_out, double *std_out, size_t capacity);

/* Plan for a raw strided copy: `rank` output dims (innermost first) of extent size[i], whose
 * elements (es bytes each: 1, 2, 4 or 8) are read from the input at stride in_stride[i]
 * (elements); the output is dense.  Runs on the same kernels as transpose / dimchg and is
 * launched with nnsb_plan_launch.  Used by the aggregator for the concatenation along an inner
 * axis (gsttensor_aggregator.c:768-787, a block interleave of frames_out frames). */
NNSB_API int nnsb_plan_create_permute (nnsb_ctx * ctx, int es, int rank, const uint32_t *size,
    const uint64_t *in_stride, nnsb_plan ** out);

/* tensor_converter video row-stride strip, gsttensor_converter.c:1087-1100: copies `rows`
 * rows of `row_bytes` from a source with pitch `src_pitch` into a dense destination.
 * Both pointers are device memory. */
NNSB_API int nnsb_strip_launch (nnsb_stream * s, const void *d_src, void *d_dst,
    size_t row_bytes, size_t src_pitch, size_t rows);

/* tensor_decoder direct_video row padding, tensordec-directvideo.c:326-339 (the inverse of the
 * strip): `rows` dense rows of `row_bytes` go to a destination with pitch `dst_pitch`; the
 * padding bytes are written as zero (the reference leaves them uninitialised).  Device memory. */
NNSB_API int nnsb_pad_launch (nnsb_stream * s, const void *d_src, void *d_dst,
    size_t row_bytes, size_t dst_pitch, size_t rows);
/* pitched D2H: the same padding done by the copy engine on the way to a host video buffer
 * (padding bytes of the host buffer are left untouched) */
NNSB_API int nnsb_d2h_2d_async (nnsb_stream * s, void *dst_host, size_t dst_pitch,
    const void *src_dev, size_t src_pitch, size_t row_bytes, size_t rows);

/* tensor_merge mode=linear, gsttensor_merge.c:506-586: output row r = block r of every input
 * back to back; block_bytes[k] = element size * prod (dim_k[0..axis]), rows = prod of the dims
 * above the axis.  n <= 16, device pointers. */
NNSB_API int nnsb_merge_launch (nnsb_stream * s, uint32_t n, const void *const *d_in,
    const uint64_t *block_bytes, void *d_out, uint64_t rows);

/* ---------------------------------------------------------------------------------------
 * Element mirrors.  GStreamer is not available in the build image, so the GObject /
 * GstBaseTransform surface of the two elements is exposed as plain functions with the same
 * names, property strings, caps strings and return conventions; the GStreamer glue
 * (nnstreamer_b200/gst/) is a 1:1 wrapper around them.
 * ------------------------------------------------------------------------------------- */
#define NNSB_FLOW_OK 0                  /* GST_FLOW_OK */
#define NNSB_FLOW_ERROR (-5)            /* GST_FLOW_ERROR */
#define NNSB_FLOW_NOT_SUPPORTED (-6)    /* GST_FLOW_NOT_SUPPORTED */
#define NNSB_ENOENT (-2)                /* property value is NULL (e.g. `option` never set) */

/* One GstMemory: a view of a block of host or device memory.
 *
 * Blocks produced by the elements are reference counted (GstMiniObject refcount of a
 * GstMemory): `owned` = this view holds one reference, dropped by nnsb_memory_release; the
 * block goes back to its pool with the last reference.  Views of caller-managed memory
 * (block == NULL, owned == 0) are never freed here.
 *
 * Device memories are STREAM-ORDERED: an element returns as soon as its kernels are queued,
 * and the block carries the CUDA event after which its contents are complete.  A consumer
 *   - that is one of these elements waits on that event on its own stream (no host sync);
 *   - that launches its own CUDA work calls nnsb_memory_wait (mem, stream) or takes the raw
 *     event from nnsb_memory_export, and afterwards nnsb_memory_mark_use so the block is not
 *     recycled under it;
 *   - that reads the bytes from the host, or through a raw pointer on an unrelated stream,
 *     calls nnsb_memory_sync first (what gst_memory_map does in the CUDA allocator glue).
 * A caller-managed device input (block == NULL) must stay untouched until the outputs
 * computed from it are ready. */
typedef struct {
  void *data;
  size_t size;
  int32_t on_device;                    /* 0 = host, 1 = device memory of the element's GPU */
  int32_t owned;                        /* 1 = holds a reference on `block` */
  void *owner;                          /* the context that allocated it */
  void *block;                          /* opaque refcounted allocation, NULL = caller-managed */
} nnsb_memory;

/* What a CUDA-aware consumer needs to use a device memory in place: the
 * hand-off to tensor_filter sub-plugins that want device pointers, e.g.
 * ext/nnstreamer/tensor_filter/tensor_filter_tensorrt.cc:390, instead of the host round trip
 * of gst_memory_map in tensor_filter.c:675. */
typedef struct {
  void *device_ptr;
  size_t size;
  int32_t device_id;
  int32_t reserved;
  void *ready_event;                    /* cudaEvent_t, NULL = contents are complete already */
} nnsb_device_view;

/* Memory life cycle.  No reference counterpart for the stream ordering; ref / share / release
 * are gst_memory_ref, gst_memory_share (nnstreamer_plugin_api_impl.c:534, tensor_demux
 * gsttensor_demux.c:516-522, tensor_mux gsttensor_mux.c:482) and gst_memory_unref. */
/* a pooled block of `size` bytes; `s` (may be NULL) = the stream that will touch it first */
NNSB_API int nnsb_memory_alloc (nnsb_ctx * ctx, size_t size, int on_device, nnsb_stream * s,
    nnsb_memory * out);
NNSB_API int nnsb_memory_ref (const nnsb_memory * mem, nnsb_memory * out);
/* sub-range [offset, offset + size) of the same block; keeps the block alive on its own */
NNSB_API int nnsb_memory_share (const nnsb_memory * mem, size_t offset, size_t size, nnsb_memory * out);
NNSB_API int nnsb_memory_release (nnsb_memory * mem);
/* stream `s` waits (on the GPU) until the contents are complete; no-op when they are */
NNSB_API int nnsb_memory_wait (const nnsb_memory * mem, nnsb_stream * s);
/* the calling thread waits until the contents are complete */
NNSB_API int nnsb_memory_sync (const nnsb_memory * mem);
/* "the work queued on `s` so far produces outs[0..n_out) and reads ins[0..n_in)": records one
 * event; outs become ready at it, and ins are not recycled before it.  Either list may be empty. */
NNSB_API int nnsb_memory_mark_io (nnsb_stream * s, nnsb_memory * outs, uint32_t n_out,
    const nnsb_memory * ins, uint32_t n_in);
NNSB_API int nnsb_memory_mark_ready (nnsb_memory * mem, nnsb_stream * s);
NNSB_API int nnsb_memory_mark_use (const nnsb_memory * mem, nnsb_stream * s);
/* device pointer + ready event of a device memory, for CUDA-aware consumers */
NNSB_API int nnsb_memory_export (const nnsb_memory * mem, nnsb_device_view * view);
/* the raw cudaStream_t behind a stream object (for consumers that launch their own work) */
NNSB_API void *nnsb_stream_handle (nnsb_stream * s);

typedef struct nnsb_transform nnsb_transform;

/* gst_tensor_transform_init, gsttensor_transform.c:288-301 (silent=TRUE, mode=unknown,
 * option=NULL, acceleration=FALSE as in a build without ORC).  The GPU context is created
 * on the first transform, so properties and caps work on a machine without a GPU. */
NNSB_API int nnsb_transform_new (int device_id, nnsb_transform ** out);
NNSB_API int nnsb_transform_free (nnsb_transform * t);
/* gst_tensor_transform_set_property, :963-1025.  Properties: "silent", "mode" (nick or
 * number), "option", "acceleration", "apply".  An invalid option restores the previous one
 * (:985-990) and returns NNSB_EINVAL; the reference only logs. */
NNSB_API int nnsb_transform_set_property (nnsb_transform * t, const char *name, const char *value);
/* gst_tensor_transform_get_property, :1030-1081 (+ "transpose-rank-limit").  Returns the
 * string length, NNSB_ENOENT when the value is NULL, NNSB_EINVAL for unknown names. */
NNSB_API int nnsb_transform_get_property (nnsb_transform * t, const char *name, char *buf, size_t buflen);
/* gst_tensor_transform_transform_caps, :1971-2061.  direction: nnsb_pad_direction of `caps`;
 * filter may be NULL.  Result written to out (caps string). */
NNSB_API int nnsb_transform_transform_caps (nnsb_transform * t, int direction, const char *caps,
    const char *filter, char *out, size_t outlen);
/* gst_tensor_transform_fixate_caps, :2066-2088 */
NNSB_API int nnsb_transform_fixate_caps (nnsb_transform * t, int direction, const char *caps,
    const char *othercaps, char *out, size_t outlen);
/* gst_tensor_transform_set_caps, :2093-2164; 0 = accepted */
NNSB_API int nnsb_transform_set_caps (nnsb_transform * t, const char *incaps, const char *outcaps);
/* gst_tensor_transform_transform_size, :2169-2186: always succeeds with *othersize = 0 */
NNSB_API int nnsb_transform_transform_size (nnsb_transform * t, int direction, const char *caps,
    size_t size, const char *othercaps, size_t *othersize);
/* gst_tensor_transform_transform, :1667-1814: n_in input memories -> *n_out output memories
 * appended to out[] (capacity NNSB_SIZE_LIMIT).  Host inputs give (pinned) host outputs with
 * the copies pipelined against the kernels, complete on return; device inputs give device
 * outputs that are stream-ordered (see nnsb_memory: the call returns once the kernels are
 * queued).  Tensors skipped by `apply` come out as new references to the input block
 * (gst_memory_ref, :1716).  Returns an NNSB_FLOW_* code. */
NNSB_API int nnsb_transform_transform (nnsb_transform * t, const nnsb_memory * in, uint32_t n_in,
    nnsb_memory * out, uint32_t *n_out);
/* the negotiated configs (filter->in_config / out_config) */
NNSB_API int nnsb_transform_get_config (nnsb_transform * t, int direction, nnsb_tensors_config * cfg);
/* element-owned memory helpers (the CUDA GstAllocator's alloc / free) */
NNSB_API int nnsb_transform_alloc (nnsb_transform * t, size_t size, int on_device, nnsb_memory * out);
/* synchronous copies between a device memory of this element and host memory: what the CUDA
 * GstAllocator's mem_map / mem_unmap do for peers that are not CUDA-aware (gst_memory_map,
 * gsttensor_transform.c:1725,1761) */
NNSB_API int nnsb_transform_download (nnsb_transform * t, void *host_dst, const nnsb_memory * dev_src);
NNSB_API int nnsb_transform_upload (nnsb_transform * t, const nnsb_memory * dev_dst, const void *host_src);
/* kernels launched by this element so far */
NNSB_API int nnsb_transform_kernel_launches (nnsb_transform * t, uint64_t *count);
/* Where device outputs come from: by default the element's own pool (the reference calls
 * gst_allocator_alloc (NULL, ...), :1758); a provider lets the DOWNSTREAM element hand out the
 * memory instead (GstBaseTransform's decide_allocation choosing the peer's pool), so that e.g.
 * tensor_aggregator receives its frames where it wants them.  The provider returns 0 and an
 * owned device memory of exactly `size` bytes, or an error (the element's pool is used then).
 * fn == NULL removes the provider. */
typedef int (*nnsb_output_provider_fn) (void *user, size_t size, nnsb_memory * out);
NNSB_API int nnsb_transform_set_output_provider (nnsb_transform * t, nnsb_output_provider_fn fn, void *user);

typedef struct nnsb_converter nnsb_converter;

/* tensor_converter, video branch only (gsttensor_converter.c:1046-1109,1390-1519,2101-2239).
 * Properties: "frames-per-tensor" (uint >= 1, default 1), "silent". */
NNSB_API int nnsb_converter_new (int device_id, nnsb_converter ** out);
NNSB_API int nnsb_converter_free (nnsb_converter * c);
NNSB_API int nnsb_converter_set_property (nnsb_converter * c, const char *name, const char *value);
NNSB_API int nnsb_converter_get_property (nnsb_converter * c, const char *name, char *buf, size_t buflen);
/* gst_tensor_converter_parse_caps / parse_video: video/x-raw caps (format GRAY8, RGB, BGR,
 * RGBx, BGRx, xRGB, xBGR, RGBA, BGRA, ARGB, ABGR) -> tensors config; out = src caps string */
NNSB_API int nnsb_converter_set_caps (nnsb_converter * c, const char *video_caps, char *out, size_t outlen);
NNSB_API int nnsb_converter_get_config (nnsb_converter * c, nnsb_tensors_config * cfg);
/* 1 when rows carry padding that chain() removes (gst_tensor_converter_video_stride, :1390) */
NNSB_API int nnsb_converter_remove_padding (nnsb_converter * c);
NNSB_API size_t nnsb_converter_frame_size (nnsb_converter * c);   /* GST_VIDEO_INFO_SIZE */
/* gst_tensor_converter_chain, _NNS_VIDEO case: takes one video frame (host or device).  The
 * row padding is stripped on the way into a device-side batch buffer (pitched H2D copy for
 * host frames, strip kernel for device frames); when frames-per-tensor frames have arrived
 * *produced = 1 and `out` holds the batched tensor in device memory (release it). */
NNSB_API int nnsb_converter_chain (nnsb_converter * c, const nnsb_memory * frame, nnsb_memory * out,
    int *produced);

/* ---------------------------------------------------------------------------------------
 * Batching queue: groups frames so that one launch covers `batch` tensors (north_star).
 * Frames are copied (H2D on a copy stream, or D2D) into one of two device batch buffers;
 * nnsb_queue_flush launches the plan once over everything queued and hands back the device
 * pointer of the outputs (valid until the flush after next).  Per-tensor semantics are
 * those of nnsb_plan_launch.
 * ------------------------------------------------------------------------------------- */
typedef struct nnsb_queue nnsb_queue;
NNSB_API int nnsb_queue_create (nnsb_ctx * ctx, nnsb_plan * plan, uint32_t batch, nnsb_queue ** out);
NNSB_API int nnsb_queue_destroy (nnsb_queue * q);
/* *full = 1 when the queue now holds `batch` frames */
NNSB_API int nnsb_queue_push (nnsb_queue * q, const void *frame, int on_device, int *full);
NNSB_API int nnsb_queue_flush (nnsb_queue * q, void **d_out, uint32_t *n_frames, nnsb_stream ** stream);

/* ---------------------------------------------------------------------------------------
 * tensor_aggregator mirror: GstTensorAggregator,
 * gst/nnstreamer/elements/gsttensor_aggregator.c.  The GstAdapter becomes a byte range of one
 * device buffer; the concatenation along an inner axis (:768-787) is one permutation launch.
 * Properties (:170-240): frames-in, frames-out, frames-flush, frames-dim, concat, silent.
 * ------------------------------------------------------------------------------------- */
typedef struct nnsb_aggregator nnsb_aggregator;
#define NNSB_CLOCK_TIME_NONE 0xffffffffffffffffull      /* GST_CLOCK_TIME_NONE */
typedef struct
{
  uint64_t pts, dts, duration;  /* GST_BUFFER_PTS / _DTS / _DURATION, nanoseconds */
} nnsb_buffer_time;
NNSB_API int nnsb_aggregator_new (int device_id, nnsb_aggregator ** out);
NNSB_API int nnsb_aggregator_free (nnsb_aggregator * a);
/* gst_tensor_aggregator_set_property / _get_property, :313-382 */
NNSB_API int nnsb_aggregator_set_property (nnsb_aggregator * a, const char *name, const char *value);
NNSB_API int nnsb_aggregator_get_property (nnsb_aggregator * a, const char *name, char *buf,
    size_t buflen);
/* gst_tensor_aggregator_parse_caps, :1026-1081: sink caps in, source caps out */
NNSB_API int nnsb_aggregator_set_caps (nnsb_aggregator * a, const char *incaps, char *outcaps,
    size_t outlen);
NNSB_API int nnsb_aggregator_get_config (nnsb_aggregator * a, int direction, nnsb_tensors_config * cfg);
/* gst_tensor_aggregator_reset, :979-990 (READY->PAUSED :958, flush-stop :420-422) */
NNSB_API int nnsb_aggregator_reset (nnsb_aggregator * a);
/* number of output buffers the next chain call with in_size bytes will produce */
NNSB_API int nnsb_aggregator_outputs_for (nnsb_aggregator * a, size_t in_size);
/* gst_tensor_aggregator_chain, :835-941.  `in` is host or device memory and may be reused on
 * return.  Every output is a device memory (owned, release with nnsb_memory_release; when
 * frames-in == frames-out and nothing is concatenated a device input is passed on unowned).
 * t_in / t_outs may be NULL.  Returns a NNSB_FLOW_* code. */
NNSB_API int nnsb_aggregator_chain (nnsb_aggregator * a, const nnsb_memory * in,
    const nnsb_buffer_time * t_in, nnsb_memory * outs, nnsb_buffer_time * t_outs, uint32_t max_out,
    uint32_t *n_out);
NNSB_API int nnsb_aggregator_kernel_launches (nnsb_aggregator * a, uint64_t *count);
/* The memory the next input buffer should be written into: a share of the adapter's ring at its
 * tail.  A producer that fills it and passes it to nnsb_aggregator_chain has pushed the buffer
 * without a copy (gst_adapter_push, :873, normally memcpy's into the adapter).  This is the role
 * of a downstream-proposed buffer pool in GStreamer; the signature is nnsb_output_provider_fn so
 * that nnsb_transform_set_output_provider (t, nnsb_aggregator_acquire, a) wires an upstream
 * tensor_transform to it.  Fails with NNSB_ENOTSUP when frames-in == frames-out. */
NNSB_API int nnsb_aggregator_acquire (void *aggregator, size_t size, nnsb_memory * out);
/* bytes the element has moved itself so far: into the adapter / out of it (concatenation) */
NNSB_API int nnsb_aggregator_copy_counters (nnsb_aggregator * a, uint64_t *copied_in, uint64_t *copied_out);

/* ---------------------------------------------------------------------------------------
 * tensor_decoder mode=direct_video mirror: the GstTensorDecoderDef
 * callbacks of ext/nnstreamer/tensor_decoder/tensordec-directvideo.c.
 * ------------------------------------------------------------------------------------- */
typedef struct nnsb_direct_video nnsb_direct_video;
NNSB_API int nnsb_direct_video_new (int device_id, nnsb_direct_video ** out);      /* dv_init :102 */
NNSB_API int nnsb_direct_video_free (nnsb_direct_video * dv);                       /* dv_exit :118 */
/* dv_setOption, :126-158: option 0 (= "option1") is the video format name; returns 1 (TRUE)
 * or 0 (FALSE) like the callback */
NNSB_API int nnsb_direct_video_set_option (nnsb_direct_video * dv, int op_num, const char *param);
/* dv_getOutCaps, :161-262: video/x-raw caps for the first tensor of `config` */
NNSB_API int nnsb_direct_video_get_out_caps (nnsb_direct_video * dv, const nnsb_tensors_config * config,
    char *caps, size_t capslen);
/* dv_getTransformSize / _get_video_xraw_bufsize, :265-287 */
NNSB_API size_t nnsb_direct_video_transform_size (const nnsb_tensors_config * config);
/* dv_decode, :290-349.  `in` = the first tensor (uint8, host or device).  If out->data is set
 * and out->size equals the video buffer size that memory is filled (host: pitched D2H, device:
 * pad kernel), else a device memory is allocated (owned).  Returns a NNSB_FLOW_* code. */
NNSB_API int nnsb_direct_video_decode (nnsb_direct_video * dv, const nnsb_tensors_config * config,
    const nnsb_memory * in, nnsb_memory * out);

/* ---------------------------------------------------------------------------------------
 * tensor_merge / tensor_split data paths.
 * ------------------------------------------------------------------------------------- */
typedef struct nnsb_merge nnsb_merge;
NNSB_API int nnsb_merge_new (int device_id, nnsb_merge ** out);
NNSB_API int nnsb_merge_free (nnsb_merge * m);
/* properties of gsttensor_merge.c:159-185: "mode" (linear), "option" ("0".."3" = merge axis),
 * "silent", "sync-mode", "sync-option" (stored; the time synchronisation is not mirrored) */
NNSB_API int nnsb_merge_set_property (nnsb_merge * m, const char *name, const char *value);
NNSB_API int nnsb_merge_get_property (nnsb_merge * m, const char *name, char *buf, size_t buflen);
/* gst_tensor_merge_get_merged_config, :380-427: the fixed caps of the n sink pads in, the
 * source caps out */
NNSB_API int nnsb_merge_set_caps (nnsb_merge * m, uint32_t n, const char *const *caps, char *out,
    size_t outlen);
NNSB_API int nnsb_merge_get_config (nnsb_merge * m, nnsb_tensors_config * cfg);
/* gst_tensor_merge_generate_mem, :463-600: one memory per sink pad (host or device) in, the
 * merged tensor (device, owned) out.  Returns a NNSB_FLOW_* code. */
NNSB_API int nnsb_merge_chain (nnsb_merge * m, const nnsb_memory * in, uint32_t n, nnsb_memory * out);

typedef struct nnsb_split nnsb_split;
NNSB_API int nnsb_split_new (int device_id, nnsb_split ** out);
NNSB_API int nnsb_split_free (nnsb_split * s);
/* gst_tensor_split_set_property / _get_property, gsttensor_split.c:566-700: "tensorseg"
 * ("2:100:100,1:100:100"), "tensorpick" ("0,2"), "silent" */
NNSB_API int nnsb_split_set_property (nnsb_split * s, const char *name, const char *value);
NNSB_API int nnsb_split_get_property (nnsb_split * s, const char *name, char *buf, size_t buflen);
NNSB_API int nnsb_split_set_caps (nnsb_split * s, const char *caps);
NNSB_API int nnsb_split_num_outputs (nnsb_split * s);
/* caps of source pad src_<idx> (gst_tensor_split_get_tensor_pad, :345-361) */
NNSB_API int nnsb_split_get_out_caps (nnsb_split * s, uint32_t idx, char *out, size_t outlen);
/* gst_tensor_split_chain, :455-530: one device memory (owned) per picked segment */
NNSB_API int nnsb_split_chain (nnsb_split * s, const nnsb_memory * in, nnsb_memory * outs,
    uint32_t max_out, uint32_t *n_out);

/* ---------------------------------------------------------------------------------------
 * Hand-off to tensor_filter without leaving HBM, and tensor_mux /
 * tensor_demux over reference-counted device memories (rank 4).
 * ------------------------------------------------------------------------------------- */
#define NNSB_FLOW_DROPPED 100           /* GST_BASE_TRANSFORM_FLOW_DROPPED (GST_FLOW_CUSTOM_SUCCESS) */
/* GstTensorMemory, tensor_typedef.h:219-223 -- `data` is a DEVICE pointer here */
typedef struct {
  void *data;
  size_t size;
} nnsb_tensor_memory;
/* The v1 `invoke` of a tensor_filter sub-plugin (GstTensorFilterFramework,
 * nnstreamer_plugin_api_filter.h:419-421) for a CUDA-aware framework: inputs and pre-allocated
 * outputs are device memory, and the callback enqueues its work on `cuda_stream` (a
 * cudaStream_t) WITHOUT synchronising -- the element has already made that stream wait for the
 * producers of the inputs, and marks the outputs ready after whatever the callback queued.
 * Returns 0 = ok, > 0 = drop this buffer, < 0 = error (tensor_filter.c:806-815). */
typedef int (*nnsb_filter_invoke_fn) (void *private_data, const nnsb_tensor_memory * input,
    uint32_t n_in, nnsb_tensor_memory * output, uint32_t n_out, void *cuda_stream);
typedef struct nnsb_filter nnsb_filter;
NNSB_API int nnsb_filter_new (int device_id, nnsb_filter ** out);
NNSB_API int nnsb_filter_free (nnsb_filter * f);
/* the framework's callback + its private_data (gst_tensor_filter_common_open_fw) */
NNSB_API int nnsb_filter_set_framework (nnsb_filter * f, nnsb_filter_invoke_fn invoke, void *private_data);
/* prop->input_meta / output_meta (getInputDimension / getOutputDimension of the model) */
NNSB_API int nnsb_filter_set_info (nnsb_filter * f, const nnsb_tensors_config * in,
    const nnsb_tensors_config * out);
/* gst_tensor_filter_transform, tensor_filter.c:632-890: size checks (:719-735), output memories
 * from the device pool (:737-757), invoke (:786).  Device inputs are used in place (zero bytes
 * copied); host inputs are uploaded once (counted).  Outputs are stream-ordered device memories.
 * Returns NNSB_FLOW_OK, NNSB_FLOW_DROPPED or NNSB_FLOW_ERROR. */
NNSB_API int nnsb_filter_transform (nnsb_filter * f, const nnsb_memory * in, uint32_t n_in,
    nnsb_memory * out, uint32_t *n_out);
/* bytes this element has copied between host and device so far, and the invoke count */
NNSB_API int nnsb_filter_copy_counters (nnsb_filter * f, uint64_t *h2d_bytes, uint64_t *d2h_bytes,
    uint64_t *invokes);

typedef struct nnsb_mux nnsb_mux;
NNSB_API int nnsb_mux_new (nnsb_mux ** out);
NNSB_API int nnsb_mux_free (nnsb_mux * m);
/* gsttensor_mux.c:156-170: "silent", "sync-mode" (nosync|slowest|basepad|refresh), "sync-option"
 * (stored; the collect-pads time synchronisation is not mirrored) */
NNSB_API int nnsb_mux_set_property (nnsb_mux * m, const char *name, const char *value);
NNSB_API int nnsb_mux_get_property (nnsb_mux * m, const char *name, char *buf, size_t buflen);
/* fixed caps of the n sink pads in pad order -> other/tensors source caps (:390-412) */
NNSB_API int nnsb_mux_set_caps (nnsb_mux * m, uint32_t n, const char *const *caps, char *out, size_t outlen);
NNSB_API int nnsb_mux_get_config (nnsb_mux * m, nnsb_tensors_config * cfg);
/* gst_tensor_mux_collected, :467,482: the memories of one collected set, in pad order, become
 * the memories of one buffer -- new references, nothing is copied */
NNSB_API int nnsb_mux_chain (nnsb_mux * m, const nnsb_memory * in, uint32_t n_in, nnsb_memory * out,
    uint32_t *n_out);

typedef struct nnsb_demux nnsb_demux;
NNSB_API int nnsb_demux_new (nnsb_demux ** out);
NNSB_API int nnsb_demux_free (nnsb_demux * d);
/* gsttensor_demux.c:144-153,590-626: "silent", "tensorpick" ("0,1:2,2+0": groups split at
 * ",.;/", indices inside a group at ":+") */
NNSB_API int nnsb_demux_set_property (nnsb_demux * d, const char *name, const char *value);
NNSB_API int nnsb_demux_get_property (nnsb_demux * d, const char *name, char *buf, size_t buflen);
NNSB_API int nnsb_demux_set_caps (nnsb_demux * d, const char *caps);
NNSB_API int nnsb_demux_num_outputs (nnsb_demux * d);
/* caps of source pad src_<idx> (gst_tensor_demux_get_tensor_config, :288-330) */
NNSB_API int nnsb_demux_get_out_caps (nnsb_demux * d, uint32_t idx, char *out, size_t outlen);
/* gst_tensor_demux_chain, :468-560: source pad i receives counts[i] memories, stored back to
 * back in outs[] -- references to the input blocks (:516-522); one blob holding several static
 * tensors is cut into shares first (nnstreamer_plugin_api_impl.c:534). */
NNSB_API int nnsb_demux_chain (nnsb_demux * d, const nnsb_memory * in, uint32_t n_in, nnsb_memory * outs,
    uint32_t max_out, uint32_t *counts, uint32_t max_src, uint32_t *n_src);

/* A linked run of elements pushed by ONE call: tensor_converter (optional) followed by n_t
 * tensor_transform instances.  This is GStreamer's push model -- the converter's chain function
 * pushes its buffer into the next element's chain function in the same streaming thread
 * (gst_pad_push; gsttensor_converter.c:1250, GstBaseTransform's default chain around
 * gsttensor_transform.c:1667) -- so a caller that pays for every call into the library (Python,
 * a JNI binding) pays once per frame, not once per element.  The elements stay owned by the
 * caller and must outlive the branch; they must be negotiated already. */
typedef struct nnsb_branch nnsb_branch;
NNSB_API int nnsb_branch_new (nnsb_converter * conv, nnsb_transform * const *t, uint32_t n_t,
    nnsb_branch ** out);
NNSB_API int nnsb_branch_free (nnsb_branch * b);
/* in: one video frame (conv != NULL) or the n_in memories of one tensor buffer.  outs receives the
 * *n_out memories the last element produced (0 while the converter is still collecting
 * frames-per-tensor frames).  Intermediate memories are released before the call returns;
 * nothing waits for the GPU.  Returns NNSB_FLOW_OK or NNSB_FLOW_ERROR. */
NNSB_API int nnsb_branch_push (nnsb_branch * b, const nnsb_memory * in, uint32_t n_in,
    nnsb_memory * outs, uint32_t *n_out);

/* No counterpart in the reference (its CPU element simply mallocs).  The same stand kernels
 * as nnsb_plan_launch(), for callers that must not allocate while frames are flowing: a
 * pipeline whose GPU memory is already fully reserved by an inference runtime's pool, a
 * streaming thread that cannot tolerate the unbounded latency of cudaMalloc, or a plug-in
 * that is handed only a CUDA stream and device pointers.  No cudaMalloc / cudaFree / pool is
 * touched: the scratch is a fixed NNSB_NOALLOC_SCRATCH_BYTES array inside the library's device
 * image (a GPU holds that much memory from the moment the image is loaded on it until exit).
 *   d_in, d_out : `batch` tensors of shape `dim` back to back, on the CURRENT CUDA device
 *   dim         : NNSB_RANK_LIMIT (8) entries, innermost first, unused ones set to 1 -- all
 *                 eight are read
 *   out_type    : NNSB_TYPE_END keeps the input type (the option's default)
 *   batch       : 1..65535 (NNSB_ERANGE above that)
 *   cuda_stream : a cudaStream_t, NULL = default stream
 * Fails with NNSB_ERANGE when batch x channels needs more scratch than the array holds.
 * WARM-UP: with CUDA's default lazy module loading the array and each kernel variant (input
 * type x output type x mode x channel-count class) are put on the device by the first call that
 * needs them, and that load does reserve device memory and take driver locks.  Make one call
 * per device and per variant at initialisation -- before the other runtime takes the rest of
 * the memory and before any stream capture (the first call on a device also resolves the
 * array's device address) -- or run with CUDA_MODULE_LOADING=EAGER.  Later calls allocate
 * nothing.  All calls on a device share the array, so calls that may overlap on the GPU race
 * with each other; keep them on one stream. */
#define NNSB_NOALLOC_SCRATCH_BYTES (1u << 20)
NNSB_API int nnsb_stand_launch_noalloc (const void *d_in, void *d_out, int in_type,
    int out_type, const uint32_t *dim, int stand_mode, int per_channel, uint32_t batch,
    void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* NNSTREAMER_B200_H */
