/*
 * oracle/ref_wrap.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C entry points around the reference's OWN functions, which build_ref.py pastes
 * above this file inside one translation unit (they are `static` in the reference):
 *   gsttensor_transform.c:1120 dimchg, :1202 typecast, :1240 arithmetic, :1443 transpose,
 *   :1512 stand, :1632 clamp, :1850 convert_dimension; tensor_data.c is linked unmodified.
 * This file only fills a GstTensorTransform the way set_option_data (:652-958) would and
 * calls them.  Written C89-style because the reference is built with -std=gnu89.
 */

static const guint ref_element_size[] = { 4, 4, 2, 2, 1, 1, 8, 4, 8, 8, 2, 0 };

/* nnstreamer_plugin_api_util_impl.c:1069-1080 */
gulong
gst_tensor_get_element_count (const tensor_dim dim)
{
  gulong count = 1;
  guint i;
  for (i = 0; i < NNS_TENSOR_RANK_LIMIT; i++)
    count *= dim[i];
  return count;
}

/* nnstreamer_plugin_api_util_impl.c:1085-1091 */
gsize
gst_tensor_get_element_size (tensor_type type)
{
  if ((int) type < 0 || type > _NNS_END)
    return 0;
  return ref_element_size[type];
}

/* nnstreamer_plugin_api_util_impl.c:155-166 */
gsize
gst_tensor_info_get_size (const GstTensorInfo * info)
{
  return gst_tensor_get_element_count (info->dimension) *
      gst_tensor_get_element_size (info->type);
}

typedef struct
{
  int op;                       /* tensor_transform_operator: 0 typecast 1 add 2 mul 3 div */
  int applying_ch;              /* -1: all channels */
  int is_float;                 /* operand literal had '.', 'e' or 'E' (:796-807) */
  int cast_type;                /* typecast target for op 0 */
  double dval;
  int64_t ival;
} ref_op_t;

static void
ref_fill_info (GstTensorInfo * info, int type, const uint32_t * dim)
{
  guint i;
  info->name = NULL;
  info->type = (tensor_type) type;
  for (i = 0; i < NNS_TENSOR_RANK_LIMIT; i++)
    info->dimension[i] = dim[i];
}

static void
ref_filter_init (GstTensorTransform * f, tensor_transform_mode mode)
{
  memset (f, 0, sizeof (*f));
  f->silent = TRUE;
  f->mode = mode;
  f->loaded = TRUE;
  f->acceleration = FALSE;
  f->operators = NULL;
  f->apply = NULL;
}

int
ref_convert_dimension (int mode, int direction_sink, int from, int to,
    const uint8_t * order, const uint32_t * in_dim, uint32_t * out_dim)
{
  GstTensorTransform f;
  GstTensorInfo in_info, out_info;
  guint i;
  gboolean ok;

  ref_filter_init (&f, (tensor_transform_mode) mode);
  if (mode == GTT_DIMCHG) {
    f.data_dimchg.from = from;
    f.data_dimchg.to = to;
  } else if (mode == GTT_TRANSPOSE) {
    for (i = 0; i < NNS_TENSOR_TRANSPOSE_RANK_LIMIT; i++)
      f.data_transpose.trans_order[i] = order[i];
  }
  ref_fill_info (&in_info, _NNS_UINT8, in_dim);
  ok = gst_tensor_transform_convert_dimension (&f,
      direction_sink ? GST_PAD_SINK : GST_PAD_SRC, 0, &in_info, &out_info);
  for (i = 0; i < NNS_TENSOR_RANK_LIMIT; i++)
    out_dim[i] = out_info.dimension[i];
  return ok ? 0 : -1;
}

int
ref_typecast (const void *in, void *out, int in_type, int out_type,
    const uint32_t * dim)
{
  GstTensorTransform f;
  GstTensorInfo in_info, out_info;

  ref_filter_init (&f, GTT_TYPECAST);
  f.data_typecast.to = (tensor_type) out_type;
  ref_fill_info (&in_info, in_type, dim);
  ref_fill_info (&out_info, out_type, dim);
  return (int) gst_tensor_transform_typecast (&f, &in_info, &out_info,
      (const uint8_t *) in, (uint8_t *) out);
}

int
ref_arithmetic (const void *in, void *out, int in_type, int out_type,
    const uint32_t * dim, int per_channel, unsigned int ch_dim, int n_ops,
    const ref_op_t * ops)
{
  GstTensorTransform f;
  GstTensorInfo in_info, out_info;
  tensor_transform_operator_s *op_s;
  GSList *nodes;
  int i, ret;

  if (n_ops <= 0)
    return -1;
  ref_filter_init (&f, GTT_ARITHMETIC);
  f.data_arithmetic.out_type = _NNS_END;
  f.data_arithmetic.per_channel_arith = per_channel;
  f.data_arithmetic.ch_dim = ch_dim;

  op_s = (tensor_transform_operator_s *) calloc (n_ops, sizeof (*op_s));
  nodes = (GSList *) calloc (n_ops, sizeof (*nodes));
  for (i = 0; i < n_ops; i++) {
    op_s[i].op = (tensor_transform_operator) ops[i].op;
    op_s[i].applying_ch = ops[i].applying_ch;
    if (ops[i].op == GTT_OP_TYPECAST) {
      /* :781-785 */
      op_s[i].value.type = (tensor_type) ops[i].cast_type;
      f.data_arithmetic.out_type = op_s[i].value.type;
    } else if (ops[i].is_float) {
      /* :796-801 */
      double val = ops[i].dval;
      gst_tensor_data_set (&op_s[i].value, _NNS_FLOAT64, &val);
    } else {
      /* :802-807 */
      int64_t val = ops[i].ival;
      gst_tensor_data_set (&op_s[i].value, _NNS_INT64, &val);
    }
    nodes[i].data = &op_s[i];
    nodes[i].next = (i + 1 < n_ops) ? &nodes[i + 1] : NULL;
  }
  f.operators = nodes;

  ref_fill_info (&in_info, in_type, dim);
  ref_fill_info (&out_info, out_type, dim);
  ret = (int) gst_tensor_transform_arithmetic (&f, &in_info, &out_info,
      (const uint8_t *) in, (uint8_t *) out);
  free (op_s);
  free (nodes);
  return ret;
}

int
ref_transpose (const void *in, void *out, int type, const uint32_t * in_dim,
    const uint8_t * order)
{
  GstTensorTransform f;
  GstTensorInfo in_info, out_info;
  guint i;

  ref_filter_init (&f, GTT_TRANSPOSE);
  for (i = 0; i < NNS_TENSOR_TRANSPOSE_RANK_LIMIT; i++)
    f.data_transpose.trans_order[i] = order[i];
  ref_fill_info (&in_info, type, in_dim);
  gst_tensor_transform_convert_dimension (&f, GST_PAD_SINK, 0, &in_info,
      &out_info);
  return (int) gst_tensor_transform_transpose (&f, &in_info, &out_info,
      (const uint8_t *) in, (uint8_t *) out);
}

int
ref_dimchg (const void *in, void *out, int type, const uint32_t * in_dim,
    int from, int to)
{
  GstTensorTransform f;
  GstTensorInfo in_info, out_info;

  ref_filter_init (&f, GTT_DIMCHG);
  f.data_dimchg.from = from;
  f.data_dimchg.to = to;
  ref_fill_info (&in_info, type, in_dim);
  gst_tensor_transform_convert_dimension (&f, GST_PAD_SINK, 0, &in_info,
      &out_info);
  return (int) gst_tensor_transform_dimchg (&f, &in_info, &out_info,
      (const uint8_t *) in, (uint8_t *) out);
}

int
ref_stand (const void *in, void *out, int in_type, int out_type,
    const uint32_t * dim, int mode, int per_channel)
{
  GstTensorTransform f;
  GstTensorInfo in_info, out_info;

  ref_filter_init (&f, GTT_STAND);
  f.data_stand.mode = (tensor_transform_stand_mode) mode;
  f.data_stand.out_type = (tensor_type) out_type;
  f.data_stand.per_channel = per_channel;
  ref_fill_info (&in_info, in_type, dim);
  ref_fill_info (&out_info, out_type, dim);
  return (int) gst_tensor_transform_stand (&f, &in_info, &out_info,
      (const uint8_t *) in, (uint8_t *) out);
}

int
ref_clamp (const void *in, void *out, int type, const uint32_t * dim,
    double min, double max)
{
  GstTensorTransform f;
  GstTensorInfo in_info, out_info;

  ref_filter_init (&f, GTT_CLAMP);
  f.data_clamp.min = min;
  f.data_clamp.max = max;
  ref_fill_info (&in_info, type, dim);
  ref_fill_info (&out_info, type, dim);
  return (int) gst_tensor_transform_clamp (&f, &in_info, &out_info,
      (const uint8_t *) in, (uint8_t *) out);
}

/* statistics helpers of tensor_data.c:314-493, exposed for the stand parity tests */
int
ref_average (const void *raw, size_t length, int type, double *result)
{
  gdouble *r = NULL;
  if (!gst_tensor_data_raw_average ((gpointer) raw, length, (tensor_type) type, &r))
    return -1;
  *result = *r;
  free (r);
  return 0;
}

int
ref_std (const void *raw, size_t length, int type, double average, double *result)
{
  gdouble *r = NULL;
  if (!gst_tensor_data_raw_std ((gpointer) raw, length, (tensor_type) type, &average, &r))
    return -1;
  *result = *r;
  free (r);
  return 0;
}

int
ref_has_float16 (void)
{
#ifdef FLOAT16_SUPPORT
  return 1;
#else
  return 0;
#endif
}
