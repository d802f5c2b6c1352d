"""ctypes binding of libnnstreamer_b200.so (the C-ABI of include/nnstreamer_b200.h).

The library is the product: there is no Python or CPU implementation behind it.  Loading
fails loudly if the shared object is missing (run ``python -c 'import __graft_entry__ as g;
g.build()'``), and every compute entry point fails with NNSB_ECUDA when no B200 is visible.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libnnstreamer_b200.so")

RANK_LIMIT = 8
SIZE_LIMIT = 16
META_RANK_LIMIT = 16
MAX_OPS = 64
META_HEADER_SIZE = 128
NOALLOC_SCRATCH_BYTES = 1 << 20

# tensor_typedef.h:133-148
TYPES = ["int32", "uint32", "int16", "uint16", "int8", "uint8", "float64", "float32",
         "int64", "uint64", "float16"]
TYPE_ID = {n: i for i, n in enumerate(TYPES)}
TYPE_END = len(TYPES)
MODES = ["dimchg", "typecast", "arithmetic", "transpose", "stand", "clamp"]
MODE_ID = {n: i for i, n in enumerate(MODES)}
PAD_SRC, PAD_SINK = 1, 2

OK, EINVAL, ENOMEM, ENOTSUP, ECUDA, ERANGE = 0, -22, -12, -95, -5, -34


class TensorInfo(C.Structure):
    _fields_ = [("type", C.c_int32), ("dim", C.c_uint32 * RANK_LIMIT)]


class TensorsConfig(C.Structure):
    _fields_ = [("num_tensors", C.c_uint32), ("info", TensorInfo * SIZE_LIMIT),
                ("format", C.c_int32), ("rate_n", C.c_int32), ("rate_d", C.c_int32)]


class MetaInfo(C.Structure):
    _fields_ = [("version", C.c_uint32), ("type", C.c_uint32),
                ("dimension", C.c_uint32 * META_RANK_LIMIT), ("format", C.c_uint32),
                ("media_type", C.c_uint32), ("sparse_nnz", C.c_uint32)]


class ArithOp(C.Structure):
    _fields_ = [("op", C.c_int32), ("applying_ch", C.c_int32), ("is_float", C.c_int32),
                ("cast_type", C.c_int32), ("dval", C.c_double), ("ival", C.c_int64)]


class Option(C.Structure):
    _fields_ = [("mode", C.c_int32), ("dimchg_from", C.c_int32), ("dimchg_to", C.c_int32),
                ("typecast_to", C.c_int32), ("arith_out_type", C.c_int32),
                ("arith_per_channel", C.c_int32), ("arith_ch_dim", C.c_uint32),
                ("n_ops", C.c_int32), ("ops", ArithOp * MAX_OPS),
                ("trans_order", C.c_uint8 * RANK_LIMIT), ("stand_mode", C.c_int32),
                ("stand_out_type", C.c_int32), ("stand_per_channel", C.c_int32),
                ("clamp_min", C.c_double), ("clamp_max", C.c_double)]


class Memory(C.Structure):
    _fields_ = [("data", C.c_void_p), ("size", C.c_size_t), ("on_device", C.c_int32),
                ("owned", C.c_int32), ("owner", C.c_void_p), ("block", C.c_void_p)]


class DeviceView(C.Structure):
    _fields_ = [("device_ptr", C.c_void_p), ("size", C.c_size_t), ("device_id", C.c_int32),
                ("reserved", C.c_int32), ("ready_event", C.c_void_p)]


class TensorMemory(C.Structure):
    _fields_ = [("data", C.c_void_p), ("size", C.c_size_t)]


FILTER_INVOKE = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(TensorMemory), C.c_uint32,
                            C.POINTER(TensorMemory), C.c_uint32, C.c_void_p)
FLOW_DROPPED = 100


class BufferTime(C.Structure):
    _fields_ = [("pts", C.c_uint64), ("dts", C.c_uint64), ("duration", C.c_uint64)]


CLOCK_TIME_NONE = 0xffffffffffffffff
FLOW_OK, FLOW_ERROR, FLOW_NOT_SUPPORTED, ENOENT = 0, -5, -6, -2


class NnsbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("nnstreamer_b200 error %d: %s" % (code, msg))
        self.code = code


_lib = None

_VP = C.c_void_p
_PROTOS = {
    # name: (restype, argtypes)
    "nnsb_last_error": (C.c_char_p, []),
    "nnsb_type_from_string": (C.c_int, [C.c_char_p]),
    "nnsb_type_to_string": (C.c_char_p, [C.c_int]),
    "nnsb_element_size": (C.c_size_t, [C.c_int]),
    "nnsb_element_count": (C.c_uint64, [C.POINTER(C.c_uint32)]),
    "nnsb_info_size": (C.c_size_t, [C.POINTER(TensorInfo)]),
    "nnsb_dim_parse": (C.c_int, [C.c_char_p, C.POINTER(C.c_uint32)]),
    "nnsb_dim_to_string": (C.c_int, [C.POINTER(C.c_uint32), C.c_uint, C.c_char_p, C.c_size_t]),
    "nnsb_dim_is_valid": (C.c_int, [C.POINTER(C.c_uint32)]),
    "nnsb_meta_from_info": (C.c_int, [C.POINTER(TensorInfo), C.POINTER(MetaInfo)]),
    "nnsb_meta_to_info": (C.c_int, [C.POINTER(MetaInfo), C.POINTER(TensorInfo)]),
    "nnsb_meta_write_header": (C.c_int, [C.POINTER(MetaInfo), _VP]),
    "nnsb_meta_parse_header": (C.c_int, [C.POINTER(MetaInfo), _VP]),
    "nnsb_meta_data_size": (C.c_size_t, [C.POINTER(MetaInfo)]),
    "nnsb_mode_from_string": (C.c_int, [C.c_char_p]),
    "nnsb_mode_to_string": (C.c_char_p, [C.c_int]),
    "nnsb_option_parse": (C.c_int, [C.c_int, C.c_char_p, C.POINTER(Option)]),
    "nnsb_convert_dimension": (C.c_int, [C.POINTER(Option), C.c_int, C.POINTER(TensorInfo),
                                         C.POINTER(TensorInfo)]),
    "nnsb_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "nnsb_ctx_create": (C.c_int, [C.c_int, C.POINTER(_VP)]),
    "nnsb_ctx_destroy": (C.c_int, [_VP]),
    "nnsb_ctx_device": (C.c_int, [_VP]),
    "nnsb_ctx_sm_count": (C.c_int, [_VP]),
    "nnsb_stream_create": (C.c_int, [_VP, C.POINTER(_VP)]),
    "nnsb_stream_wrap": (C.c_int, [_VP, _VP, C.POINTER(_VP)]),
    "nnsb_stream_sync": (C.c_int, [_VP]),
    "nnsb_stream_destroy": (C.c_int, [_VP]),
    "nnsb_event_create": (C.c_int, [_VP, C.POINTER(_VP)]),
    "nnsb_event_record": (C.c_int, [_VP, _VP]),
    "nnsb_event_sync": (C.c_int, [_VP]),
    "nnsb_stream_wait_event": (C.c_int, [_VP, _VP]),
    "nnsb_event_elapsed_ms": (C.c_int, [_VP, _VP, C.POINTER(C.c_float)]),
    "nnsb_event_destroy": (C.c_int, [_VP]),
    "nnsb_dev_alloc": (C.c_int, [_VP, C.c_size_t, C.POINTER(_VP)]),
    "nnsb_dev_free": (C.c_int, [_VP, _VP]),
    "nnsb_dev_trim": (C.c_int, [_VP]),
    "nnsb_dev_memset_async": (C.c_int, [_VP, _VP, C.c_int, C.c_size_t]),
    "nnsb_pinned_alloc": (C.c_int, [_VP, C.c_size_t, C.POINTER(_VP)]),
    "nnsb_pinned_free": (C.c_int, [_VP, _VP]),
    "nnsb_h2d_async": (C.c_int, [_VP, _VP, _VP, C.c_size_t]),
    "nnsb_d2h_async": (C.c_int, [_VP, _VP, _VP, C.c_size_t]),
    "nnsb_d2d_async": (C.c_int, [_VP, _VP, _VP, C.c_size_t]),
    "nnsb_h2d_2d_async": (C.c_int, [_VP, _VP, C.c_size_t, _VP, C.c_size_t, C.c_size_t, C.c_size_t]),
    "nnsb_plan_create": (C.c_int, [_VP, C.POINTER(Option), C.POINTER(TensorInfo), C.c_int,
                                   C.POINTER(_VP)]),
    "nnsb_plan_destroy": (C.c_int, [_VP]),
    "nnsb_plan_out_info": (C.c_int, [_VP, C.POINTER(TensorInfo)]),
    "nnsb_plan_in_size": (C.c_size_t, [_VP]),
    "nnsb_plan_out_size": (C.c_size_t, [_VP]),
    "nnsb_plan_launch": (C.c_int, [_VP, _VP, _VP, C.c_uint32, _VP]),
    "nnsb_plan_is_elementwise": (C.c_int, [_VP]),
    "nnsb_plan_chunk_granule": (C.c_uint64, [_VP]),
    "nnsb_plan_launch_elems": (C.c_int, [_VP, _VP, _VP, C.c_uint64, C.c_uint64, _VP]),
    "nnsb_plan_last_launch_count": (C.c_int, [_VP]),
    "nnsb_plan_describe": (C.c_int, [_VP, C.c_char_p, C.c_size_t]),
    "nnsb_plan_create_permute": (C.c_int, [_VP, C.c_int, C.c_int, C.POINTER(C.c_uint32),
                                           C.POINTER(C.c_uint64), C.POINTER(_VP)]),
    "nnsb_plan_stand_stats": (C.c_int, [_VP, _VP, C.c_uint32, C.POINTER(C.c_double),
                                        C.POINTER(C.c_double), C.c_size_t]),
    "nnsb_strip_launch": (C.c_int, [_VP, _VP, _VP, C.c_size_t, C.c_size_t, C.c_size_t]),
    "nnsb_transform_new": (C.c_int, [C.c_int, C.POINTER(_VP)]),
    "nnsb_transform_free": (C.c_int, [_VP]),
    "nnsb_transform_set_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p]),
    "nnsb_transform_get_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p, C.c_size_t]),
    "nnsb_transform_transform_caps": (C.c_int, [_VP, C.c_int, C.c_char_p, C.c_char_p, C.c_char_p,
                                                C.c_size_t]),
    "nnsb_transform_fixate_caps": (C.c_int, [_VP, C.c_int, C.c_char_p, C.c_char_p, C.c_char_p,
                                             C.c_size_t]),
    "nnsb_transform_set_caps": (C.c_int, [_VP, C.c_char_p, C.c_char_p]),
    "nnsb_transform_transform_size": (C.c_int, [_VP, C.c_int, C.c_char_p, C.c_size_t, C.c_char_p,
                                                C.POINTER(C.c_size_t)]),
    "nnsb_transform_transform": (C.c_int, [_VP, C.POINTER(Memory), C.c_uint32, C.POINTER(Memory),
                                           C.POINTER(C.c_uint32)]),
    "nnsb_transform_get_config": (C.c_int, [_VP, C.c_int, C.POINTER(TensorsConfig)]),
    "nnsb_transform_alloc": (C.c_int, [_VP, C.c_size_t, C.c_int, C.POINTER(Memory)]),
    "nnsb_memory_release": (C.c_int, [C.POINTER(Memory)]),
    "nnsb_memory_alloc": (C.c_int, [_VP, C.c_size_t, C.c_int, _VP, C.POINTER(Memory)]),
    "nnsb_memory_ref": (C.c_int, [C.POINTER(Memory), C.POINTER(Memory)]),
    "nnsb_memory_share": (C.c_int, [C.POINTER(Memory), C.c_size_t, C.c_size_t, C.POINTER(Memory)]),
    "nnsb_memory_wait": (C.c_int, [C.POINTER(Memory), _VP]),
    "nnsb_memory_sync": (C.c_int, [C.POINTER(Memory)]),
    "nnsb_memory_mark_io": (C.c_int, [_VP, C.POINTER(Memory), C.c_uint32, C.POINTER(Memory), C.c_uint32]),
    "nnsb_memory_mark_ready": (C.c_int, [C.POINTER(Memory), _VP]),
    "nnsb_memory_mark_use": (C.c_int, [C.POINTER(Memory), _VP]),
    "nnsb_memory_export": (C.c_int, [C.POINTER(Memory), C.POINTER(DeviceView)]),
    "nnsb_stream_handle": (_VP, [_VP]),
    "nnsb_dev_alloc_async": (C.c_int, [_VP, C.c_size_t, C.POINTER(_VP)]),
    "nnsb_dev_free_async": (C.c_int, [_VP, _VP]),
    "nnsb_ctx_device_mallocs": (C.c_uint64, [_VP]),
    "nnsb_transform_download": (C.c_int, [_VP, _VP, C.POINTER(Memory)]),
    "nnsb_transform_upload": (C.c_int, [_VP, C.POINTER(Memory), _VP]),
    "nnsb_transform_kernel_launches": (C.c_int, [_VP, C.POINTER(C.c_uint64)]),
    "nnsb_converter_new": (C.c_int, [C.c_int, C.POINTER(_VP)]),
    "nnsb_converter_free": (C.c_int, [_VP]),
    "nnsb_converter_set_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p]),
    "nnsb_converter_get_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p, C.c_size_t]),
    "nnsb_converter_set_caps": (C.c_int, [_VP, C.c_char_p, C.c_char_p, C.c_size_t]),
    "nnsb_converter_get_config": (C.c_int, [_VP, C.POINTER(TensorsConfig)]),
    "nnsb_converter_remove_padding": (C.c_int, [_VP]),
    "nnsb_converter_frame_size": (C.c_size_t, [_VP]),
    "nnsb_converter_chain": (C.c_int, [_VP, C.POINTER(Memory), C.POINTER(Memory), C.POINTER(C.c_int)]),
    "nnsb_aggregator_new": (C.c_int, [C.c_int, C.POINTER(_VP)]),
    "nnsb_aggregator_free": (C.c_int, [_VP]),
    "nnsb_aggregator_set_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p]),
    "nnsb_aggregator_get_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p, C.c_size_t]),
    "nnsb_aggregator_set_caps": (C.c_int, [_VP, C.c_char_p, C.c_char_p, C.c_size_t]),
    "nnsb_aggregator_get_config": (C.c_int, [_VP, C.c_int, C.POINTER(TensorsConfig)]),
    "nnsb_aggregator_reset": (C.c_int, [_VP]),
    "nnsb_aggregator_outputs_for": (C.c_int, [_VP, C.c_size_t]),
    "nnsb_aggregator_chain": (C.c_int, [_VP, C.POINTER(Memory), C.POINTER(BufferTime),
                                        C.POINTER(Memory), C.POINTER(BufferTime), C.c_uint32,
                                        C.POINTER(C.c_uint32)]),
    "nnsb_aggregator_kernel_launches": (C.c_int, [_VP, C.POINTER(C.c_uint64)]),
    "nnsb_aggregator_acquire": (C.c_int, [_VP, C.c_size_t, C.POINTER(Memory)]),
    "nnsb_aggregator_copy_counters": (C.c_int, [_VP, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "nnsb_transform_set_output_provider": (C.c_int, [_VP, _VP, _VP]),
    "nnsb_pad_launch": (C.c_int, [_VP, _VP, _VP, C.c_size_t, C.c_size_t, C.c_size_t]),
    "nnsb_d2h_2d_async": (C.c_int, [_VP, _VP, C.c_size_t, _VP, C.c_size_t, C.c_size_t, C.c_size_t]),
    "nnsb_direct_video_new": (C.c_int, [C.c_int, C.POINTER(_VP)]),
    "nnsb_direct_video_free": (C.c_int, [_VP]),
    "nnsb_direct_video_set_option": (C.c_int, [_VP, C.c_int, C.c_char_p]),
    "nnsb_direct_video_get_out_caps": (C.c_int, [_VP, C.POINTER(TensorsConfig), C.c_char_p,
                                                 C.c_size_t]),
    "nnsb_direct_video_transform_size": (C.c_size_t, [C.POINTER(TensorsConfig)]),
    "nnsb_direct_video_decode": (C.c_int, [_VP, C.POINTER(TensorsConfig), C.POINTER(Memory),
                                           C.POINTER(Memory)]),
    "nnsb_merge_launch": (C.c_int, [_VP, C.c_uint32, C.POINTER(_VP), C.POINTER(C.c_uint64), _VP,
                                    C.c_uint64]),
    "nnsb_merge_new": (C.c_int, [C.c_int, C.POINTER(_VP)]),
    "nnsb_merge_free": (C.c_int, [_VP]),
    "nnsb_merge_set_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p]),
    "nnsb_merge_get_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p, C.c_size_t]),
    "nnsb_merge_set_caps": (C.c_int, [_VP, C.c_uint32, C.POINTER(C.c_char_p), C.c_char_p, C.c_size_t]),
    "nnsb_merge_get_config": (C.c_int, [_VP, C.POINTER(TensorsConfig)]),
    "nnsb_merge_chain": (C.c_int, [_VP, C.POINTER(Memory), C.c_uint32, C.POINTER(Memory)]),
    "nnsb_split_new": (C.c_int, [C.c_int, C.POINTER(_VP)]),
    "nnsb_split_free": (C.c_int, [_VP]),
    "nnsb_split_set_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p]),
    "nnsb_split_get_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p, C.c_size_t]),
    "nnsb_split_set_caps": (C.c_int, [_VP, C.c_char_p]),
    "nnsb_split_num_outputs": (C.c_int, [_VP]),
    "nnsb_split_get_out_caps": (C.c_int, [_VP, C.c_uint32, C.c_char_p, C.c_size_t]),
    "nnsb_split_chain": (C.c_int, [_VP, C.POINTER(Memory), C.POINTER(Memory), C.c_uint32,
                                   C.POINTER(C.c_uint32)]),
    "nnsb_filter_new": (C.c_int, [C.c_int, C.POINTER(_VP)]),
    "nnsb_filter_free": (C.c_int, [_VP]),
    "nnsb_filter_set_framework": (C.c_int, [_VP, FILTER_INVOKE, _VP]),
    "nnsb_filter_set_info": (C.c_int, [_VP, C.POINTER(TensorsConfig), C.POINTER(TensorsConfig)]),
    "nnsb_filter_transform": (C.c_int, [_VP, C.POINTER(Memory), C.c_uint32, C.POINTER(Memory),
                                        C.POINTER(C.c_uint32)]),
    "nnsb_filter_copy_counters": (C.c_int, [_VP, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                            C.POINTER(C.c_uint64)]),
    "nnsb_mux_new": (C.c_int, [C.POINTER(_VP)]),
    "nnsb_mux_free": (C.c_int, [_VP]),
    "nnsb_mux_set_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p]),
    "nnsb_mux_get_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p, C.c_size_t]),
    "nnsb_mux_set_caps": (C.c_int, [_VP, C.c_uint32, C.POINTER(C.c_char_p), C.c_char_p, C.c_size_t]),
    "nnsb_mux_get_config": (C.c_int, [_VP, C.POINTER(TensorsConfig)]),
    "nnsb_mux_chain": (C.c_int, [_VP, C.POINTER(Memory), C.c_uint32, C.POINTER(Memory),
                                 C.POINTER(C.c_uint32)]),
    "nnsb_branch_new": (C.c_int, [_VP, C.POINTER(_VP), C.c_uint32, C.POINTER(_VP)]),
    "nnsb_branch_free": (C.c_int, [_VP]),
    "nnsb_branch_push": (C.c_int, [_VP, C.POINTER(Memory), C.c_uint32, C.POINTER(Memory),
                                   C.POINTER(C.c_uint32)]),
    "nnsb_demux_new": (C.c_int, [C.POINTER(_VP)]),
    "nnsb_demux_free": (C.c_int, [_VP]),
    "nnsb_demux_set_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p]),
    "nnsb_demux_get_property": (C.c_int, [_VP, C.c_char_p, C.c_char_p, C.c_size_t]),
    "nnsb_demux_set_caps": (C.c_int, [_VP, C.c_char_p]),
    "nnsb_demux_num_outputs": (C.c_int, [_VP]),
    "nnsb_demux_get_out_caps": (C.c_int, [_VP, C.c_uint32, C.c_char_p, C.c_size_t]),
    "nnsb_demux_chain": (C.c_int, [_VP, C.POINTER(Memory), C.c_uint32, C.POINTER(Memory), C.c_uint32,
                                   C.POINTER(C.c_uint32), C.c_uint32, C.POINTER(C.c_uint32)]),
    "nnsb_queue_create": (C.c_int, [_VP, _VP, C.c_uint32, C.POINTER(_VP)]),
    "nnsb_queue_destroy": (C.c_int, [_VP]),
    "nnsb_queue_push": (C.c_int, [_VP, _VP, C.c_int, C.POINTER(C.c_int)]),
    "nnsb_queue_flush": (C.c_int, [_VP, C.POINTER(_VP), C.POINTER(C.c_uint32), C.POINTER(_VP)]),
    "nnsb_stand_launch_noalloc": (C.c_int, [_VP, _VP, C.c_int, C.c_int, C.POINTER(C.c_uint32),
                                            C.c_int, C.c_int, C.c_uint32, _VP]),
}


def lib():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise ImportError(
                "%s is missing: the CUDA extension has not been built "
                "(python -c 'import __graft_entry__ as g; g.build()'). There is no fallback."
                % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def declared_symbols():
    return sorted(_PROTOS)


def check(code):
    if code != OK:
        raise NnsbError(code, lib().nnsb_last_error().decode("utf-8", "replace"))
    return code


def make_info(type_name, dim):
    info = TensorInfo()
    info.type = TYPE_ID[type_name] if isinstance(type_name, str) else int(type_name)
    d = list(dim) + [1] * (RANK_LIMIT - len(dim))
    for i in range(RANK_LIMIT):
        info.dim[i] = int(d[i])
    return info


def parse_option(mode, option):
    """nnsb_option_parse; returns (code, Option)."""
    opt = Option()
    m = MODE_ID[mode] if isinstance(mode, str) else int(mode)
    code = lib().nnsb_option_parse(m, option.encode() if option is not None else None, C.byref(opt))
    return code, opt
