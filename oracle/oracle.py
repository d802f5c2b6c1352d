"""oracle/oracle.py -- TEST INFRASTRUCTURE ONLY.

ctypes front-end of the two CPU checkers of the tensor_transform hot path:

  * ``port``  -- oracle/nns_oracle.c, our plain-C restatement of the reference algorithm
                 (both acceleration=false "C" semantics and acceleration=true "ORC" semantics);
  * ``ref``   -- oracle/_ref/libnns_ref.so, the reference's OWN functions compiled from
                 /root/reference by oracle/build_ref.py (acceleration=false only; built in
                 the development container, shipped prebuilt to the GPU box).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` leg
may import this module.  The product (nnstreamer_b200/) never does: it has no CPU path.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SRC = os.path.join(HERE, "nns_oracle.c")
PORT_SO = os.path.join(HERE, "libnns_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libnns_ref.so")

# tensor_typedef.h:133-148
TYPES = ["int32", "uint32", "int16", "uint16", "int8", "uint8", "float64", "float32",
         "int64", "uint64", "float16"]
TYPE_ID = {n: i for i, n in enumerate(TYPES)}
NP_DTYPE = {n: np.dtype(n) for n in TYPES}
OP = {"typecast": 0, "add": 1, "mul": 2, "div": 3}
RANK = 8


class Op(C.Structure):
    """Mirrors oracle_op_t / ref_op_t / nnsb_arith_op (identical layout on purpose)."""
    _fields_ = [("op", C.c_int32), ("applying_ch", C.c_int32), ("is_float", C.c_int32),
                ("cast_type", C.c_int32), ("dval", C.c_double), ("ival", C.c_int64)]


def build_port(force=False):
    """Compile oracle/nns_oracle.c -> oracle/libnns_oracle.so (gcc, no reference needed)."""
    if (not force and os.path.isfile(PORT_SO)
            and os.path.getmtime(PORT_SO) >= os.path.getmtime(PORT_SRC)):
        return PORT_SO
    cmd = ["gcc", "-O2", "-std=gnu11", "-ffp-contract=off", "-fPIC", "-shared", "-Wall",
           "-o", PORT_SO, PORT_SRC, "-lm"]
    subprocess.check_call(cmd)
    return PORT_SO


def build_ref(force=False):
    """Compile the reference's own code when /root/reference is present; else keep prebuilt."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("nns_build_ref", os.path.join(HERE, "build_ref.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.build(force=force)


def dims8(dim):
    d = list(dim) + [1] * (RANK - len(dim))
    return (C.c_uint32 * RANK)(*d)


def make_ops(ops):
    """ops: list of (name, value[, channel]) / ("typecast", "float32").
    Numeric values follow the reference's literal rule: a Python float is a 'double literal',
    a Python int an 'int64 literal' (gsttensor_transform.c:796-807)."""
    arr = (Op * max(1, len(ops)))()
    for i, o in enumerate(ops):
        name = o[0]
        arr[i].op = OP[name]
        arr[i].applying_ch = -1
        arr[i].cast_type = len(TYPES)
        if name == "typecast":
            arr[i].cast_type = TYPE_ID[o[1]]
        else:
            v = o[1]
            if isinstance(v, float):
                arr[i].is_float = 1
                arr[i].dval = v
            else:
                arr[i].is_float = 0
                arr[i].ival = int(v)
            if len(o) > 2 and o[2] is not None:
                arr[i].applying_ch = int(o[2])
    return arr


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _count(dim):
    n = 1
    for d in dim:
        n *= int(d)
    return n


def convert_dims(mode, dim, **kw):
    """Output dimension of one tensor (innermost-first list of 8)."""
    d = list(dim) + [1] * (RANK - len(dim))
    if mode == "transpose":
        order = kw["order"]
        out = list(d)
        for i in range(4):
            out[i] = d[order[i]]
        return out
    if mode == "dimchg":
        f, t = kw["from_"], kw["to"]
        out = list(d)
        if f < t:
            out[f:t] = d[f + 1:t + 1]
            out[t] = d[f]
        elif f > t:
            out[t] = d[f]
            out[t + 1:f + 1] = d[t:f]
        return out
    return d


class Port:
    """The C restatement."""
    kind = "port"

    def __init__(self):
        self.lib = C.CDLL(build_port())
        L = self.lib
        L.oracle_typecast.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int]
        L.oracle_arithmetic.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                        C.POINTER(C.c_uint32), C.c_int, C.c_uint, C.c_int,
                                        C.POINTER(Op), C.c_int]
        L.oracle_transpose.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_uint32),
                                       C.POINTER(C.c_uint8)]
        L.oracle_dimchg.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_uint32),
                                    C.c_int, C.c_int]
        L.oracle_stand.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                   C.POINTER(C.c_uint32), C.c_int, C.c_int]
        L.oracle_stand_stats.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint32), C.c_int,
                                         C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.oracle_clamp.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_uint64, C.c_double,
                                   C.c_double]
        L.oracle_stride_strip.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]

        L.oracle_aggregator_concat.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32,
                                               C.c_uint64]
        L.oracle_direct_video.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
        L.oracle_merge_linear.argtypes = [C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_uint32),
                                          C.c_uint64, C.c_int, C.c_void_p]

    def typecast(self, x, out_type, accel=False):
        x = np.ascontiguousarray(x)
        out = np.empty(x.size, dtype=NP_DTYPE[out_type])
        self.lib.oracle_typecast(_ptr(x), _ptr(out), x.size, TYPE_ID[x.dtype.name],
                                 TYPE_ID[out_type], int(accel))
        return out

    def arithmetic(self, x, dim, ops, per_channel=False, ch_dim=0, accel=False):
        x = np.ascontiguousarray(x)
        out_type = x.dtype.name
        for o in ops:
            if o[0] == "typecast":
                out_type = o[1]
        out = np.empty(x.size, dtype=NP_DTYPE[out_type])
        r = self.lib.oracle_arithmetic(_ptr(x), _ptr(out), TYPE_ID[x.dtype.name],
                                       TYPE_ID[out_type], dims8(dim), int(per_channel),
                                       int(ch_dim), len(ops), make_ops(ops), int(accel))
        assert r == 0
        return out

    def transpose(self, x, dim, order):
        x = np.ascontiguousarray(x)
        out = np.empty_like(x)
        self.lib.oracle_transpose(_ptr(x), _ptr(out), TYPE_ID[x.dtype.name], dims8(dim),
                                  (C.c_uint8 * 4)(*order))
        return out

    def dimchg(self, x, dim, from_, to):
        x = np.ascontiguousarray(x)
        out = np.empty_like(x)
        r = self.lib.oracle_dimchg(_ptr(x), _ptr(out), TYPE_ID[x.dtype.name], dims8(dim), from_, to)
        if r != 0:
            raise ValueError("dimchg from > to: GST_FLOW_ERROR")
        return out

    def stand(self, x, dim, mode="default", out_type=None, per_channel=False):
        x = np.ascontiguousarray(x)
        out_type = out_type or x.dtype.name
        out = np.empty(x.size, dtype=NP_DTYPE[out_type])
        r = self.lib.oracle_stand(_ptr(x), _ptr(out), TYPE_ID[x.dtype.name], TYPE_ID[out_type],
                                  dims8(dim), 0 if mode == "default" else 1, int(per_channel))
        assert r == 0
        return out

    def stand_stats(self, x, dim, per_channel=False):
        x = np.ascontiguousarray(x)
        nch = int(dim[0]) if per_channel else 1
        mean = np.zeros(nch, dtype=np.float64)
        std = np.zeros(nch, dtype=np.float64)
        self.lib.oracle_stand_stats(_ptr(x), TYPE_ID[x.dtype.name], dims8(dim), int(per_channel),
                                    mean.ctypes.data_as(C.POINTER(C.c_double)),
                                    std.ctypes.data_as(C.POINTER(C.c_double)))
        return mean, std

    def clamp(self, x, lo, hi):
        x = np.ascontiguousarray(x)
        out = np.empty_like(x)
        self.lib.oracle_clamp(_ptr(x), _ptr(out), TYPE_ID[x.dtype.name], x.size, lo, hi)
        return out

    def stride_strip(self, src, color, width, height):
        src = np.ascontiguousarray(src, dtype=np.uint8)
        out = np.empty(color * width * height, dtype=np.uint8)
        self.lib.oracle_stride_strip(_ptr(src), _ptr(out), color, width, height)
        return out


    def direct_video(self, src, color, width, height):
        """dv_decode (tensordec-directvideo.c:322-339); padding bytes come back as zero."""
        src = np.ascontiguousarray(src, dtype=np.uint8)
        pitch = (color * width + 3) // 4 * 4
        out = np.zeros(pitch * height, dtype=np.uint8)
        self.lib.oracle_direct_video(_ptr(src), _ptr(out), color, width, height)
        return out

    def merge_linear(self, tensors, dims, direction):
        """gst_tensor_merge_generate_mem (gsttensor_merge.c:506-586): tensors = list of arrays
        of one type, dims = their rank-4 dims (innermost first)."""
        arrs = [np.ascontiguousarray(t) for t in tensors]
        es = arrs[0].dtype.itemsize
        n = len(arrs)
        ptrs = (C.c_void_p * n)(*[a.ctypes.data for a in arrs])
        flat = (C.c_uint32 * (4 * n))(*[int(d) for dim in dims for d in (list(dim) + [1] * 4)[:4]])
        out = np.empty(sum(a.size for a in arrs), dtype=arrs[0].dtype)
        r = self.lib.oracle_merge_linear(n, ptrs, flat, es, int(direction), _ptr(out))
        assert r == 0
        return out

    def aggregator_concat(self, src, frame_size, frames_out, block_size):
        """gst_tensor_aggregator_concat (gsttensor_aggregator.c:768-787) on raw bytes."""
        src = np.ascontiguousarray(src).view(np.uint8).reshape(-1)
        assert src.size == frame_size * frames_out
        out = np.empty_like(src)
        r = self.lib.oracle_aggregator_concat(_ptr(src), _ptr(out), frame_size, frames_out, block_size)
        assert r == 0
        return out


class AggregatorModel:
    """CPU model of GstTensorAggregator's data path (gsttensor_aggregator.c:835-941): a byte
    adapter, frames_out frames per output, concatenation along frames-dim, frames_flush frames
    flushed.  dim / type describe the INCOMING tensor (the sink caps)."""

    def __init__(self, port_, dim, type_name, frames_in=1, frames_out=1, frames_flush=0,
                 frames_dim=7, concat=True):
        self.port = port_
        self.dim = dims8(dim)
        self.es = np.dtype(type_name).itemsize
        self.fi, self.fo, self.ff, self.fd = frames_in, frames_out, frames_flush, frames_dim
        self.concat = concat
        assert self.dim[frames_dim] % frames_in == 0          # :1066-1070
        self.out_dim = list(self.dim)
        self.out_dim[frames_dim] = self.dim[frames_dim] // frames_in * frames_out
        self.frame_dim = list(self.out_dim)
        self.frame_dim[frames_dim] //= frames_out             # :813-815
        self.adapter = bytearray()

    def reset(self):
        self.adapter = bytearray()

    def _push_out(self, data, frame_size):
        need = self.concat and self.fo > 1 and any(d > 1 for d in self.frame_dim[self.fd + 1:])
        assert frame_size == self.es * int(np.prod(self.frame_dim, dtype=np.uint64))
        if not need:
            return np.frombuffer(bytes(data), dtype=np.uint8).copy()
        block = self.es * int(np.prod(self.frame_dim[:self.fd + 1], dtype=np.uint64))
        return self.port.aggregator_concat(np.frombuffer(bytes(data), dtype=np.uint8), frame_size,
                                           self.fo, block)

    def chain(self, buf):
        buf = np.ascontiguousarray(buf).view(np.uint8).reshape(-1)
        frame_size = buf.size // self.fi
        if self.fi == self.fo:
            return [self._push_out(buf, frame_size)]
        self.adapter += buf.tobytes()
        out_size = frame_size * self.fo
        outs = []
        while len(self.adapter) >= out_size:
            avail = len(self.adapter)
            outs.append(self._push_out(self.adapter[:out_size], frame_size))
            flush = frame_size * self.ff if self.ff > 0 else out_size
            flush = min(flush, avail)
            del self.adapter[:flush]
        return outs


class Ref:
    """The reference's own code (acceleration=false semantics only)."""
    kind = "reference"

    def __init__(self):
        path = build_ref()
        if not path or not os.path.isfile(path):
            raise FileNotFoundError("oracle/_ref/libnns_ref.so is not built and /root/reference is absent")
        self.lib = C.CDLL(path)
        L = self.lib
        L.ref_typecast.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_uint32)]
        L.ref_arithmetic.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                     C.POINTER(C.c_uint32), C.c_int, C.c_uint, C.c_int,
                                     C.POINTER(Op)]
        L.ref_transpose.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_uint32),
                                    C.POINTER(C.c_uint8)]
        L.ref_dimchg.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_uint32),
                                 C.c_int, C.c_int]
        L.ref_stand.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_uint32),
                                C.c_int, C.c_int]
        L.ref_clamp.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_uint32),
                                C.c_double, C.c_double]
        L.ref_convert_dimension.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int,
                                            C.POINTER(C.c_uint8), C.POINTER(C.c_uint32),
                                            C.POINTER(C.c_uint32)]
        L.ref_average.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_double)]
        L.ref_std.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_double, C.POINTER(C.c_double)]

    def typecast(self, x, out_type, accel=False):
        assert not accel
        x = np.ascontiguousarray(x)
        out = np.empty(x.size, dtype=NP_DTYPE[out_type])
        self.lib.ref_typecast(_ptr(x), _ptr(out), TYPE_ID[x.dtype.name], TYPE_ID[out_type],
                              dims8([x.size]))
        return out

    def arithmetic(self, x, dim, ops, per_channel=False, ch_dim=0, accel=False):
        assert not accel
        x = np.ascontiguousarray(x)
        out_type = x.dtype.name
        for o in ops:
            if o[0] == "typecast":
                out_type = o[1]
        out = np.empty(x.size, dtype=NP_DTYPE[out_type])
        r = self.lib.ref_arithmetic(_ptr(x), _ptr(out), TYPE_ID[x.dtype.name], TYPE_ID[out_type],
                                    dims8(dim), int(per_channel), int(ch_dim), len(ops),
                                    make_ops(ops))
        assert r == 0
        return out

    def transpose(self, x, dim, order):
        x = np.ascontiguousarray(x)
        out = np.empty_like(x)
        self.lib.ref_transpose(_ptr(x), _ptr(out), TYPE_ID[x.dtype.name], dims8(dim),
                               (C.c_uint8 * 4)(*order))
        return out

    def dimchg(self, x, dim, from_, to):
        """Only from == 0 (or from == to) is memory-safe in the reference."""
        assert from_ == 0 or from_ == to
        x = np.ascontiguousarray(x)
        out = np.empty_like(x)
        r = self.lib.ref_dimchg(_ptr(x), _ptr(out), TYPE_ID[x.dtype.name], dims8(dim), from_, to)
        if r != 0:
            raise ValueError("dimchg: GST_FLOW_ERROR")
        return out

    def stand(self, x, dim, mode="default", out_type=None, per_channel=False):
        x = np.ascontiguousarray(x)
        out_type_id = TYPE_ID[out_type or x.dtype.name]
        out = np.empty(x.size, dtype=NP_DTYPE[out_type or x.dtype.name])
        r = self.lib.ref_stand(_ptr(x), _ptr(out), TYPE_ID[x.dtype.name], out_type_id, dims8(dim),
                               0 if mode == "default" else 1, int(per_channel))
        assert r == 0
        return out

    def clamp(self, x, lo, hi):
        x = np.ascontiguousarray(x)
        out = np.empty_like(x)
        self.lib.ref_clamp(_ptr(x), _ptr(out), TYPE_ID[x.dtype.name], dims8([x.size]), lo, hi)
        return out

    def convert_dimension(self, mode, dim, sink=True, from_=0, to=0, order=(0, 1, 2, 3)):
        modes = {"dimchg": 0, "typecast": 1, "arithmetic": 2, "transpose": 3, "stand": 4, "clamp": 5}
        out = (C.c_uint32 * RANK)()
        self.lib.ref_convert_dimension(modes[mode], int(sink), from_, to, (C.c_uint8 * 4)(*order),
                                       dims8(dim), out)
        return list(out)


_port = None
_ref = None


def port():
    global _port
    if _port is None:
        _port = Port()
    return _port


def ref():
    global _ref
    if _ref is None:
        _ref = Ref()
    return _ref


def ref_available():
    try:
        ref()
        return True
    except (FileNotFoundError, OSError):
        return False
