"""Python mirror of the reference elements' interface on top of the C-ABI element functions.

TensorTransform  <->  tensor_transform  (gst/nnstreamer/elements/gsttensor_transform.c)
TensorConverter  <->  tensor_converter, video branch (gsttensor_converter.c)
BatchQueue            the device-side batching queue

Same property names and value strings, same caps strings; `chain`/`transform` take and return
lists of numpy arrays (host memories) or DeviceMemory handles.  Nothing is computed here: every
call goes to libnnstreamer_b200.so.
"""
import ctypes as C

import numpy as np

from . import _lib as L

_CAPS_BUF = 4096


class DeviceMemory:
    """A device-resident GstMemory stand-in: a reference-counted, stream-ordered view of a pooled
    block (or of caller-managed device memory)."""

    def __init__(self, mem):
        self.mem = mem

    @property
    def ptr(self):
        """gst_memory_map: waits until the producer's kernels have finished, then the pointer."""
        L.check(L.lib().nnsb_memory_sync(C.byref(self.mem)))
        return self.mem.data

    @property
    def ptr_async(self):
        """The pointer without waiting: only for work ordered after `wait(stream)`."""
        return self.mem.data

    @property
    def size(self):
        return self.mem.size

    def wait(self, stream):
        """Make `stream` (an nnsb_stream handle) wait on the GPU for the contents."""
        L.check(L.lib().nnsb_memory_wait(C.byref(self.mem), stream))

    def sync(self):
        L.check(L.lib().nnsb_memory_sync(C.byref(self.mem)))

    def mark_use(self, stream):
        L.check(L.lib().nnsb_memory_mark_use(C.byref(self.mem), stream))

    def mark_ready(self, stream):
        L.check(L.lib().nnsb_memory_mark_ready(C.byref(self.mem), stream))

    def ref(self):
        out = L.Memory()
        L.check(L.lib().nnsb_memory_ref(C.byref(self.mem), C.byref(out)))
        return DeviceMemory(out)

    def share(self, offset, size):
        """gst_memory_share: a sub-range that keeps the block alive on its own."""
        out = L.Memory()
        L.check(L.lib().nnsb_memory_share(C.byref(self.mem), int(offset), int(size), C.byref(out)))
        return DeviceMemory(out)

    def export(self):
        """(device pointer, size, device id, cudaEvent_t or None) for CUDA-aware consumers."""
        v = L.DeviceView()
        L.check(L.lib().nnsb_memory_export(C.byref(self.mem), C.byref(v)))
        return v.device_ptr, v.size, v.device_id, v.ready_event

    @property
    def __cuda_array_interface__(self):
        """uint8 view for torch / cupy (`torch.as_tensor(mem, device="cuda")`); the producer's event
        is waited for here because the interface cannot carry one."""
        self.sync()
        return {"shape": (int(self.mem.size),), "typestr": "|u1", "data": (int(self.mem.data), False),
                "version": 3, "strides": None}

    def release(self):
        L.lib().nnsb_memory_release(C.byref(self.mem))


def _host_memory(arr):
    m = L.Memory()
    m.data = arr.ctypes.data_as(C.c_void_p)
    m.size = arr.nbytes
    m.on_device = 0
    m.owned = 0
    m.owner = None
    m.block = None
    return m


def device_memory(ptr, size):
    m = L.Memory()
    m.data = C.c_void_p(ptr)
    m.size = size
    m.on_device = 1
    m.owned = 0
    m.owner = None
    m.block = None
    return DeviceMemory(m)


class TensorTransform:
    """gst_element_factory_make ("tensor_transform")"""

    def __init__(self, device_id=0, **props):
        self._lib = L.lib()
        self.handle = C.c_void_p()
        L.check(self._lib.nnsb_transform_new(int(device_id), C.byref(self.handle)))
        for k, v in props.items():
            self.set_property(k.replace("_", "-"), v)

    def close(self):
        if self.handle:
            self._lib.nnsb_transform_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- GObject properties ----------------------------------------------------------------
    def set_property(self, name, value):
        """Returns True when accepted.  Like g_object_set, an invalid `option` raises nothing:
        the previous option stays in force (gsttensor_transform.c:985-990)."""
        if isinstance(value, bool):
            value = "true" if value else "false"
        val = None if value is None else str(value).encode()
        code = self._lib.nnsb_transform_set_property(self.handle, name.encode(), val)
        if code != 0 and name != "option":
            L.check(code)
        return code == 0

    def get_property(self, name):
        buf = C.create_string_buffer(_CAPS_BUF)
        code = self._lib.nnsb_transform_get_property(self.handle, name.encode(), buf, _CAPS_BUF)
        if code == L.ENOENT:
            return None
        if code < 0:
            L.check(code)
        s = buf.value.decode()
        if name in ("silent", "acceleration"):
            return s == "true"
        if name == "transpose-rank-limit":
            return int(s)
        return s

    # -- GstBaseTransform vfuncs -----------------------------------------------------------
    def transform_caps(self, direction, caps, filter_caps=None):
        buf = C.create_string_buffer(_CAPS_BUF)
        d = L.PAD_SINK if direction == "sink" else L.PAD_SRC
        code = self._lib.nnsb_transform_transform_caps(
            self.handle, d, caps.encode(), filter_caps.encode() if filter_caps else None, buf, _CAPS_BUF)
        if code < 0:
            L.check(code)
        return buf.value.decode()

    def fixate_caps(self, direction, caps, othercaps):
        buf = C.create_string_buffer(_CAPS_BUF)
        d = L.PAD_SINK if direction == "sink" else L.PAD_SRC
        code = self._lib.nnsb_transform_fixate_caps(self.handle, d, caps.encode(), othercaps.encode(),
                                                    buf, _CAPS_BUF)
        if code < 0:
            L.check(code)
        return buf.value.decode()

    def set_caps(self, incaps, outcaps):
        return self._lib.nnsb_transform_set_caps(self.handle, incaps.encode(), outcaps.encode()) == 0

    def negotiate(self, incaps, peer_caps="other/tensor; other/tensors"):
        """What BaseTransform does with a fixed sink caps: transform_caps -> fixate -> set_caps.
        Returns the fixed src caps string."""
        out = self.fixate_caps("sink", incaps, peer_caps)
        if not self.set_caps(incaps, out):
            raise L.NnsbError(L.EINVAL, self._lib.nnsb_last_error().decode())
        return out

    def config(self, direction="src"):
        cfg = L.TensorsConfig()
        d = L.PAD_SINK if direction == "sink" else L.PAD_SRC
        L.check(self._lib.nnsb_transform_get_config(self.handle, d, C.byref(cfg)))
        return cfg

    def transform(self, memories):
        """gst_tensor_transform_transform: list of numpy arrays (host) or DeviceMemory ->
        (flow_return, list of outputs).  Host outputs are numpy arrays (copied out of the
        element's pinned buffers); device outputs are DeviceMemory."""
        n = len(memories)
        ins = (L.Memory * max(1, n))()
        keep = []
        for i, m in enumerate(memories):
            if isinstance(m, DeviceMemory):
                ins[i] = m.mem
            else:
                a = np.ascontiguousarray(m)
                keep.append(a)
                ins[i] = _host_memory(a)
        outs = (L.Memory * L.SIZE_LIMIT)()
        n_out = C.c_uint32(0)
        flow = self._lib.nnsb_transform_transform(self.handle, ins, n, outs, C.byref(n_out))
        results = []
        for i in range(n_out.value):
            m = outs[i]
            if m.on_device:
                results.append(DeviceMemory(m))
            else:
                buf = (C.c_uint8 * m.size).from_address(m.data)
                results.append(np.frombuffer(buf, dtype=np.uint8).copy())
                if m.owned:
                    self._lib.nnsb_memory_release(C.byref(m))
        return flow, results

    def transform_raw(self, ins, n, outs, n_out):
        """No-copy variant for benchmarks: ctypes Memory arrays in, Memory array out."""
        return self._lib.nnsb_transform_transform(self.handle, ins, n, outs, C.byref(n_out))

    def alloc(self, size, on_device=False):
        m = L.Memory()
        L.check(self._lib.nnsb_transform_alloc(self.handle, int(size), int(bool(on_device)), C.byref(m)))
        return m

    @property
    def kernel_launches(self):
        c = C.c_uint64(0)
        L.check(self._lib.nnsb_transform_kernel_launches(self.handle, C.byref(c)))
        return int(c.value)

    def link_to_aggregator(self, aggregator):
        """Device outputs are written where the downstream tensor_aggregator wants them (its ring):
        what a downstream-proposed buffer pool does in GStreamer.  None unlinks."""
        if aggregator is None:
            L.check(self._lib.nnsb_transform_set_output_provider(self.handle, None, None))
        else:
            fn = C.cast(self._lib.nnsb_aggregator_acquire, C.c_void_p)
            L.check(self._lib.nnsb_transform_set_output_provider(self.handle, fn, aggregator.handle))


class TensorConverter:
    """gst_element_factory_make ("tensor_converter"), video/x-raw input only."""

    def __init__(self, device_id=0, **props):
        self._lib = L.lib()
        self.handle = C.c_void_p()
        L.check(self._lib.nnsb_converter_new(int(device_id), C.byref(self.handle)))
        for k, v in props.items():
            self.set_property(k.replace("_", "-"), v)

    def close(self):
        if self.handle:
            self._lib.nnsb_converter_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_property(self, name, value):
        if isinstance(value, bool):
            value = "true" if value else "false"
        L.check(self._lib.nnsb_converter_set_property(self.handle, name.encode(), str(value).encode()))

    def get_property(self, name):
        buf = C.create_string_buffer(256)
        code = self._lib.nnsb_converter_get_property(self.handle, name.encode(), buf, 256)
        if code < 0:
            L.check(code)
        s = buf.value.decode()
        return int(s) if name == "frames-per-tensor" else (s == "true")

    def set_caps(self, video_caps):
        buf = C.create_string_buffer(_CAPS_BUF)
        code = self._lib.nnsb_converter_set_caps(self.handle, video_caps.encode(), buf, _CAPS_BUF)
        if code < 0:
            L.check(code)
        return buf.value.decode()

    @property
    def remove_padding(self):
        return bool(self._lib.nnsb_converter_remove_padding(self.handle))

    @property
    def frame_size(self):
        return self._lib.nnsb_converter_frame_size(self.handle)

    def config(self):
        cfg = L.TensorsConfig()
        L.check(self._lib.nnsb_converter_get_config(self.handle, C.byref(cfg)))
        return cfg

    def chain(self, frame):
        """One video frame in (numpy uint8 array or DeviceMemory); returns (flow, DeviceMemory
        or None): a tensor comes out every frames-per-tensor frames."""
        if isinstance(frame, DeviceMemory):
            m = frame.mem
        else:
            a = np.ascontiguousarray(frame, dtype=np.uint8)
            m = _host_memory(a)
        out = L.Memory()
        produced = C.c_int(0)
        flow = self._lib.nnsb_converter_chain(self.handle, C.byref(m), C.byref(out), C.byref(produced))
        return flow, (DeviceMemory(out) if produced.value else None)


class TensorAggregator:
    """gst_element_factory_make ("tensor_aggregator"): frames-in / frames-out / frames-flush /
    frames-dim / concat, with the adapter and the concatenation on the device."""

    _UINT_PROPS = ("frames-in", "frames-out", "frames-flush", "frames-dim")

    def __init__(self, device_id=0, **props):
        self._lib = L.lib()
        self.handle = C.c_void_p()
        L.check(self._lib.nnsb_aggregator_new(int(device_id), C.byref(self.handle)))
        for k, v in props.items():
            self.set_property(k.replace("_", "-"), v)

    def close(self):
        if self.handle:
            self._lib.nnsb_aggregator_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_property(self, name, value):
        if isinstance(value, bool):
            value = "true" if value else "false"
        L.check(self._lib.nnsb_aggregator_set_property(self.handle, name.encode(), str(value).encode()))

    def get_property(self, name):
        buf = C.create_string_buffer(256)
        code = self._lib.nnsb_aggregator_get_property(self.handle, name.encode(), buf, 256)
        if code < 0:
            L.check(code)
        s = buf.value.decode()
        return int(s) if name in self._UINT_PROPS else (s == "true")

    def set_caps(self, caps):
        """Sink caps in, source caps out (gst_tensor_aggregator_parse_caps)."""
        buf = C.create_string_buffer(_CAPS_BUF)
        code = self._lib.nnsb_aggregator_set_caps(self.handle, caps.encode(), buf, _CAPS_BUF)
        if code < 0:
            L.check(code)
        return buf.value.decode()

    def config(self, direction):
        cfg = L.TensorsConfig()
        L.check(self._lib.nnsb_aggregator_get_config(self.handle, int(direction), C.byref(cfg)))
        return cfg

    def reset(self):
        L.check(self._lib.nnsb_aggregator_reset(self.handle))

    def outputs_for(self, nbytes):
        return self._lib.nnsb_aggregator_outputs_for(self.handle, int(nbytes))

    def chain(self, buf, pts=None, dts=None, duration=None):
        """One buffer in (numpy array or DeviceMemory); returns (flow, [DeviceMemory...],
        [(pts, dts, duration)...]) with None for GST_CLOCK_TIME_NONE."""
        if isinstance(buf, DeviceMemory):
            m = buf.mem
        else:
            a = np.ascontiguousarray(buf)
            m = _host_memory(a)
        none = L.CLOCK_TIME_NONE
        t_in = L.BufferTime(none if pts is None else int(pts), none if dts is None else int(dts),
                            none if duration is None else int(duration))
        room = max(1, self.outputs_for(m.size))
        outs = (L.Memory * room)()
        t_outs = (L.BufferTime * room)()
        n_out = C.c_uint32(0)
        flow = self._lib.nnsb_aggregator_chain(self.handle, C.byref(m), C.byref(t_in), outs, t_outs,
                                               room, C.byref(n_out))
        mems, times = [], []
        for i in range(n_out.value):
            o = L.Memory()
            C.memmove(C.byref(o), C.byref(outs[i]), C.sizeof(L.Memory))
            mems.append(DeviceMemory(o))
            t = t_outs[i]
            times.append(tuple(None if v == none else int(v) for v in (t.pts, t.dts, t.duration)))
        return flow, mems, times

    @property
    def kernel_launches(self):
        c = C.c_uint64(0)
        L.check(self._lib.nnsb_aggregator_kernel_launches(self.handle, C.byref(c)))
        return int(c.value)

    def acquire(self, nbytes):
        """The device memory the next input buffer should be written into (a share of the ring)."""
        m = L.Memory()
        L.check(self._lib.nnsb_aggregator_acquire(self.handle, int(nbytes), C.byref(m)))
        return DeviceMemory(m)

    @property
    def bytes_copied(self):
        """(bytes copied into the adapter, bytes moved by the concatenation) so far."""
        a, b = C.c_uint64(0), C.c_uint64(0)
        L.check(self._lib.nnsb_aggregator_copy_counters(self.handle, C.byref(a), C.byref(b)))
        return int(a.value), int(b.value)


class DirectVideoDecoder:
    """tensor_decoder mode=direct_video (tensordec-directvideo.c): tensor -> video/x-raw frame,
    rows padded to a multiple of 4 bytes."""

    def __init__(self, device_id=0, option1=None):
        self._lib = L.lib()
        self.handle = C.c_void_p()
        L.check(self._lib.nnsb_direct_video_new(int(device_id), C.byref(self.handle)))
        if option1 is not None and not self.set_option(0, option1):
            raise ValueError("direct_video: unknown format %r" % (option1,))

    def close(self):
        if self.handle:
            self._lib.nnsb_direct_video_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, op_num, param):
        return bool(self._lib.nnsb_direct_video_set_option(
            self.handle, int(op_num), None if param is None else str(param).encode()))

    @staticmethod
    def make_config(channels, width, height, rate=(0, 1), type_name="uint8"):
        cfg = L.TensorsConfig()
        cfg.num_tensors = 1
        cfg.info[0] = L.make_info(type_name, [channels, width, height, 1])
        cfg.rate_n, cfg.rate_d = rate
        return cfg

    def out_caps(self, cfg):
        buf = C.create_string_buffer(_CAPS_BUF)
        code = self._lib.nnsb_direct_video_get_out_caps(self.handle, C.byref(cfg), buf, _CAPS_BUF)
        if code < 0:
            L.check(code)
        return buf.value.decode()

    def transform_size(self, cfg):
        return self._lib.nnsb_direct_video_transform_size(C.byref(cfg))

    def decode(self, cfg, tensor, out=None):
        """tensor: numpy uint8 array or DeviceMemory.  out: None (a device memory is allocated),
        a numpy uint8 array of transform_size bytes, or a DeviceMemory to fill.  Returns
        (flow, DeviceMemory or the numpy array)."""
        if isinstance(tensor, DeviceMemory):
            m = tensor.mem
        else:
            a = np.ascontiguousarray(tensor, dtype=np.uint8)
            m = _host_memory(a)
        if out is None:
            o = L.Memory()
        elif isinstance(out, DeviceMemory):
            o = out.mem
        else:
            o = _host_memory(out)
        flow = self._lib.nnsb_direct_video_decode(self.handle, C.byref(cfg), C.byref(m), C.byref(o))
        if out is None:
            return flow, (DeviceMemory(o) if flow == 0 else None)
        return flow, out


class TensorMerge:
    """tensor_merge mode=linear: N single-tensor pads concatenated along axis `option`."""

    def __init__(self, device_id=0, **props):
        self._lib = L.lib()
        self.handle = C.c_void_p()
        L.check(self._lib.nnsb_merge_new(int(device_id), C.byref(self.handle)))
        for k, v in props.items():
            self.set_property(k.replace("_", "-"), v)

    def close(self):
        if self.handle:
            self._lib.nnsb_merge_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_property(self, name, value):
        if isinstance(value, bool):
            value = "true" if value else "false"
        L.check(self._lib.nnsb_merge_set_property(self.handle, name.encode(), str(value).encode()))

    def get_property(self, name):
        buf = C.create_string_buffer(256)
        code = self._lib.nnsb_merge_get_property(self.handle, name.encode(), buf, 256)
        if code == L.ENOENT:
            return None
        if code < 0:
            L.check(code)
        return buf.value.decode()

    def set_caps(self, caps_list):
        arr = (C.c_char_p * len(caps_list))(*[c.encode() for c in caps_list])
        buf = C.create_string_buffer(_CAPS_BUF)
        code = self._lib.nnsb_merge_set_caps(self.handle, len(caps_list), arr, buf, _CAPS_BUF)
        if code < 0:
            L.check(code)
        return buf.value.decode()

    def chain(self, tensors):
        """One memory per pad (numpy arrays or DeviceMemory); returns (flow, DeviceMemory)."""
        n = len(tensors)
        ins = (L.Memory * n)()
        keep = []
        for i, t in enumerate(tensors):
            if isinstance(t, DeviceMemory):
                ins[i] = t.mem
            else:
                a = np.ascontiguousarray(t)
                keep.append(a)
                ins[i] = _host_memory(a)
        out = L.Memory()
        flow = self._lib.nnsb_merge_chain(self.handle, ins, n, C.byref(out))
        return flow, (DeviceMemory(out) if flow == 0 else None)


class TensorSplit:
    """tensor_split: tensorseg / tensorpick, one device memory per picked segment."""

    def __init__(self, device_id=0, **props):
        self._lib = L.lib()
        self.handle = C.c_void_p()
        L.check(self._lib.nnsb_split_new(int(device_id), C.byref(self.handle)))
        for k, v in props.items():
            self.set_property(k.replace("_", "-"), v)

    def close(self):
        if self.handle:
            self._lib.nnsb_split_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_property(self, name, value):
        if isinstance(value, bool):
            value = "true" if value else "false"
        L.check(self._lib.nnsb_split_set_property(self.handle, name.encode(), str(value).encode()))

    def get_property(self, name):
        buf = C.create_string_buffer(1024)
        code = self._lib.nnsb_split_get_property(self.handle, name.encode(), buf, 1024)
        if code < 0:
            L.check(code)
        return buf.value.decode()

    def set_caps(self, caps):
        L.check(self._lib.nnsb_split_set_caps(self.handle, caps.encode()))

    @property
    def num_outputs(self):
        return self._lib.nnsb_split_num_outputs(self.handle)

    def out_caps(self, idx):
        buf = C.create_string_buffer(_CAPS_BUF)
        code = self._lib.nnsb_split_get_out_caps(self.handle, int(idx), buf, _CAPS_BUF)
        if code < 0:
            L.check(code)
        return buf.value.decode()

    def chain(self, tensor):
        if isinstance(tensor, DeviceMemory):
            m = tensor.mem
        else:
            a = np.ascontiguousarray(tensor)
            m = _host_memory(a)
        room = max(1, self.num_outputs)
        outs = (L.Memory * room)()
        n_out = C.c_uint32(0)
        flow = self._lib.nnsb_split_chain(self.handle, C.byref(m), outs, room, C.byref(n_out))
        mems = []
        for i in range(n_out.value):
            o = L.Memory()
            C.memmove(C.byref(o), C.byref(outs[i]), C.sizeof(L.Memory))
            mems.append(DeviceMemory(o))
        return flow, mems


class TensorFilter:
    """tensor_filter's transform with a CUDA-aware framework behind it: `invoke(inputs, outputs,
    cuda_stream)` receives lists of (device pointer, size) and the raw cudaStream_t it must
    enqueue its work on (no synchronisation inside); returns 0, >0 (drop) or <0 (error)."""

    def __init__(self, invoke, in_config, out_config, device_id=0):
        self._lib = L.lib()
        self.handle = C.c_void_p()
        L.check(self._lib.nnsb_filter_new(int(device_id), C.byref(self.handle)))
        self._invoke = invoke

        def trampoline(_priv, inp, n_in, outp, n_out, stream):
            try:
                ins = [(inp[i].data, inp[i].size) for i in range(n_in)]
                outs = [(outp[i].data, outp[i].size) for i in range(n_out)]
                return int(self._invoke(ins, outs, stream) or 0)
            except Exception:                       # noqa: BLE001 -- must not unwind into C
                import traceback
                traceback.print_exc()
                return -1

        self._cb = L.FILTER_INVOKE(trampoline)      # keep the thunk alive
        L.check(self._lib.nnsb_filter_set_framework(self.handle, self._cb, None))
        L.check(self._lib.nnsb_filter_set_info(self.handle, C.byref(in_config), C.byref(out_config)))

    @staticmethod
    def make_config(infos, rate=(0, 1)):
        """infos: list of (type name, dim list)."""
        cfg = L.TensorsConfig()
        cfg.num_tensors = len(infos)
        for i, (t, d) in enumerate(infos):
            cfg.info[i] = L.make_info(t, d)
        cfg.rate_n, cfg.rate_d = rate
        return cfg

    def close(self):
        if self.handle:
            self._lib.nnsb_filter_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def transform(self, memories):
        n = len(memories)
        ins = (L.Memory * max(1, n))()
        keep = []
        for i, m in enumerate(memories):
            if isinstance(m, DeviceMemory):
                ins[i] = m.mem
            else:
                a = np.ascontiguousarray(m)
                keep.append(a)
                ins[i] = _host_memory(a)
        outs = (L.Memory * L.SIZE_LIMIT)()
        n_out = C.c_uint32(0)
        flow = self._lib.nnsb_filter_transform(self.handle, ins, n, outs, C.byref(n_out))
        res = []
        for i in range(n_out.value):
            o = L.Memory()
            C.memmove(C.byref(o), C.byref(outs[i]), C.sizeof(L.Memory))
            res.append(DeviceMemory(o))
        return flow, res

    @property
    def counters(self):
        """(bytes uploaded, bytes downloaded, invoke calls) by this element so far."""
        a, b, c = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        L.check(self._lib.nnsb_filter_copy_counters(self.handle, C.byref(a), C.byref(b), C.byref(c)))
        return int(a.value), int(b.value), int(c.value)


def _as_memory_array(memories):
    n = len(memories)
    ins = (L.Memory * max(1, n))()
    keep = []
    for i, m in enumerate(memories):
        if isinstance(m, DeviceMemory):
            ins[i] = m.mem
        else:
            a = np.ascontiguousarray(m)
            keep.append(a)
            ins[i] = _host_memory(a)
    return ins, keep


def _wrap_out(m):
    """An output memory of mux / demux: device -> DeviceMemory, host view -> the ctypes Memory."""
    o = L.Memory()
    C.memmove(C.byref(o), C.byref(m), C.sizeof(L.Memory))
    return DeviceMemory(o) if o.on_device else o


class Branch:
    """A linked run of elements pushed by one call (GStreamer's push model: the converter's chain
    pushes into the transform's chain in the same streaming thread): `TensorConverter` (optional)
    followed by `TensorTransform` instances, all negotiated already.  One C-ABI call per frame
    instead of one per element -- what a Python streaming thread needs to keep up with the GPU."""

    def __init__(self, converter, transforms):
        self._lib = L.lib()
        self._elements = (converter, list(transforms))      # kept alive with the branch
        arr = (C.c_void_p * max(1, len(transforms)))(*[t.handle for t in transforms])
        self.handle = C.c_void_p()
        L.check(self._lib.nnsb_branch_new(converter.handle if converter is not None else None,
                                          arr, len(transforms), C.byref(self.handle)))
        self._outs = (L.Memory * L.SIZE_LIMIT)()
        self._n_out = C.c_uint32(0)

    def close(self):
        if self.handle:
            self._lib.nnsb_branch_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def push(self, memories):
        """One frame (a DeviceMemory or numpy array) or the memories of one tensor buffer in ->
        (flow_return, list of DeviceMemory / numpy outputs; empty while the converter collects)."""
        if not isinstance(memories, (list, tuple)):
            memories = [memories]
        n = len(memories)
        ins = (L.Memory * max(1, n))()
        keep = []
        for i, m in enumerate(memories):
            if isinstance(m, DeviceMemory):
                ins[i] = m.mem
            else:
                a = np.ascontiguousarray(m)
                keep.append(a)
                ins[i] = _host_memory(a)
        flow = self._lib.nnsb_branch_push(self.handle, ins, n, self._outs, C.byref(self._n_out))
        results = []
        for i in range(self._n_out.value):
            m = L.Memory()
            C.memmove(C.byref(m), C.byref(self._outs[i]), C.sizeof(L.Memory))
            if m.on_device:
                results.append(DeviceMemory(m))
            else:
                buf = (C.c_uint8 * m.size).from_address(m.data)
                results.append(np.frombuffer(buf, dtype=np.uint8).copy())
                if m.owned:
                    self._lib.nnsb_memory_release(C.byref(m))
        return flow, results

    def push_raw(self, ins, n, outs, n_out):
        """No-copy variant: ctypes Memory arrays in and out, the caller releases the outputs."""
        return self._lib.nnsb_branch_push(self.handle, ins, n, outs, n_out)


class TensorMux:
    """tensor_mux: the memories of all sink pads, in pad order, become one other/tensors buffer
    (references, nothing is copied)."""

    def __init__(self, **props):
        self._lib = L.lib()
        self.handle = C.c_void_p()
        L.check(self._lib.nnsb_mux_new(C.byref(self.handle)))
        for k, v in props.items():
            self.set_property(k.replace("_", "-"), v)

    def close(self):
        if self.handle:
            self._lib.nnsb_mux_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_property(self, name, value):
        if isinstance(value, bool):
            value = "true" if value else "false"
        L.check(self._lib.nnsb_mux_set_property(self.handle, name.encode(), str(value).encode()))

    def get_property(self, name):
        buf = C.create_string_buffer(256)
        code = self._lib.nnsb_mux_get_property(self.handle, name.encode(), buf, 256)
        if code == L.ENOENT:
            return None
        if code < 0:
            L.check(code)
        return buf.value.decode()

    def set_caps(self, caps_list):
        arr = (C.c_char_p * len(caps_list))(*[c.encode() for c in caps_list])
        buf = C.create_string_buffer(_CAPS_BUF)
        code = self._lib.nnsb_mux_set_caps(self.handle, len(caps_list), arr, buf, _CAPS_BUF)
        if code < 0:
            L.check(code)
        return buf.value.decode()

    def chain(self, memories):
        ins, _keep = _as_memory_array(memories)
        outs = (L.Memory * L.SIZE_LIMIT)()
        n_out = C.c_uint32(0)
        flow = self._lib.nnsb_mux_chain(self.handle, ins, len(memories), outs, C.byref(n_out))
        return flow, [_wrap_out(outs[i]) for i in range(n_out.value)]


class TensorDemux:
    """tensor_demux: `tensorpick` groups -> one list of memory references per source pad."""

    def __init__(self, **props):
        self._lib = L.lib()
        self.handle = C.c_void_p()
        L.check(self._lib.nnsb_demux_new(C.byref(self.handle)))
        for k, v in props.items():
            self.set_property(k.replace("_", "-"), v)

    def close(self):
        if self.handle:
            self._lib.nnsb_demux_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_property(self, name, value):
        if isinstance(value, bool):
            value = "true" if value else "false"
        L.check(self._lib.nnsb_demux_set_property(self.handle, name.encode(), str(value).encode()))

    def get_property(self, name):
        buf = C.create_string_buffer(1024)
        code = self._lib.nnsb_demux_get_property(self.handle, name.encode(), buf, 1024)
        if code < 0:
            L.check(code)
        return buf.value.decode()

    def set_caps(self, caps):
        L.check(self._lib.nnsb_demux_set_caps(self.handle, caps.encode()))

    @property
    def num_outputs(self):
        return self._lib.nnsb_demux_num_outputs(self.handle)

    def out_caps(self, idx):
        buf = C.create_string_buffer(_CAPS_BUF)
        code = self._lib.nnsb_demux_get_out_caps(self.handle, int(idx), buf, _CAPS_BUF)
        if code < 0:
            L.check(code)
        return buf.value.decode()

    def chain(self, memories):
        """Returns (flow, [[memories of src_0], [memories of src_1], ...])."""
        ins, _keep = _as_memory_array(memories)
        room = 64
        outs = (L.Memory * room)()
        counts = (C.c_uint32 * L.SIZE_LIMIT)()
        n_src = C.c_uint32(0)
        flow = self._lib.nnsb_demux_chain(self.handle, ins, len(memories), outs, room, counts,
                                          L.SIZE_LIMIT, C.byref(n_src))
        pads, k = [], 0
        for i in range(n_src.value):
            pads.append([_wrap_out(outs[k + j]) for j in range(counts[i])])
            k += counts[i]
        return flow, pads


class BatchQueue:
    """nnsb_queue: groups frames so that one launch covers `batch` tensors."""

    def __init__(self, dev, plan, batch):
        self._lib = L.lib()
        self.dev = dev
        self.plan = plan
        self.batch = int(batch)
        self.handle = C.c_void_p()
        L.check(self._lib.nnsb_queue_create(dev.ctx, plan.handle, self.batch, C.byref(self.handle)))

    def push(self, frame):
        full = C.c_int(0)
        if isinstance(frame, int):
            L.check(self._lib.nnsb_queue_push(self.handle, C.c_void_p(frame), 1, C.byref(full)))
        else:
            a = np.ascontiguousarray(frame)
            self._keep = a
            L.check(self._lib.nnsb_queue_push(self.handle, a.ctypes.data_as(C.c_void_p), 0, C.byref(full)))
        return bool(full.value)

    def flush(self):
        """Launch over everything queued; returns (device pointer, frames, stream handle)."""
        d_out = C.c_void_p()
        n = C.c_uint32(0)
        stream = C.c_void_p()
        L.check(self._lib.nnsb_queue_flush(self.handle, C.byref(d_out), C.byref(n), C.byref(stream)))
        return d_out.value, int(n.value), stream

    def close(self):
        if self.handle:
            self._lib.nnsb_queue_destroy(self.handle)
            self.handle = C.c_void_p()
