"""Thin object wrappers over the C-ABI runtime: context, stream, memory, plans.

numpy is used only to hold HOST buffers; all arithmetic happens in the CUDA kernels.
"""
import ctypes as C

import numpy as np

from . import _lib as L


class Device:
    """One nnsb_ctx + one nnsb_stream (what one element instance owns)."""

    def __init__(self, device_id=0):
        self._lib = L.lib()
        self.ctx = C.c_void_p()
        L.check(self._lib.nnsb_ctx_create(int(device_id), C.byref(self.ctx)))
        self.stream = C.c_void_p()
        L.check(self._lib.nnsb_stream_create(self.ctx, C.byref(self.stream)))
        self.device_id = int(device_id)

    # -- lifecycle -----------------------------------------------------------------------
    def close(self):
        if self.stream:
            self._lib.nnsb_stream_destroy(self.stream)
            self.stream = C.c_void_p()
        if self.ctx:
            self._lib.nnsb_ctx_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def sm_count(self):
        return self._lib.nnsb_ctx_sm_count(self.ctx)

    # -- memory --------------------------------------------------------------------------
    def alloc(self, nbytes):
        p = C.c_void_p()
        L.check(self._lib.nnsb_dev_alloc(self.ctx, max(1, int(nbytes)), C.byref(p)))
        return p.value

    def free(self, ptr):
        if ptr:
            L.check(self._lib.nnsb_dev_free(self.ctx, C.c_void_p(ptr)))

    def pinned(self, nbytes, dtype=np.uint8):
        """A page-locked host array (numpy view over pooled cudaHostAlloc memory)."""
        p = C.c_void_p()
        nbytes = max(1, int(nbytes))
        L.check(self._lib.nnsb_pinned_alloc(self.ctx, nbytes, C.byref(p)))
        buf = (C.c_uint8 * nbytes).from_address(p.value)
        arr = np.frombuffer(buf, dtype=np.uint8, count=nbytes).view(dtype)
        return arr, p.value

    def pinned_free(self, ptr):
        if ptr:
            L.check(self._lib.nnsb_pinned_free(self.ctx, C.c_void_p(ptr)))

    def h2d(self, dst, host, stream=None):
        host = np.ascontiguousarray(host)
        L.check(self._lib.nnsb_h2d_async(stream or self.stream, C.c_void_p(dst),
                                         host.ctypes.data_as(C.c_void_p), host.nbytes))
        return host

    def d2h(self, host, src, stream=None):
        L.check(self._lib.nnsb_d2h_async(stream or self.stream, host.ctypes.data_as(C.c_void_p),
                                         C.c_void_p(src), host.nbytes))

    def sync(self, stream=None):
        L.check(self._lib.nnsb_stream_sync(stream or self.stream))

    def upload(self, host):
        """Allocate + copy + sync. Returns the device pointer."""
        host = np.ascontiguousarray(host)
        p = self.alloc(host.nbytes)
        if host.nbytes:
            self.h2d(p, host)
            self.sync()
        return p

    def download(self, ptr, count, dtype):
        out = np.empty(int(count), dtype=dtype)
        if out.nbytes:
            self.d2h(out, ptr)
            self.sync()
        return out

    # -- events --------------------------------------------------------------------------
    def event(self):
        e = C.c_void_p()
        L.check(self._lib.nnsb_event_create(self.ctx, C.byref(e)))
        return e

    def record(self, e, stream=None):
        L.check(self._lib.nnsb_event_record(e, stream or self.stream))

    def elapsed_ms(self, e0, e1):
        L.check(self._lib.nnsb_event_sync(e1))
        ms = C.c_float()
        L.check(self._lib.nnsb_event_elapsed_ms(e0, e1, C.byref(ms)))
        return float(ms.value)

    def event_destroy(self, e):
        self._lib.nnsb_event_destroy(e)

    # -- plans ---------------------------------------------------------------------------
    def plan(self, mode, option, in_type, in_dim, acceleration=False):
        return Plan(self, mode, option, in_type, in_dim, acceleration)

    def strip(self, d_src, d_dst, row_bytes, src_pitch, rows, stream=None):
        L.check(self._lib.nnsb_strip_launch(stream or self.stream, C.c_void_p(d_src),
                                            C.c_void_p(d_dst), row_bytes, src_pitch, rows))


    def pad(self, d_src, d_dst, row_bytes, dst_pitch, rows, stream=None):
        L.check(self._lib.nnsb_pad_launch(stream or self.stream, C.c_void_p(d_src),
                                          C.c_void_p(d_dst), row_bytes, dst_pitch, rows))


class Plan:
    """nnsb_plan: parsed (mode, option) bound to one input tensor info."""

    def __init__(self, dev, mode, option, in_type, in_dim, acceleration=False):
        self.dev = dev
        self._lib = dev._lib
        code, self.opt = L.parse_option(mode, option)
        L.check(code)
        self.in_info = L.make_info(in_type, in_dim)
        self.handle = C.c_void_p()
        L.check(self._lib.nnsb_plan_create(dev.ctx, C.byref(self.opt), C.byref(self.in_info),
                                           int(bool(acceleration)), C.byref(self.handle)))
        self.out_info = L.TensorInfo()
        L.check(self._lib.nnsb_plan_out_info(self.handle, C.byref(self.out_info)))
        self.in_size = self._lib.nnsb_plan_in_size(self.handle)
        self.out_size = self._lib.nnsb_plan_out_size(self.handle)
        self.out_type = L.TYPES[self.out_info.type]
        self.out_dim = list(self.out_info.dim)

    def launch(self, d_in, d_out, batch=1, stream=None):
        L.check(self._lib.nnsb_plan_launch(self.handle, C.c_void_p(d_in), C.c_void_p(d_out),
                                           int(batch), stream or self.dev.stream))

    @property
    def last_launch_count(self):
        return self._lib.nnsb_plan_last_launch_count(self.handle)

    def describe(self):
        """One line naming the kernel path the plan compiled to."""
        buf = C.create_string_buffer(256)
        n = self._lib.nnsb_plan_describe(self.handle, buf, len(buf))
        if n < 0:
            L.check(n)
        return buf.value.decode()

    def stand_stats(self, batch, nch):
        n = int(batch) * int(nch)
        mean = (C.c_double * n)()
        std = (C.c_double * n)()
        L.check(self._lib.nnsb_plan_stand_stats(self.handle, self.dev.stream, int(batch), mean,
                                                std, n))
        return np.array(mean), np.array(std)

    def run_host(self, x, batch=1):
        """Convenience for tests: host array in -> host array out (upload, launch, download)."""
        x = np.ascontiguousarray(x)
        assert x.nbytes == self.in_size * batch, (x.nbytes, self.in_size, batch)
        d_in = self.dev.upload(x)
        d_out = self.dev.alloc(self.out_size * batch)
        try:
            self.launch(d_in, d_out, batch)
            out = self.dev.download(d_out, self.out_size * batch // np.dtype(self.out_type).itemsize,
                                    np.dtype(self.out_type))
        finally:
            self.dev.free(d_in)
            self.dev.free(d_out)
        return out

    def close(self):
        if self.handle:
            self._lib.nnsb_plan_destroy(self.handle)
            self.handle = C.c_void_p()
