"""The allocation-free stand entry point, nnsb_stand_launch_noalloc().

Who needs it: users whose frame-processing path must not call cudaMalloc -- the GPU's memory
may already be fully reserved by an inference runtime's pool (a late allocation can fail),
the streaming thread may have a per-frame deadline (cudaMalloc has unbounded latency), or
the caller is a plug-in that only receives a CUDA stream and device pointers.  The entry
point runs the same kernels as a stand plan but keeps its scratch in a fixed 1 MiB array
inside the library instead of the context's device pool.

What is checked here, through the C-ABI on a real GPU:
  * two calls in a row (different data, so stale scratch from the first call would show up in
    the second) each reproduce nnsb_plan_launch() bit for bit -- same kernels, same launch
    geometry, hence the tolerance is zero, tighter than the 1e-6 the stand parity tests use;
  * a batch x channel count that does not fit the array is refused with NNSB_ERANGE;
  * the file sorts last (pytest collects files alphabetically) so a `-x` run has already
    finished every other test, and it creates everything it uses, so it also passes alone in
    a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _dims(dim):
    d = list(dim) + [1] * (8 - len(dim))
    return (C.c_uint32 * 8)(*d)


def _noalloc(L, lib, dev, stream, x, dim, mode, per_channel, batch, out_type=None):
    """Upload x, run the allocation-free entry point on `stream`, download the result."""
    in_type = L.TYPE_ID[x.dtype.name]
    o_type = L.TYPE_ID[out_type] if out_type else L.TYPE_END
    o_np = np.dtype(out_type) if out_type else x.dtype
    d_in = dev.upload(x)                       # buffers come from the caller, as documented
    d_out = dev.alloc(x.size * o_np.itemsize)
    code = lib.nnsb_stand_launch_noalloc(C.c_void_p(d_in), C.c_void_p(d_out), in_type, o_type,
                                         _dims(dim), 0 if mode == "default" else 1,
                                         int(per_channel), batch, None)   # NULL = default stream
    assert code == 0, lib.nnsb_last_error()
    L.check(lib.nnsb_stream_sync(stream))
    out = dev.download(d_out, x.size, o_np)
    dev.free(d_in)
    dev.free(d_out)
    return out


@pytest.mark.parametrize("mode,per_channel,out_type", [
    ("default", False, None),
    ("default", True, None),
    ("dc-average", True, "float64"),
])
def test_noalloc_matches_plan_launch_twice_in_a_row(mode, per_channel, out_type):
    from nnstreamer_b200 import Device
    from nnstreamer_b200 import _lib as L
    lib = L.lib()
    dim, batch = [3, 64, 48, 1], 4
    n = 3 * 64 * 48
    rng = np.random.default_rng(123)
    with Device(0) as dev:
        # the entry point works on the default stream of the current device; wrap that
        # stream once so the test can wait for it through the C-ABI
        default_stream = C.c_void_p()
        L.check(lib.nnsb_stream_wrap(dev.ctx, None, C.byref(default_stream)))
        opt = mode + ((":" + out_type) if out_type else "") + (",per-channel:true" if per_channel else "")
        plan = dev.plan("stand", opt, "float32", dim)
        for call in range(2):
            # second call: other values and another scale, so leftovers would be visible
            x = (rng.random(n * batch) * (10.0 + 90.0 * call) - 5.0 * call).astype(np.float32)
            want = plan.run_host(x, batch=batch)
            dev.sync()
            got = _noalloc(L, lib, dev, default_stream, x, dim, mode, per_channel, batch, out_type)
            assert got.dtype == want.dtype
            assert got.tobytes() == want.tobytes(), "call %d differs from nnsb_plan_launch" % call
        plan.close()
        lib.nnsb_stream_destroy(default_stream)


def test_noalloc_rejects_what_does_not_fit():
    from nnstreamer_b200 import Device
    from nnstreamer_b200 import _lib as L
    lib = L.lib()
    with Device(0) as dev:
        # 70000 channels x batch 2: the {mean, std} table alone is 2.2 MB > 1 MiB
        dim = [70000, 2, 1, 1]
        d_in = dev.alloc(70000 * 2 * 2 * 4)
        d_out = dev.alloc(70000 * 2 * 2 * 4)
        code = lib.nnsb_stand_launch_noalloc(C.c_void_p(d_in), C.c_void_p(d_out),
                                             L.TYPE_ID["float32"], L.TYPE_END, _dims(dim), 0, 1, 2,
                                             None)
        assert code == L.ERANGE
        msg = lib.nnsb_last_error().decode()
        assert "scratch" in msg and str(L.NOALLOC_SCRATCH_BYTES) in msg
        dev.free(d_in)
        dev.free(d_out)
