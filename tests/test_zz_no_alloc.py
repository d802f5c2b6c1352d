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
    geometry, hence tolerance zero -- AND match the CPU reference within the 1e-6 the stand
    parity tests use;
  * the branch with the smaller grid (taken when the plan path's block partials would not fit
    the array) sums in another order, so it is compared with the CPU reference only, within
    the same tolerance; so is an odd channel count (the non-specialised kernels);
  * after one warm-up call per variant, further calls leave cudaMemGetInfo exactly where it
    was (CUDA loads device code lazily: the FIRST call is allowed to, and does, take memory);
  * what does not fit (scratch, batch > 65535) is refused with NNSB_ERANGE and a clear message;
  * the file sorts last (pytest collects files alphabetically) so a `-x` run has already
    finished every other test, and it creates everything it uses, so it also passes alone in
    a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def default_grid(monkeypatch):
    """The bit-identity checks assume the library's default launch geometry."""
    monkeypatch.delenv("NNSB_BLOCKS_PER_SM", raising=False)


def _dims(dim):
    d = list(dim) + [1] * (8 - len(dim))
    return (C.c_uint32 * 8)(*d)


def _launch(L, lib, d_in, d_out, in_type, o_type, dim, mode, per_channel, batch):
    return lib.nnsb_stand_launch_noalloc(C.c_void_p(d_in), C.c_void_p(d_out), in_type, o_type,
                                         _dims(dim), 0 if mode == "default" else 1,
                                         int(per_channel), batch, None)   # NULL = default stream


def _noalloc(L, lib, dev, stream, x, dim, mode, per_channel, batch, out_type=None):
    """Upload x, run the allocation-free entry point on `stream`, download the result."""
    in_type = L.TYPE_ID[x.dtype.name]
    o_type = L.TYPE_ID[out_type] if out_type else L.TYPE_END
    o_np = np.dtype(out_type) if out_type else x.dtype
    d_in = dev.upload(x)                       # buffers come from the caller, as documented
    d_out = dev.alloc(x.size * o_np.itemsize)
    code = _launch(L, lib, d_in, d_out, in_type, o_type, dim, mode, per_channel, batch)
    assert code == 0, lib.nnsb_last_error()
    L.check(lib.nnsb_stream_sync(stream))
    out = dev.download(d_out, x.size, o_np)
    dev.free(d_in)
    dev.free(d_out)
    return out


def _close_to_reference(got, want, scale):
    g = got.astype(np.float64)
    w = want.astype(np.float64)
    tol = 1e-6 * np.abs(w) + 1e-6 * scale       # the tolerance of tests/test_gpu_parity.py
    bad = np.abs(g - w) > tol
    assert not bad.any(), "%d of %d beyond tolerance; worst %g" % (bad.sum(), bad.size,
                                                                   np.abs(g - w).max())


def _reference(port, x, dim, mode, out_type, per_channel, batch):
    n = x.size // batch
    return np.concatenate([port.stand(x[i * n:(i + 1) * n], dim, mode, out_type, per_channel)
                           for i in range(batch)])


@pytest.mark.parametrize("mode,per_channel,out_type", [
    ("default", False, None),
    ("default", True, None),
    ("dc-average", True, "float64"),
])
def test_noalloc_matches_plan_launch_twice_in_a_row(port, mode, per_channel, out_type):
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
            _close_to_reference(got, _reference(port, x, dim, mode, out_type, per_channel, batch),
                                100.0 if mode == "dc-average" else 1.0)
        plan.close()
        lib.nnsb_stream_destroy(default_stream)


def test_noalloc_small_grid_branch_and_odd_channels_match_the_reference(port):
    """batch 16 x 8 channels x 4.26 M elements: the plan path's grid would keep 520 blocks per
    tensor = 1.06 MB of partial sums, more than the array holds, so the entry point drops to four
    blocks per SM.  That is another summation order than nnsb_plan_launch's, hence the check is
    against the CPU reference.  Then 5 channels (no specialised kernel) on a small tensor."""
    from nnstreamer_b200 import Device
    from nnstreamer_b200 import _lib as L
    lib = L.lib()
    rng = np.random.default_rng(77)
    with Device(0) as dev:
        default_stream = C.c_void_p()
        L.check(lib.nnsb_stream_wrap(dev.ctx, None, C.byref(default_stream)))
        dim, batch = [8, 1024, 520, 1], 16
        n = int(np.prod(dim))
        assert 16 * batch * 8 * (-(-n // 8192)) > L.NOALLOC_SCRATCH_BYTES     # default grid: too big
        # every frame gets its own offset and spread, every channel its own bias
        x = (rng.standard_normal(n * batch, dtype=np.float32).reshape(batch, -1, 8)
             * np.linspace(0.5, 40.0, batch, dtype=np.float32)[:, None, None]
             + np.arange(8, dtype=np.float32)[None, None, :] * 3
             + np.linspace(-100.0, 100.0, batch, dtype=np.float32)[:, None, None]).reshape(-1)
        for mode, pc in (("default", True), ("dc-average", True)):
            got = _noalloc(L, lib, dev, default_stream, x, dim, mode, pc, batch)
            _close_to_reference(got, _reference(port, x, dim, mode, None, pc, batch),
                                200.0 if mode == "dc-average" else 1.0)
        dim, batch = [5, 33, 17, 1], 3
        n = int(np.prod(dim))
        x = (rng.random(n * batch) * 50 - 7).astype(np.float32)
        for call in range(2):
            got = _noalloc(L, lib, dev, default_stream, x + call, dim, "default", True, batch)
            _close_to_reference(got, _reference(port, x + call, dim, "default", None, True, batch), 1.0)
        lib.nnsb_stream_destroy(default_stream)


def test_noalloc_warmed_up_calls_take_no_device_memory():
    """CUDA loads device code lazily: the first call of a variant may (and does) put the array
    and the kernels on the GPU.  From the second call on, free device memory must not move."""
    import torch
    from nnstreamer_b200 import Device
    from nnstreamer_b200 import _lib as L
    lib = L.lib()
    dim, batch = [3, 320, 240, 1], 8
    n = int(np.prod(dim)) * batch
    with Device(0) as dev:
        default_stream = C.c_void_p()
        L.check(lib.nnsb_stream_wrap(dev.ctx, None, C.byref(default_stream)))
        x = np.random.default_rng(5).random(n).astype(np.float32)
        d_in = dev.upload(x)
        d_out = dev.alloc(n * 4)
        f32 = L.TYPE_ID["float32"]
        assert _launch(L, lib, d_in, d_out, f32, L.TYPE_END, dim, "default", True, batch) == 0   # warm-up
        L.check(lib.nnsb_stream_sync(default_stream))
        torch.cuda.synchronize()
        torch.cuda.mem_get_info(0)                      # torch's own first-use allocations
        before = torch.cuda.mem_get_info(0)[0]
        for _ in range(50):
            assert _launch(L, lib, d_in, d_out, f32, L.TYPE_END, dim, "default", True, batch) == 0
        L.check(lib.nnsb_stream_sync(default_stream))
        after = torch.cuda.mem_get_info(0)[0]
        assert after == before, "free device memory moved by %d bytes" % (before - after)
        dev.free(d_in)
        dev.free(d_out)
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
        # more tensors than one launch has grid rows for
        code = lib.nnsb_stand_launch_noalloc(C.c_void_p(d_in), C.c_void_p(d_out),
                                             L.TYPE_ID["float32"], L.TYPE_END, _dims([1, 1, 1, 1]),
                                             0, 0, 65536, None)
        assert code == L.ERANGE and "65535" in lib.nnsb_last_error().decode()
        dev.free(d_in)
        dev.free(d_out)
