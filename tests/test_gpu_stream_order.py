"""Stream-ordered device memories (include/nnstreamer_b200.h, nnsb_memory): elements return as
soon as their kernels are queued; blocks carry the event after which they are complete, pooled
blocks are only recycled once the work that touched them has finished, and references / shares
keep a block alive (gst_memory_ref / gst_memory_share, nnstreamer_plugin_api_impl.c:534).

Everything goes through the C-ABI; results are compared with the CPU reference."""
import ctypes as C
import threading

import numpy as np
import pytest

from helpers import bits_equal
from nnstreamer_b200 import (Branch, DeviceMemory, TensorConverter, TensorFilter, TensorTransform,
                             device_memory)
from nnstreamer_b200 import _lib as L

pytestmark = pytest.mark.gpu

VIDEO = "video/x-raw, format=(string)RGB, width=(int)%d, height=(int)%d, framerate=(fraction)30/1"


def caps_for(dim, type_name):
    return "other/tensor, dimension=(string)%s, type=(string)%s, framerate=(fraction)30/1" % (
        ":".join(str(d) for d in dim), type_name)


def upload(elem, x):
    """A pooled device memory of `elem` holding x."""
    m = elem.alloc(x.nbytes, on_device=True)
    assert elem._lib.nnsb_transform_upload(elem.handle, C.byref(m), x.ctypes.data_as(C.c_void_p)) == 0
    return DeviceMemory(m)


def test_chained_elements_do_not_need_a_host_sync(dev, port):
    """converter -> typecast -> arithmetic -> transpose, 40 frames queued back to back on a
    1080p stream, only the last output is waited for; every 8th output is kept and checked."""
    w, h = 1920, 1080
    conv = TensorConverter()
    c1 = conv.set_caps(VIDEO % (w, h))
    t1 = TensorTransform(mode="typecast", option="float32")
    t2 = TensorTransform(mode="arithmetic", option="add:-127.5,div:127.5")
    t3 = TensorTransform(mode="transpose", option="1:2:0:3")
    t3.negotiate(t2.negotiate(t1.negotiate(c1)))
    rng = np.random.default_rng(5)
    frames = [rng.integers(0, 256, 3 * w * h, dtype=np.uint8) for _ in range(5)]
    src = TensorTransform(mode="typecast", option="uint8")
    kept = []
    for i in range(40):
        f = upload(src, frames[i % 5])
        flow, m0 = conv.chain(f)
        assert flow == 0
        flow, (m1,) = t1.transform([m0])
        flow, (m2,) = t2.transform([m1])
        flow, (m3,) = t3.transform([m2])
        assert flow == 0
        for m in (f, m0, m1, m2):                    # dropped while the kernels may still be queued
            m.release()
        if i % 8 == 7:
            kept.append((i, m3))
        else:
            m3.release()
    ops = [("typecast", "float32"), ("add", -127.5), ("div", 127.5)]
    for i, m3 in kept:
        want = port.transpose(port.arithmetic(frames[i % 5], [3, w, h, 1], ops), [3, w, h, 1], (1, 2, 0, 3))
        got = dev.download(m3.ptr, 3 * w * h, np.float32)       # .ptr waits for the producer
        assert bits_equal(got, want), "frame %d" % i
        m3.release()
    # nothing above forced the pools to grow without bound: a handful of blocks per element
    for e in (t1, t2, t3):
        e.close()
    conv.close()
    src.close()


def test_branch_pushes_a_frame_through_all_elements_in_one_call(dev, port):
    """nnsb_branch: converter (frames-per-tensor 2, padded 642-wide rows) -> typecast -> arithmetic
    -> transpose with ONE call per frame; the result equals the element-by-element chain's, the
    converter's "not yet" comes back as zero outputs, nothing waits for the GPU in between."""
    w, h = 642, 480
    conv = TensorConverter(frames_per_tensor=2)
    c1 = conv.set_caps(VIDEO % (w, h))
    t1 = TensorTransform(mode="typecast", option="float32")
    t2 = TensorTransform(mode="arithmetic", option="add:-127.5,div:127.5")
    t3 = TensorTransform(mode="transpose", option="1:2:0:3")
    t3.negotiate(t2.negotiate(t1.negotiate(c1)))
    br = Branch(conv, [t1, t2, t3])
    rng = np.random.default_rng(11)
    pitch = (3 * w + 3) // 4 * 4
    src = TensorTransform(mode="typecast", option="uint8")
    frames, outs = [], []
    for i in range(6):
        f = rng.integers(0, 256, pitch * h, dtype=np.uint8)
        frames.append(f)
        m = upload(src, f)
        flow, res = br.push(m)
        m.release()
        assert flow == 0
        assert len(res) == (1 if i % 2 else 0)
        outs.extend(res)
    ops = [("typecast", "float32"), ("add", -127.5), ("div", 127.5)]
    for k, m3 in enumerate(outs):
        dense = np.concatenate([port.stride_strip(frames[2 * k + j], 3, w, h) for j in range(2)])
        want = port.transpose(port.arithmetic(dense, [3, w, h, 2], ops), [3, w, h, 2], (1, 2, 0, 3))
        got = dev.download(m3.ptr, 3 * w * h * 2, np.float32)
        assert bits_equal(got, want), "tensor %d" % k
        m3.release()
    # host frames go the same way (the converter uploads them)
    flow, res = br.push(frames[0])
    assert flow == 0 and res == []
    flow, res = br.push(frames[1])
    assert flow == 0 and len(res) == 1
    dense = np.concatenate([port.stride_strip(frames[j], 3, w, h) for j in range(2)])
    want = port.transpose(port.arithmetic(dense, [3, w, h, 2], ops), [3, w, h, 2], (1, 2, 0, 3))
    assert bits_equal(dev.download(res[0].ptr, 3 * w * h * 2, np.float32), want)
    res[0].release()
    # transforms only: a tensor buffer in, a tensor buffer out
    b2 = Branch(None, [t1, t2])
    x = rng.integers(0, 256, 3 * w * h * 2, dtype=np.uint8)
    mx = upload(src, x)
    flow, res = b2.push([mx])
    assert flow == 0 and len(res) == 1
    assert bits_equal(dev.download(res[0].ptr, x.size, np.float32), port.arithmetic(x, [3, w, h, 2], ops))
    res[0].release()
    mx.release()
    # a size mismatch is a flow error, like the element's
    flow, res = br.push(np.zeros(10, np.uint8))
    assert flow != 0 and res == []
    for b in (br, b2):
        b.close()
    for e in (t1, t2, t3, src):
        e.close()
    conv.close()


def test_pool_never_recycles_a_block_that_is_still_being_read(dev, port):
    """8 threads x their own element: alloc / launch / release in a tight loop without waiting.
    A block handed out again while a queued kernel still reads or writes it would corrupt the
    outputs; every output is checked bit for bit."""
    dim = [3, 640, 360, 1]
    n = int(np.prod(dim))
    errors = []

    def worker(seed):
        try:
            rng = np.random.default_rng(seed)
            t = TensorTransform(mode="arithmetic", option="typecast:float32,add:%d,mul:0.5" % seed)
            t.negotiate(caps_for(dim, "uint8"))
            xs = [rng.integers(0, 256, n, dtype=np.uint8) for _ in range(4)]
            wants = [port.arithmetic(x, dim, [("typecast", "float32"), ("add", seed), ("mul", 0.5)]) for x in xs]
            outs = []
            for i in range(120):
                f = upload(t, xs[i % 4])
                flow, (y,) = t.transform([f])
                assert flow == 0
                f.release()                          # input block goes back while the kernel is queued
                outs.append((i, y))
                if len(outs) == 6:                   # keep a few in flight, check the oldest
                    j, o = outs.pop(0)
                    got = np.empty(n, dtype=np.float32)
                    assert t._lib.nnsb_transform_download(t.handle, got.ctypes.data_as(C.c_void_p),
                                                          C.byref(o.mem)) == 0
                    if not bits_equal(got, wants[j % 4]):
                        errors.append("thread %d frame %d" % (seed, j))
                    o.release()
            for _, o in outs:
                o.release()
            t.close()
        except Exception as exc:                     # noqa: BLE001
            errors.append("thread %d: %r" % (seed, exc))

    ths = [threading.Thread(target=worker, args=(s,)) for s in range(1, 9)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not errors, errors[:5]


def test_steady_state_reuses_pooled_blocks(dev):
    """After warm-up a long asynchronous stream of frames makes no further cudaMalloc."""
    import torch
    dim = [3, 640, 480, 1]
    n = int(np.prod(dim))
    t = TensorTransform(mode="typecast", option="float32")
    t.negotiate(caps_for(dim, "uint8"))
    src = upload(t, np.random.default_rng(1).integers(0, 256, n, dtype=np.uint8))
    last = None
    for _ in range(50):                              # same pattern as below: the pool reaches its depth
        flow, (y,) = t.transform([src])
        if last is not None:
            last.release()
        last = y
    torch.cuda.synchronize()
    before = torch.cuda.mem_get_info(0)[0]
    for _ in range(3000):
        flow, (y,) = t.transform([src])
        if last is not None:
            last.release()
        last = y
    last.sync()
    after = torch.cuda.mem_get_info(0)[0]
    last.release()
    src.release()
    t.close()
    assert before == after, "device memory moved by %d bytes" % (before - after)


def test_memory_outlives_its_element(dev, port):
    """a memory still held downstream when its element is freed must stay valid."""
    dim = [3, 64, 48, 1]
    x = np.random.default_rng(2).integers(0, 256, int(np.prod(dim)), dtype=np.uint8)
    t = TensorTransform(mode="typecast", option="float32")
    t.negotiate(caps_for(dim, "uint8"))
    flow, (y,) = t.transform([upload(t, x)])
    assert flow == 0
    part = y.share(4 * 100, 4 * 50)
    t.close()                                        # context destroyed, pools trimmed
    junk = [dev.alloc(x.size * 4) for _ in range(4)]
    for p in junk:
        L.check(L.lib().nnsb_dev_memset_async(dev.stream, C.c_void_p(p), 0x5A, x.size * 4))
    dev.sync()
    want = port.typecast(x, "float32")
    assert bits_equal(dev.download(y.ptr, x.size, np.float32), want)
    y.release()
    assert bits_equal(dev.download(part.ptr, 50, np.float32), want[100:150])
    part.release()
    for p in junk:
        dev.free(p)


def test_converter_frames_per_tensor_changed_mid_batch(dev, port):
    """raising frames-per-tensor while a batch is partly filled used to write past the
    block sized for the old count.  642x480 RGB (padded rows, the reference's own test size)."""
    w, h = 642, 480
    conv = TensorConverter(frames_per_tensor=2)
    conv.set_caps(VIDEO % (w, h))
    pitch = (3 * w + 3) // 4 * 4
    rng = np.random.default_rng(9)
    frames = [rng.integers(0, 256, pitch * h, dtype=np.uint8) for _ in range(8)]
    dense = [port.stride_strip(f, 3, w, h) for f in frames]
    flow, out = conv.chain(frames[0])
    assert flow == 0 and out is None                 # 1 of 2
    conv.set_property("frames-per-tensor", 4)        # grow: the waiting frame moves along
    assert conv.config().info[0].dim[3] == 4
    got = None
    for i in (1, 2, 3):
        flow, out = conv.chain(frames[i])
        assert flow == 0
        if out is not None:
            got = out
    assert got is not None and got.size == 4 * 3 * w * h
    assert dev.download(got.ptr, got.size, np.uint8).tobytes() == np.concatenate(dense[0:4]).tobytes()
    got.release()
    # shrink with frames waiting: 3 of 4 are in, new size 2 -> the oldest one is kept
    for i in (4, 5, 6):
        flow, out = conv.chain(frames[i])
        assert flow == 0 and out is None
    conv.set_property("frames-per-tensor", 2)
    flow, out = conv.chain(frames[7])
    assert flow == 0 and out is not None and out.size == 2 * 3 * w * h
    assert dev.download(out.ptr, out.size, np.uint8).tobytes() == np.concatenate([dense[4], dense[7]]).tobytes()
    out.release()
    conv.close()


def test_pageable_host_inputs_go_through_the_pinned_staging_ring(dev, port):
    """An upstream GstMemory is ordinary malloc'd memory.  24.9 MB tensors from pageable numpy
    arrays: the element-wise path (chunks on two streams, staged by helper threads) and a
    permutation (whole-tensor upload through the ring), both against the CPU reference; twice, so
    that recycled staging slots and events are exercised, and with an odd-sized tail."""
    dim = [3, 1920, 1080, 4]
    n = int(np.prod(dim))
    rng = np.random.default_rng(21)
    t = TensorTransform(mode="arithmetic", option="typecast:float32,per-channel:true@0,add:-1@0,div:3@1,mul:2@2")
    t.negotiate(caps_for(dim, "uint8"))
    ops = [("typecast", "float32"), ("add", -1, 0), ("div", 3, 1), ("mul", 2, 2)]
    for _ in range(2):
        x = rng.integers(0, 256, n, dtype=np.uint8)          # pageable
        flow, (y,) = t.transform([x])
        assert flow == 0
        assert bits_equal(y.view(np.float32), port.arithmetic(x, dim, ops, True, 0))
    t.close()
    dim = [3, 1283, 1009, 2]                                 # 7.7 MB, not a multiple of the 4 MiB pieces
    n = int(np.prod(dim))
    t = TensorTransform(mode="transpose", option="1:2:0:3")
    t.negotiate(caps_for(dim, "uint8"))
    for _ in range(2):
        x = rng.integers(0, 256, n, dtype=np.uint8)
        flow, (y,) = t.transform([x])
        assert flow == 0
        assert bits_equal(y, port.transpose(x, dim, (1, 2, 0, 3)))
    t.close()


# ---------------------------------------------------------------------------------------
# hand-off to tensor_filter
# ---------------------------------------------------------------------------------------
def test_filter_consumes_the_chain_output_without_leaving_hbm(dev, port):
    """converter -> typecast -> arithmetic -> transpose -> tensor_filter whose sub-plugin is a CUDA
    kernel (this library's clamp, enqueued by the callback on the stream it is given).  The
    filter element copies zero bytes; the result equals the CPU reference chain + clamp."""
    w, h = 640, 360
    conv = TensorConverter()
    c1 = conv.set_caps(VIDEO % (w, h))
    t1 = TensorTransform(mode="typecast", option="float32")
    t2 = TensorTransform(mode="arithmetic", option="add:-127.5,div:127.5")
    t3 = TensorTransform(mode="transpose", option="1:2:0:3")
    t3.negotiate(t2.negotiate(t1.negotiate(c1)))
    out_dim = [w, h, 3, 1]
    n = 3 * w * h
    lib = L.lib()
    plan = dev.plan("clamp", "-0.25:0.5", "float32", out_dim)
    wrapped = {}

    def invoke(inputs, outputs, cuda_stream):
        # a CUDA-aware sub-plugin: enqueue on the element's stream, never synchronise
        if cuda_stream not in wrapped:
            s = C.c_void_p()
            L.check(lib.nnsb_stream_wrap(dev.ctx, C.c_void_p(cuda_stream), C.byref(s)))
            wrapped[cuda_stream] = s
        assert inputs[0][1] == 4 * n and outputs[0][1] == 4 * n
        plan.launch(inputs[0][0], outputs[0][0], 1, stream=wrapped[cuda_stream])
        return 0

    cfg = TensorFilter.make_config([("float32", out_dim)], rate=(30, 1))
    filt = TensorFilter(invoke, cfg, cfg)
    rng = np.random.default_rng(11)
    src = TensorTransform(mode="typecast", option="uint8")
    results = []
    frames = [rng.integers(0, 256, n, dtype=np.uint8) for _ in range(6)]
    for x in frames:
        f = upload(src, x)
        flow, m0 = conv.chain(f)
        flow, (m1,) = t1.transform([m0])
        flow, (m2,) = t2.transform([m1])
        flow, (m3,) = t3.transform([m2])
        flow, (y,) = filt.transform([m3])
        assert flow == 0
        for m in (f, m0, m1, m2, m3):
            m.release()
        results.append(y)
    assert filt.counters == (0, 0, 6)                # zero bytes up, zero bytes down, 6 invokes
    ops = [("typecast", "float32"), ("add", -127.5), ("div", 127.5)]
    for x, y in zip(frames, results):
        chw = port.transpose(port.arithmetic(x, [3, w, h, 1], ops), [3, w, h, 1], (1, 2, 0, 3))
        want = port.clamp(chw, -0.25, 0.5)
        assert bits_equal(dev.download(y.ptr, n, np.float32), want)
        y.release()
    # size check of tensor_filter.c:719-735, and the drop / error conventions
    flow, outs = filt.transform([np.zeros(10, dtype=np.float32)])
    assert flow == L.FLOW_ERROR and not outs
    filt2 = TensorFilter(lambda i, o, s: 1, cfg, cfg)
    flow, outs = filt2.transform([np.zeros(n, dtype=np.float32)])
    assert flow == L.FLOW_DROPPED and not outs
    assert filt2.counters[0] == 4 * n                # a host tensor is uploaded, and counted
    filt3 = TensorFilter(lambda i, o, s: -3, cfg, cfg)
    flow, outs = filt3.transform([np.zeros(n, dtype=np.float32)])
    assert flow == L.FLOW_ERROR
    for s in wrapped.values():
        lib.nnsb_stream_destroy(s)
    plan.close()
    for e in (filt, filt2, filt3, conv, t1, t2, t3, src):
        e.close()


def test_filter_sub_plugin_can_be_a_torch_op_on_the_exported_pointer(dev, port):
    """What a TensorRT / torch sub-plugin would do: wrap the device pointers it is handed and run
    on the stream it is given (here: per-plane mean of the CHW tensor with torch)."""
    import torch
    w, h = 320, 240
    n = 3 * w * h
    t = TensorTransform(mode="arithmetic", option="typecast:float32,div:255.0")
    t.negotiate(caps_for([w, h, 3, 1], "uint8"))

    class View:
        def __init__(self, ptr, nbytes):
            self.__cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<f4",
                                             "data": (int(ptr), False), "version": 3, "strides": None}

    def invoke(inputs, outputs, cuda_stream):
        with torch.cuda.stream(torch.cuda.ExternalStream(cuda_stream)):
            x = torch.as_tensor(View(*inputs[0]), device="cuda:0").view(3, h * w)
            y = torch.as_tensor(View(*outputs[0]), device="cuda:0")
            y.copy_(x.mean(dim=1))
        return 0

    filt = TensorFilter(invoke, TensorFilter.make_config([("float32", [w, h, 3, 1])]),
                        TensorFilter.make_config([("float32", [3, 1, 1, 1])]))
    x = np.random.default_rng(4).integers(0, 256, n, dtype=np.uint8)
    flow, (m,) = t.transform([upload(t, x)])
    flow, (y,) = filt.transform([m])
    assert flow == 0
    ptr, size, device_id, event = y.export()
    assert size == 12 and device_id == 0
    got = dev.download(y.ptr, 3, np.float32)
    want = port.arithmetic(x, [w, h, 3, 1], [("typecast", "float32"), ("div", 255.0)]).reshape(3, -1)
    np.testing.assert_allclose(got, want.astype(np.float64).mean(axis=1), rtol=1e-5)
    assert filt.counters[:2] == (0, 0)
    m.release()
    y.release()
    filt.close()
    t.close()
