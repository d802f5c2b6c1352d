"""GPU: the element mirrors end to end (set properties -> negotiate caps -> push buffers),
following tests/nnstreamer_plugins/unittest_plugins.cc (typecast :955-1118, arithmetic
:1123-1864, per-channel :1869, option change :1951, flexible :2050) and the SSAT pipelines
(multi-tensor + apply; converter 642x480).  Outputs are compared with the oracle bit for bit."""
import ctypes as C

import numpy as np
import pytest

from helpers import bits_equal, first_mismatch
from nnstreamer_b200 import BatchQueue, TensorConverter, TensorTransform, device_memory
from nnstreamer_b200 import _lib as L
from test_oracle_golden import TYPECAST_PAIRS

pytestmark = pytest.mark.gpu


def caps1(dtype, dim, rate="0/1"):
    return ("other/tensors, format=(string)static, num_tensors=(int)1, dimensions=(string)%s, "
            "types=(string)%s, framerate=(fraction)%s" % (":".join(str(d) for d in dim), dtype, rate))


def push(t, arrays, out_types):
    flow, outs = t.transform(arrays)
    assert flow == L.FLOW_OK, L.lib().nnsb_last_error()
    return [o.view(np.dtype(ty)) for o, ty in zip(outs, out_types)]


@pytest.mark.parametrize("accel", [False, True])
@pytest.mark.parametrize("pair", TYPECAST_PAIRS, ids=["%s_%s" % p for p in TYPECAST_PAIRS])
def test_typecast_harness(pair, accel):
    """TEST_TRANSFORM_TYPECAST: 3 buffers of 5 elements, value = (i + 1) * (b + 1), EXPECT_EQ."""
    it, ot = pair
    t = TensorTransform(mode="typecast", option=ot, acceleration=accel)
    out_caps = t.negotiate(caps1(it, [5]))
    assert "types=(string)%s" % ot in out_caps
    for b in range(3):
        x = np.array([(i + 1) * (b + 1) for i in range(5)], dtype=it)
        (y,) = push(t, [x], [ot])
        assert y.nbytes == 5 * np.dtype(ot).itemsize
        assert bits_equal(y, np.array([(i + 1) * (b + 1) for i in range(5)], dtype=ot))
    t.close()


@pytest.mark.parametrize("accel", [False, True])
def test_arithmetic_harness(port, accel):
    ulp = lambda a, b, dt: np.abs(a.view(dt).astype(np.int64) - b.view(dt).astype(np.int64)).max()
    i = np.arange(5)
    cases = [  # (in type, option, oracle ops, out type, expected formula)
        ("float32", "add:.5", [("add", .5)], "float32", lambda b: (i + 1) * (b + 1) + .2 + .5),
        ("float64", "mul:.5", [("mul", .5)], "float64", lambda b: ((i + 1) * (b + 1) + .2) * .5),
        ("uint8", "typecast:float32,add:.5,mul:0.2", [("typecast", "float32"), ("add", .5), ("mul", 0.2)],
         "float32", lambda b: ((i + 1) * (b + 1) + .5) * .2),
        ("uint8", "typecast:float64,add:0.2,add:0.1,typecast:uint16",
         [("typecast", "float64"), ("add", 0.2), ("add", 0.1)], "float64", lambda b: (i + 1) * (b + 1) + .3),
        ("uint8", "typecast:int32,mul:2,div:2,add:-1",
         [("typecast", "int32"), ("mul", 2), ("div", 2), ("add", -1)], "int32", lambda b: (i + 1) * (b + 1) - 1),
    ]
    for it, option, ops, ot, expected in cases:
        t = TensorTransform(mode="arithmetic", option=option, acceleration=accel)
        out_caps = t.negotiate(caps1(it, [5]))
        assert "types=(string)%s" % ot in out_caps      # trailing typecast:uint16 is ignored
        for b in range(3):
            base = (i + 1) * (b + 1)
            x = (base + .2).astype(it) if it.startswith("float") else base.astype(it)
            (y,) = push(t, [x], [ot])
            assert bits_equal(y, port.arithmetic(x, [5], ops, accel=accel)), option
            exp = expected(b).astype(ot)
            if ot.startswith("float"):
                assert ulp(y, exp, np.int32 if ot == "float32" else np.int64) <= 4     # EXPECT_FLOAT_EQ
            else:
                assert bits_equal(y, exp)
        t.close()


def test_arithmetic_per_channel_harness():
    """arithmeticPerChannel (:1869-1946): 1 2 3 4 5 -> 0 2 6 12 20."""
    opt = ("typecast:int32,per-channel:true@0,mul:1@0,mul:2@1,mul:3@2,mul:4@3,mul:5@4,"
           "add:-1@0,add:-2@1,add:-3@2,add:-4@3,add:-5@4")
    t = TensorTransform(mode="arithmetic", option=opt, acceleration=False)
    t.negotiate(caps1("uint8", [5]))
    for b in range(3):
        x = np.array([(i + 1) * (b + 1) for i in range(5)], dtype=np.uint8)
        (y,) = push(t, [x], ["int32"])
        assert list(y) == [((i + 1) * (b + 1)) * (i + 1) - (i + 1) for i in range(5)]
    t.close()


def test_change_option_at_runtime(port):
    """arithmeticChangeOptionString (:1951-2045)."""
    t = TensorTransform(mode="arithmetic", option="add:.5")
    t.negotiate(caps1("float32", [5]))
    x = np.array([(i + 1) * (i * 3 + 1) + .2 for i in range(5)], dtype=np.float32)
    (y,) = push(t, [x], ["float32"])
    assert bits_equal(y, port.arithmetic(x, [5], [("add", .5)]))
    t.set_property("option", "mul:20")
    x = np.array([(i + 1) * (i * 3 + 1) + .9 for i in range(5)], dtype=np.float32)
    (y,) = push(t, [x], ["float32"])
    assert bits_equal(y, port.arithmetic(x, [5], [("mul", 20)]))
    t.close()


def test_flexible_tensor(port):
    """arithmeticFlexTensor (:2050-2142): 128-byte header in, header out with the new type."""
    lib = L.lib()
    t = TensorTransform(mode="arithmetic", option="typecast:float32,add:.5,mul:0.2")
    flex = "other/tensors, format=(string)flexible, framerate=(fraction)0/1"
    assert t.set_caps(flex, flex)
    in_info = L.make_info("uint8", [5])
    meta = L.MetaInfo()
    assert lib.nnsb_meta_from_info(C.byref(in_info), C.byref(meta)) == 0
    hdr = (C.c_uint8 * 128)()
    assert lib.nnsb_meta_write_header(C.byref(meta), hdr) == 0
    for b in range(3):
        x = np.array([(i + 1) * (b + 1) for i in range(5)], dtype=np.uint8)
        buf = np.concatenate([np.frombuffer(bytes(hdr), dtype=np.uint8), x])
        flow, outs = t.transform([buf])
        assert flow == L.FLOW_OK and len(outs) == 1
        out = outs[0]
        assert out.nbytes == 128 + 5 * 4
        m2 = L.MetaInfo()
        assert lib.nnsb_meta_parse_header(C.byref(m2), out[:128].ctypes.data_as(C.c_void_p)) == 0
        assert m2.type == L.TYPE_ID["float32"] and m2.dimension[0] == 5 and m2.dimension[1] == 1
        y = out[128:].view(np.float32)
        assert bits_equal(y, port.arithmetic(x, [5], [("typecast", "float32"), ("add", .5), ("mul", 0.2)]))
    t.close()


def test_multi_tensor_with_apply(port):
    """tensor_mux + apply=0,2 (transform_typecast/runTest.sh:105-119): tensor 1 passes through."""
    t = TensorTransform(mode="typecast", option="float32", apply="0,2")
    caps = ('other/tensors, format=(string)static, num_tensors=(int)3, '
            'dimensions=(string)"3:4:4:1\\,2:2:2:1\\,5:1:1:1", types=(string)"uint8\\,uint8\\,uint8", '
            'framerate=(fraction)0/1')
    t.negotiate(caps)
    rng = np.random.default_rng(0)
    a, b, c = (rng.integers(0, 256, n, dtype=np.uint8) for n in (48, 8, 5))
    ya, yb, yc = push(t, [a, b, c], ["float32", "uint8", "float32"])
    assert bits_equal(ya, port.typecast(a, "float32"))
    assert bits_equal(yb, b)
    assert bits_equal(yc, port.typecast(c, "float32"))
    # the same three tensors delivered as ONE memory blob (gst_tensor_buffer_from_config)
    ya, yb, yc = push(t, [np.concatenate([a, b, c])], ["float32", "uint8", "float32"])
    assert bits_equal(ya, port.typecast(a, "float32")) and bits_equal(yb, b)
    assert bits_equal(yc, port.typecast(c, "float32"))
    # wrong number of memories is a flow error (:1700)
    flow, outs = t.transform([a, b])
    assert flow == L.FLOW_ERROR and outs == []
    t.close()


@pytest.mark.parametrize("mode,option,it,dim,oracle", [
    ("transpose", "1:2:0:3", "float32", [3, 40, 30, 2], lambda p, x, d: p.transpose(x, d, (1, 2, 0, 3))),
    ("dimchg", "0:2", "uint8", [3, 16, 16, 1], lambda p, x, d: p.dimchg(x, d, 0, 2)),
    ("clamp", "-33.3:+77.7", "float32", [50, 20], lambda p, x, d: p.clamp(x, -33.3, 77.7)),
    ("typecast", "int16", "float32", [50, 20], lambda p, x, d: p.typecast(x, "int16")),
])
def test_other_modes_through_the_element(port, mode, option, it, dim, oracle):
    t = TensorTransform(mode=mode, option=option)
    out_caps = t.negotiate(caps1(it, dim))
    rng = np.random.default_rng(1)
    n = int(np.prod(dim))
    x = (rng.standard_normal(n) * 60).astype(it) if it.startswith("float") else \
        rng.integers(0, 256, n, dtype=it)
    want = oracle(port, x, dim)
    (y,) = push(t, [x], [want.dtype.name])
    assert bits_equal(y, want), first_mismatch(y, want)
    cfg = t.config("src")
    from oracle import oracle as O
    if mode in ("transpose", "dimchg"):
        exp_dim = O.convert_dims(mode, dim, order=(1, 2, 0, 3), from_=0, to=2)
        assert list(cfg.info[0].dim) == exp_dim
    t.close()


def test_dimchg_from_gt_to_negotiates_then_fails_at_transform():
    """gsttensor_transform.c:1179-1188: caps negotiate fine, transform returns GST_FLOW_ERROR."""
    t = TensorTransform(mode="dimchg", option="2:0")
    out = t.negotiate(caps1("uint8", [3, 4, 5, 1]))
    assert "dimensions=(string)5:3:4:1" in out
    flow, outs = t.transform([np.zeros(60, dtype=np.uint8)])
    assert flow == L.FLOW_ERROR and outs == []
    t.close()


def test_stand_through_the_element(port):
    t = TensorTransform(mode="stand", option="default,per-channel:true")
    t.negotiate(caps1("float32", [3, 32, 24, 1]))
    rng = np.random.default_rng(2)
    x = (rng.random(3 * 32 * 24) * 10).astype(np.float32)
    (y,) = push(t, [x], ["float32"])
    w = port.stand(x, [3, 32, 24, 1], "default", None, True)
    assert np.all(np.abs(y.astype(np.float64) - w) <= 1e-6 * np.abs(w) + 1e-6)
    t.close()


def test_large_host_tensor_takes_the_pipelined_path(port):
    """>= 8 MiB of traffic: chunks alternate between two streams (H2D / kernel / D2H overlap).
    Per-channel arithmetic keeps its channel phase across chunk boundaries."""
    dim = [3, 1920, 1080, 1]
    n = int(np.prod(dim))
    rng = np.random.default_rng(3)
    x = rng.integers(0, 256, n, dtype=np.uint8)
    opt = ("typecast:float32,per-channel:true@0,add:-123.675@0,add:-116.28@1,add:-103.53@2,"
           "div:58.395@0,div:57.12@1,div:57.375@2")
    ops = [("typecast", "float32"), ("add", -123.675, 0), ("add", -116.28, 1), ("add", -103.53, 2),
           ("div", 58.395, 0), ("div", 57.12, 1), ("div", 57.375, 2)]
    t = TensorTransform(mode="arithmetic", option=opt)
    t.negotiate(caps1("uint8", dim, "30/1"))
    (y,) = push(t, [x], ["float32"])
    assert t.kernel_launches > 1                      # chunked
    # the oracle takes ~0.2 s per 1080p frame: compare the whole frame once
    assert bits_equal(y, port.arithmetic(x, dim, ops, True, 0))
    t2 = TensorTransform(mode="clamp", option="10:200")
    t2.negotiate(caps1("uint8", [n]))
    (y2,) = push(t2, [x], ["uint8"])
    assert bits_equal(y2, np.clip(x, 10, 200))
    t.close()
    t2.close()


def test_device_memory_in_device_memory_out(dev, port):
    """Chained CUDA-aware elements: no host round trip between typecast and arithmetic."""
    dim = [3, 64, 48, 1]
    n = int(np.prod(dim))
    rng = np.random.default_rng(4)
    x = rng.integers(0, 256, n, dtype=np.uint8)
    t1 = TensorTransform(mode="typecast", option="float32")
    mid_caps = t1.negotiate(caps1("uint8", dim))
    t2 = TensorTransform(mode="arithmetic", option="add:-127.5,div:127.5")
    t2.negotiate(mid_caps)
    d_x = dev.upload(x)
    flow, (m1,) = t1.transform([device_memory(d_x, n)])
    assert flow == L.FLOW_OK and m1.mem.on_device and m1.size == 4 * n
    flow, (m2,) = t2.transform([m1])
    assert flow == L.FLOW_OK and m2.mem.on_device
    y = dev.download(m2.ptr, n, np.float32)
    assert bits_equal(y, port.arithmetic(x, dim, [("typecast", "float32"), ("add", -127.5), ("div", 127.5)]))
    m1.release()
    m2.release()
    dev.free(d_x)
    t1.close()
    t2.close()


def test_element_memory_upload_download_round_trip():
    """The CUDA GstAllocator's mem_map / mem_unmap path: device memory <-> host bytes."""
    import ctypes as C
    t = TensorTransform(mode="typecast", option="float32")
    m = t.alloc(4096, on_device=True)
    src = np.arange(4096, dtype=np.uint8)
    assert L.lib().nnsb_transform_upload(t.handle, C.byref(m), src.ctypes.data_as(C.c_void_p)) == 0
    dst = np.zeros(4096, dtype=np.uint8)
    assert L.lib().nnsb_transform_download(t.handle, dst.ctypes.data_as(C.c_void_p), C.byref(m)) == 0
    assert bits_equal(src, dst)
    L.lib().nnsb_memory_release(C.byref(m))
    t.close()


# ---- tensor_converter ----------------------------------------------------------------------
@pytest.mark.parametrize("fmt,c,w,h", [("RGB", 3, 642, 480), ("GRAY8", 1, 642, 480), ("RGB", 3, 640, 480),
                                       ("BGRx", 4, 642, 480), ("RGB", 3, 1918, 64)])
def test_converter_video_frames(dev, port, fmt, c, w, h):
    """nnstreamer_converter/runTest.sh:34-59: 640 and 642 wide frames; 642 is the padded case."""
    conv = TensorConverter(frames_per_tensor=3)
    conv.set_caps("video/x-raw, format=(string)%s, width=(int)%d, height=(int)%d, framerate=(fraction)30/1"
                  % (fmt, w, h))
    row, pitch = c * w, (c * w + 3) // 4 * 4
    rng = np.random.default_rng(w + c)
    frames = [rng.integers(0, 256, pitch * h, dtype=np.uint8) for _ in range(6)]
    dense = [port.stride_strip(f, c, w, h) if pitch != row else f for f in frames]
    got = []
    for k, f in enumerate(frames):
        if k % 2 == 0:
            flow, out = conv.chain(f)                                # host frame: pitched H2D
        else:
            d = dev.upload(f)                                        # device frame: strip kernel
            flow, out = conv.chain(device_memory(d, f.nbytes))
            dev.free(d)
        assert flow == L.FLOW_OK
        assert (out is not None) == (k % 3 == 2)                     # a tensor every 3 frames
        if out is not None:
            assert out.size == 3 * row * h
            got.append(dev.download(out.ptr, out.size, np.uint8))
            out.release()
    assert bits_equal(got[0], np.concatenate(dense[0:3]))
    assert bits_equal(got[1], np.concatenate(dense[3:6]))
    # wrong frame size is a flow error
    flow, out = conv.chain(frames[0][:-4])
    assert flow == L.FLOW_ERROR
    conv.close()


def test_converter_to_transform_chain_on_device(dev, port):
    """videotestsrc-like frames -> tensor_converter frames-per-tensor=4 -> tensor_transform
    (BASELINE configs[0] chain), everything after the upload stays in HBM."""
    w, h, n_frames = 224, 224, 4
    conv = TensorConverter(frames_per_tensor=n_frames)
    caps = conv.set_caps("video/x-raw, format=(string)RGB, width=(int)%d, height=(int)%d, framerate=(fraction)30/1" % (w, h))
    t = TensorTransform(mode="arithmetic", option="typecast:float32,add:-127.5,div:127.5")
    t.negotiate(caps)
    rng = np.random.default_rng(9)
    frames = [rng.integers(0, 256, 3 * w * h, dtype=np.uint8) for _ in range(n_frames)]
    out = None
    for f in frames:
        flow, out = conv.chain(f)
        assert flow == L.FLOW_OK
    assert out is not None
    flow, (y,) = t.transform([out])
    assert flow == L.FLOW_OK
    got = dev.download(y.ptr, 3 * w * h * n_frames, np.float32)
    want = port.arithmetic(np.concatenate(frames), [3, w, h, n_frames],
                           [("typecast", "float32"), ("add", -127.5), ("div", 127.5)])
    assert bits_equal(got, want)
    out.release()
    y.release()
    conv.close()
    t.close()


# ---- batching queue ----------------------------------------------------------------------
def test_batching_queue_groups_frames(dev, port):
    dim = [3, 64, 48, 1]
    n = int(np.prod(dim))
    plan = dev.plan("stand", "default", "float32", dim)
    q = BatchQueue(dev, plan, 4)
    rng = np.random.default_rng(6)
    frames = [(rng.random(n) * (k + 1)).astype(np.float32) for k in range(10)]
    got = []
    for k, f in enumerate(frames):
        full = q.push(f if k % 2 == 0 else dev.upload(f))
        if full:
            d_out, cnt, stream = q.flush()
            assert cnt == 4
            dev.sync(stream)
            got.append(dev.download(d_out, cnt * n, np.float32))
    d_out, cnt, stream = q.flush()                  # partial batch at end of stream
    assert cnt == 2
    dev.sync(stream)
    got.append(dev.download(d_out, cnt * n, np.float32))
    assert q.flush()[1] == 0
    y = np.concatenate(got)
    w = np.concatenate([port.stand(f, dim, "default") for f in frames])   # per-frame statistics
    assert np.all(np.abs(y.astype(np.float64) - w) <= 1e-6 * np.abs(w) + 1e-6)
    q.close()
    plan.close()
