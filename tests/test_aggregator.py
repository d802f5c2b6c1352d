"""tensor_aggregator: oracle vs the reference's known-answer vectors and
host-side logic on CPU; the device adapter + concatenation against the oracle on the GPU.

Reference: gst/nnstreamer/elements/gsttensor_aggregator.c; golden vectors extracted from
tests/nnstreamer_plugins/unittest_plugins.cc (testTensorAggregator aggregate1..10, flushData) by
tests/golden/make_aggregator_golden.py."""
import os

import numpy as np
import pytest

import nnstreamer_b200 as nb
from nnstreamer_b200 import NnsbError, TensorAggregator
from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = np.load(os.path.join(HERE, "golden", "aggregator_golden.npz"))
CASES = ["aggregate%d" % i for i in range(1, 11)]


def caps_for(dim, type_name="int32", rate="0/1"):
    return "other/tensor, dimension=(string)%s, type=(string)%s, framerate=(fraction)%s" % (
        ":".join(str(d) for d in dim), type_name, rate)


@pytest.fixture(scope="module")
def port():
    return O.port()


# ---------------------------------------------------------------------------------------
# CPU: the oracle reproduces the reference's own test vectors
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_unit_tests(port, name):
    fi, fo, ff, fd, concat = (int(v) for v in GOLD[name + "/props"])
    dim = [int(d) for d in GOLD[name + "/dim"]]
    model = O.AggregatorModel(port, dim, "int32", fi, fo, ff, fd, bool(concat))
    outs = []
    for i in range(2):
        outs += model.chain(GOLD["frames"][i])
    got = np.concatenate(outs).view(np.int32)
    assert np.array_equal(got, GOLD[name + "/expected"])
    assert len(outs) == (2 if name + "/per_buffer" in GOLD else 1)


def test_oracle_flush_data(port):
    fi, fo, ff, fd, concat = (int(v) for v in GOLD["flush/props"])
    data = GOLD["flush/data"]
    model = O.AggregatorModel(port, [10], "int32", fi, fo, ff, fd, bool(concat))
    outs = model.chain(data)
    assert len(outs) == 1 and np.array_equal(outs[0].view(np.int32), data[:6])
    model.reset()                                   # flush-start / flush-stop
    outs = model.chain(data)
    assert len(outs) == 1 and np.array_equal(outs[0].view(np.int32), data[:6])


def test_oracle_concat_is_a_transpose(port):
    """The block interleave equals moving the frame axis inside the axes above frames-dim."""
    rng = np.random.default_rng(7)
    dim = [3, 4, 5, 6]
    for fd in range(4):
        for fo in (2, 3):
            frames = rng.integers(0, 255, (fo,) + tuple(reversed(dim)), dtype=np.uint8)
            block = int(np.prod(dim[:fd + 1]))
            got = port.aggregator_concat(frames.reshape(-1), int(np.prod(dim)), fo, block)
            hi = int(np.prod(dim[fd + 1:]))
            want = frames.reshape(fo, hi, block).transpose(1, 0, 2).reshape(-1)
            assert np.array_equal(got, want)


# ---------------------------------------------------------------------------------------
# CPU: properties and caps of the element mirror (no device work)
# ---------------------------------------------------------------------------------------
def test_properties_defaults_and_round_trip():
    a = TensorAggregator()
    # gsttensor_aggregator.c:76-101
    assert a.get_property("frames-in") == 1 and a.get_property("frames-out") == 1
    assert a.get_property("frames-flush") == 0 and a.get_property("frames-dim") == 7
    assert a.get_property("concat") is True and a.get_property("silent") is True
    # unittest_plugins.cc:2158-2218
    for name in ("frames-in", "frames-out", "frames-flush", "frames-dim"):
        v = a.get_property(name)
        a.set_property(name, v + 1 if name != "frames-dim" else 3)
        assert a.get_property(name) == (v + 1 if name != "frames-dim" else 3)
    a.set_property("concat", False)
    assert a.get_property("concat") is False
    for name, bad in (("frames-in", 0), ("frames-out", 0), ("frames-dim", 8), ("frames-in", "x"),
                      ("no-such", 1)):
        with pytest.raises(NnsbError):
            a.set_property(name, bad)
    a.close()


def test_caps_arithmetic():
    # gsttensor_aggregator.c:1060-1074: 2:200:200:1, frames_out=10 -> dim 3 or dim 2
    a = TensorAggregator(frames_out=10, frames_dim=3)
    out = a.set_caps(caps_for([2, 200, 200, 1], "uint8", "30/1"))
    assert "dimension=(string)2:200:200:10:1:1:1:1" in out and "framerate=(fraction)30/1" in out
    a.close()
    a = TensorAggregator(frames_out=10, frames_dim=2)
    assert "dimension=(string)2:200:2000:1:1:1:1:1" in a.set_caps(caps_for([2, 200, 200, 1], "uint8"))
    a.close()
    # dis-aggregation: 4 frames in, 1 out
    a = TensorAggregator(frames_in=4, frames_out=1, frames_dim=3)
    assert "dimension=(string)3:8:8:1:1:1:1:1" in a.set_caps(caps_for([3, 8, 8, 4], "float32"))
    assert list(a.config(nb._lib.PAD_SINK).info[0].dim)[:4] == [3, 8, 8, 4]
    assert list(a.config(nb._lib.PAD_SRC).info[0].dim)[:4] == [3, 8, 8, 1]
    a.close()
    # "Cannot update dimension in output tensor": dim not divisible by frames-in
    a = TensorAggregator(frames_in=3, frames_dim=3)
    with pytest.raises(NnsbError) as e:
        a.set_caps(caps_for([3, 8, 8, 4]))
    assert "Cannot update dimension" in str(e.value)
    a.close()
    # "Cannot flush frames": in=5 out=10 flush=20
    a = TensorAggregator(frames_in=5, frames_out=10, frames_flush=20, frames_dim=3)
    with pytest.raises(NnsbError) as e:
        a.set_caps(caps_for([3, 8, 8, 5]))
    assert "Cannot flush frames" in str(e.value)
    a.close()
    # not a tensor caps / not fixed
    a = TensorAggregator()
    with pytest.raises(NnsbError):
        a.set_caps("video/x-raw, format=(string)RGB, width=(int)4, height=(int)4")
    a.close()


def test_outputs_for_counts_the_chain_loop():
    a = TensorAggregator(frames_in=3, frames_out=4, frames_flush=2, frames_dim=3)
    a.set_caps(caps_for([2, 2, 2, 3], "uint8"))
    # the documentation's example (gsttensor_aggregator.md:20-41): 3 frames per buffer
    assert a.outputs_for(24) == 0           # 3 frames < 4
    a.close()


# ---------------------------------------------------------------------------------------
# GPU: the device adapter and the concatenation launch
# ---------------------------------------------------------------------------------------
def _run_device(agg, dev, buffers, on_device=False):
    outs = []
    for b in buffers:
        if on_device:
            d = dev.upload(np.ascontiguousarray(b).view(np.uint8).reshape(-1))
            flow, mems, _ = agg.chain(nb.device_memory(d, b.nbytes))
        else:
            flow, mems, _ = agg.chain(b)
        assert flow == 0, nb.lib().nnsb_last_error()
        for m in mems:
            outs.append(dev.download(m.ptr, m.size, np.uint8))
            m.release()
    return outs


@pytest.fixture(scope="module")
def dev():
    d = nb.Device(0)
    yield d
    d.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_device_matches_reference_unit_tests(dev, name):
    fi, fo, ff, fd, concat = (int(v) for v in GOLD[name + "/props"])
    dim = [int(d) for d in GOLD[name + "/dim"]]
    for on_device in (False, True):
        a = TensorAggregator(frames_in=fi, frames_out=fo, frames_flush=ff, frames_dim=fd,
                             concat=bool(concat))
        a.set_caps(caps_for(dim))
        outs = _run_device(a, dev, [GOLD["frames"][0], GOLD["frames"][1]], on_device)
        got = np.concatenate(outs).view(np.int32)
        assert np.array_equal(got, GOLD[name + "/expected"]), name
        a.close()


@pytest.mark.gpu
def test_device_flush_data(dev):
    fi, fo, ff, fd, concat = (int(v) for v in GOLD["flush/props"])
    data = GOLD["flush/data"]
    a = TensorAggregator(frames_in=fi, frames_out=fo, frames_flush=ff, frames_dim=fd)
    a.set_caps(caps_for([10]))
    outs = _run_device(a, dev, [data])
    assert len(outs) == 1 and np.array_equal(outs[0].view(np.int32), data[:6])
    a.reset()
    outs = _run_device(a, dev, [data])
    assert len(outs) == 1 and np.array_equal(outs[0].view(np.int32), data[:6])
    a.close()


AGG_CONFIGS = [
    # dim (incoming), type, frames_in, frames_out, frames_flush, frames_dim, concat, n_buffers
    ([3, 16, 12, 1], "uint8", 1, 4, 0, 3, True, 9),
    ([3, 16, 12, 1], "uint8", 1, 4, 2, 3, True, 9),         # sliding window
    ([3, 16, 12, 3], "float32", 3, 4, 2, 3, True, 7),       # the documentation's 3/4/2 example
    ([3, 16, 12, 1], "float32", 1, 5, 0, 2, True, 11),      # concat along height (no reorder: dim 3 is 1)
    ([3, 16, 12, 2], "float32", 1, 3, 0, 2, True, 7),       # concat along height with 2 outer frames
    ([3, 16, 12, 2], "int16", 1, 3, 1, 1, True, 8),         # concat along width
    ([4, 16, 12, 2], "uint8", 1, 2, 0, 0, True, 6),         # concat along channels
    ([3, 16, 12, 2], "float64", 1, 3, 0, 1, False, 6),      # concat=false: frames stay back to back
    ([3, 8, 8, 8], "uint8", 8, 2, 0, 3, True, 3),           # dis-aggregation 8 -> 4 x 2
    ([3, 8, 8, 4], "float32", 4, 1, 0, 3, True, 3),         # dis-aggregation 4 -> 4 x 1
    ([3, 8, 8, 4], "float32", 4, 4, 0, 3, True, 3),         # frames-in == frames-out
    ([3, 8, 6, 4], "float32", 2, 2, 0, 1, True, 3),         # == with concat along an inner axis
    ([5, 7, 3, 4], "uint8", 4, 3, 4, 3, True, 9),           # flush more than one output
]


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", AGG_CONFIGS, ids=lambda c: "%s-%s-in%d-out%d-fl%d-d%d-%s" % (
    "x".join(map(str, c[0])), c[1], c[2], c[3], c[4], c[5], "cat" if c[6] else "nocat"))
@pytest.mark.parametrize("on_device", [False, True], ids=["host", "device"])
def test_device_matches_oracle_model(dev, port, cfg, on_device):
    dim, type_name, fi, fo, ff, fd, concat, n_buf = cfg
    rng = np.random.default_rng(sum(dim) + fo)
    n = int(np.prod(dim))
    bufs = [rng.integers(0, 200, n).astype(type_name) for _ in range(n_buf)]
    model = O.AggregatorModel(port, dim, type_name, fi, fo, ff, fd, concat)
    want = []
    for b in bufs:
        want += model.chain(b)
    a = TensorAggregator(frames_in=fi, frames_out=fo, frames_flush=ff, frames_dim=fd, concat=concat)
    out_caps = a.set_caps(caps_for(dim, type_name))
    assert "dimension=(string)%s" % ":".join(str(d) for d in model.out_dim) in out_caps
    got = _run_device(a, dev, bufs, on_device)
    a.close()
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)


@pytest.mark.gpu
def test_device_timestamps(dev):
    """:873-912: pts of an output = pts of the buffer its first frame came from, plus the frames
    into that buffer at the caps framerate; duration scaled by frames_out / frames_in."""
    a = TensorAggregator(frames_in=2, frames_out=3, frames_dim=3)
    a.set_caps(caps_for([4, 4, 4, 2], "uint8", "10/1"))
    sec = 1000000000
    buf = np.zeros(4 * 4 * 4 * 2, dtype=np.uint8)
    times = []
    for i in range(6):
        flow, mems, ts = a.chain(buf, pts=i * sec // 5, dts=i * sec // 5, duration=sec // 5)
        assert flow == 0
        for m in mems:
            m.release()
        times += ts
    a.close()
    # outputs start at frames 0, 3, 6, 9 -> 0.0 s, 0.3 s, 0.6 s, 0.9 s at 10 frames/s
    assert [t[0] for t in times] == [0, 3 * sec // 10, 6 * sec // 10, 9 * sec // 10]
    assert all(t[2] == (sec // 5) * 3 // 2 for t in times)


@pytest.mark.gpu
def test_device_large_concat_uses_the_permutation_kernels(dev, port):
    """1080p RGB frames concatenated along the width: one launch per output, bit-exact."""
    dim = [3, 1920, 1080, 1]
    rng = np.random.default_rng(5)
    bufs = [rng.integers(0, 256, int(np.prod(dim)), dtype=np.uint8) for _ in range(4)]
    model = O.AggregatorModel(port, dim, "uint8", 1, 4, 0, 1, True)
    want = []
    for b in bufs:
        want += model.chain(b)
    a = TensorAggregator(frames_out=4, frames_dim=1)
    a.set_caps(caps_for(dim, "uint8"))
    got = _run_device(a, dev, bufs)
    assert a.kernel_launches == 1
    a.close()
    assert len(got) == 1 and np.array_equal(got[0], want[0])


@pytest.mark.gpu
def test_zero_copy_append_and_concat_through_the_output_provider(dev, port):
    """tensor_transform ! tensor_aggregator with the aggregator handing out the memory the
    transform writes into (its ring): appending along the outermost axis moves no bytes at all --
    outputs are shares of the ring -- and the width concatenation moves each byte exactly once
    (the permutation launch).  Sliding window (frames-flush 2 of 4) so that outputs overlap."""
    import ctypes as C

    from nnstreamer_b200 import DeviceMemory, TensorTransform
    dim = [3, 320, 240, 1]
    n = int(np.prod(dim))
    rng = np.random.default_rng(17)
    frames = [rng.integers(0, 256, n, dtype=np.uint8) for _ in range(10)]
    conv = [port.typecast(f, "float32") for f in frames]
    for frames_dim, concat_bytes in ((3, 0), (1, 1)):
        t = TensorTransform(mode="typecast", option="float32")
        out_caps = t.negotiate(caps_for(dim, "uint8", "30/1"))
        agg = TensorAggregator(frames_out=4, frames_flush=2, frames_dim=frames_dim)
        agg.set_caps(out_caps)
        t.link_to_aggregator(agg)
        model = O.AggregatorModel(port, dim, "float32", 1, 4, 2, frames_dim, True)
        got, want = [], []
        for f, c in zip(frames, conv):
            m = t.alloc(n, on_device=True)
            assert t._lib.nnsb_transform_upload(t.handle, C.byref(m), f.ctypes.data_as(C.c_void_p)) == 0
            src = DeviceMemory(m)
            flow, (y,) = t.transform([src])
            assert flow == 0
            src.release()
            flow, mems, _ = agg.chain(y)
            assert flow == 0, nb.lib().nnsb_last_error()
            y.release()
            got += mems
            want += model.chain(c)
        assert len(got) == len(want) == 4
        copied_in, copied_out = agg.bytes_copied
        assert copied_in == 0, "the frames were written in place"
        assert copied_out == concat_bytes * sum(w.nbytes for w in want)
        for g, w in zip(got, want):
            assert dev.download(g.ptr, g.size, np.uint8).tobytes() == w.tobytes()
            g.release()
        agg.close()
        t.close()


@pytest.mark.gpu
def test_ring_is_replaced_while_outputs_still_reference_it(dev, port):
    """Outputs are shares of the adapter's block: keep ALL of them alive over many buffers, so
    the adapter has to move on to fresh blocks; every output must still hold its own bytes."""
    dim = [4, 16, 16, 1]
    n = int(np.prod(dim))
    rng = np.random.default_rng(3)
    agg = TensorAggregator(frames_out=2, frames_flush=1, frames_dim=3)
    agg.set_caps(caps_for(dim, "int32"))
    model = O.AggregatorModel(port, dim, "int32", 1, 2, 1, 3, True)
    got, want = [], []
    for i in range(40):
        b = rng.integers(-1000, 1000, n, dtype=np.int32)
        flow, mems, _ = agg.chain(b)
        assert flow == 0
        got += mems
        want += model.chain(b)
    assert len(got) == len(want) == 39
    for g, w in zip(got, want):
        assert dev.download(g.ptr, g.size, np.uint8).tobytes() == w.tobytes()
    agg.close()                                      # the element goes first, the shares stay valid
    for g, w in zip(got, want):
        assert dev.download(g.ptr, g.size, np.uint8).tobytes() == w.tobytes()
        g.release()
