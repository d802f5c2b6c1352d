"""tensor_merge (mode=linear) and tensor_split data paths on the device.

Reference: gst/nnstreamer/elements/gsttensor_merge.c (generate_mem :463-600, merged config
:380-427) and gsttensor_split.c (get_splited :413-450).  The reference's own goldens for merge
(tests/nnstreamer_merge/generateTest.py:36-103) are concatenations of float32 tensors along the
channel ([3,2,4] x 100 x 50), width ([100,200,300]) and batch ([1,2,3]) axes -- reproduced here
with the same shapes; its split tests (tests/nnstreamer_split/runTest.sh) cut 3:100:100 RGB frames
into 1:100:100 / 2:100:100 byte ranges."""
import numpy as np
import pytest

import nnstreamer_b200 as nb
from nnstreamer_b200 import NnsbError, TensorMerge, TensorSplit
from oracle import oracle as O


def caps_for(dim, type_name="float32", rate="0/1"):
    return "other/tensor, dimension=(string)%s, type=(string)%s, framerate=(fraction)%s" % (
        ":".join(str(d) for d in dim), type_name, rate)


@pytest.fixture(scope="module")
def port():
    return O.port()


@pytest.fixture(scope="module")
def dev():
    d = nb.Device(0)
    yield d
    d.close()


# the reference generator's three cases + a height merge: (axis, list of dims)
REF_CASES = [
    (0, [[3, 100, 50, 1], [2, 100, 50, 1], [4, 100, 50, 1]]),        # channel.golden
    (1, [[3, 100, 50, 1], [3, 200, 50, 1], [3, 300, 50, 1]]),        # width.golden
    (3, [[3, 100, 50, 1], [3, 100, 50, 2], [3, 100, 50, 3]]),        # batch.golden
    (2, [[3, 100, 20, 2], [3, 100, 30, 2]]),
]


def np_merge(tensors, dims, axis):
    """numpy statement of the goldens: concatenate along `axis` (dims are innermost-first)."""
    shaped = [t.reshape(tuple(reversed(d))) for t, d in zip(tensors, dims)]
    return np.concatenate(shaped, axis=3 - axis).reshape(-1)


# ---------------------------------------------------------------------------------------
# CPU
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("case", REF_CASES, ids=lambda c: "axis%d" % c[0])
def test_oracle_merge_matches_the_reference_goldens(port, case):
    axis, dims = case
    rng = np.random.default_rng(axis)
    tensors = [rng.uniform(0.0, 10.0, int(np.prod(d))).astype(np.float32) for d in dims]
    got = port.merge_linear(tensors, dims, axis)
    assert np.array_equal(got, np_merge(tensors, dims, axis))


def test_merge_properties_and_caps():
    m = TensorMerge()
    assert m.get_property("option") is None and m.get_property("sync-mode") == "nosync"
    with pytest.raises(NnsbError):
        m.set_caps([caps_for([3, 4, 4, 1])])                # mode / option not set
    m.set_property("mode", "LINEAR")                        # g_ascii_strcasecmp
    with pytest.raises(NnsbError):
        m.set_property("mode", "cubic")
    with pytest.raises(NnsbError):
        m.set_property("option", "4")
    m.set_property("option", "2")
    out = m.set_caps([caps_for([3, 100, 100, 1], "uint8", "30/1")] * 3)
    assert "dimension=(string)3:100:300:1:1:1:1:1" in out and "framerate=(fraction)30/1" in out
    with pytest.raises(NnsbError) as e:                     # type mismatch: negotiation error
        m.set_caps([caps_for([3, 4, 4, 1], "uint8"), caps_for([3, 4, 4, 1], "int8")])
    assert "different type" in str(e.value)
    with pytest.raises(NnsbError) as e:                     # other dims must agree
        m.set_caps([caps_for([3, 4, 4, 1]), caps_for([3, 5, 4, 1])])
    assert "differs in dimension 1" in str(e.value)
    m.close()


def test_split_properties_and_caps():
    s = TensorSplit(tensorseg="2:100:100,1:100:100")
    assert s.get_property("tensorseg") == "2:100:100:1:1:1:1:1,1:100:100:1:1:1:1:1"
    assert s.num_outputs == 2
    s.set_caps(caps_for([3, 100, 100, 1], "uint8", "30/1"))
    assert "dimension=(string)2:100:100:1:1:1:1:1" in s.out_caps(0)
    assert "type=(string)uint8" in s.out_caps(1) and "framerate=(fraction)30/1" in s.out_caps(1)
    s.set_property("tensorpick", "1")
    assert s.get_property("tensorpick") == "1" and s.num_outputs == 1
    s.close()
    # separators of g_strsplit_set (",.;/")
    s = TensorSplit(tensorseg="1:4;2:4/3:4")
    assert s.num_outputs == 3
    s.close()


# ---------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("case", REF_CASES, ids=lambda c: "axis%d" % c[0])
@pytest.mark.parametrize("on_device", [False, True], ids=["host", "device"])
def test_device_merge_matches_oracle_on_the_reference_shapes(dev, port, case, on_device):
    axis, dims = case
    rng = np.random.default_rng(10 + axis)
    tensors = [rng.uniform(0.0, 10.0, int(np.prod(d))).astype(np.float32) for d in dims]
    want = port.merge_linear(tensors, dims, axis)
    m = TensorMerge(mode="linear", option=str(axis))
    m.set_caps([caps_for(d) for d in dims])
    ins = tensors
    if on_device:
        ins = [nb.device_memory(dev.upload(t), t.nbytes) for t in tensors]
    flow, out = m.chain(ins)
    assert flow == 0, nb.lib().nnsb_last_error()
    got = dev.download(out.ptr, out.size // 4, np.float32)
    out.release()
    m.close()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


MERGE_EXTRA = [
    # type, axis, dims: odd byte counts, many inputs, blocks smaller than a 16-byte chunk
    ("uint8", 0, [[3, 7, 5, 1], [1, 7, 5, 1], [2, 7, 5, 1]]),
    ("uint8", 0, [[1, 33, 9, 2]] * 16),
    ("uint8", 1, [[3, 5, 11, 1], [3, 9, 11, 1]]),
    ("int16", 1, [[3, 17, 6, 2], [3, 1, 6, 2], [3, 30, 6, 2]]),
    ("float64", 2, [[2, 6, 3, 4], [2, 6, 5, 4]]),
    ("uint8", 2, [[3, 642, 100, 1], [3, 642, 140, 1]]),
    ("uint8", 0, [[3, 1920, 1080, 1], [1, 1920, 1080, 1]]),          # RGB + alpha plane -> RGBA
]


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", MERGE_EXTRA, ids=lambda c: "%s-axis%d-n%d-%s" % (
    c[0], c[1], len(c[2]), "x".join(map(str, c[2][0]))))
def test_device_merge_edge_shapes(dev, port, cfg):
    type_name, axis, dims = cfg
    rng = np.random.default_rng(len(dims) + axis)
    tensors = [rng.integers(0, 100, int(np.prod(d))).astype(type_name) for d in dims]
    want = port.merge_linear(tensors, dims, axis)
    m = TensorMerge(mode="linear", option=str(axis))
    m.set_caps([caps_for(d, type_name) for d in dims])
    flow, out = m.chain(tensors)
    assert flow == 0, nb.lib().nnsb_last_error()
    got = dev.download(out.ptr, out.size, np.uint8)
    out.release()
    m.close()
    assert np.array_equal(got, want.view(np.uint8))


@pytest.mark.gpu
def test_device_merge_wrong_number_of_memories(dev):
    m = TensorMerge(mode="linear", option="0")
    m.set_caps([caps_for([3, 4, 4, 1]), caps_for([1, 4, 4, 1])])
    flow, out = m.chain([np.zeros(48, np.float32)])
    assert flow != 0 and out is None
    m.close()


@pytest.mark.gpu
@pytest.mark.parametrize("on_device", [False, True], ids=["host", "device"])
def test_device_split_reference_shapes(dev, on_device):
    """tests/nnstreamer_split/runTest.sh: 3:100:100 frames cut into tensorseg byte ranges."""
    x = np.random.default_rng(2).integers(0, 256, 3 * 100 * 100, dtype=np.uint8)
    src = nb.device_memory(dev.upload(x), x.nbytes) if on_device else x
    for seg, pick, want_ranges in (
            ("1:100:100,2:100:100", None, [(0, 10000), (10000, 30000)]),
            ("1:100:100,1:100:100,1:100:100", None, [(0, 10000), (10000, 20000), (20000, 30000)]),
            ("1:100:100,1:100:100,1:100:100", "0,2", [(0, 10000), (20000, 30000)]),
            ("2:100:100,1:100:100", "1", [(20000, 30000)])):
        s = TensorSplit(tensorseg=seg)
        if pick:
            s.set_property("tensorpick", pick)
        s.set_caps(caps_for([3, 100, 100, 1], "uint8"))
        flow, mems = s.chain(src)
        assert flow == 0 and len(mems) == len(want_ranges)
        for m, (a, b) in zip(mems, want_ranges):
            assert m.size == b - a
            assert np.array_equal(dev.download(m.ptr, m.size, np.uint8), x[a:b])
            m.release()
        s.close()


@pytest.mark.gpu
def test_device_split_errors(dev):
    s = TensorSplit()
    s.set_caps(caps_for([3, 4, 4, 1], "uint8"))
    flow, mems = s.chain(np.zeros(48, np.uint8))            # "No rule to split incoming buffers."
    assert flow != 0 and not mems
    s.set_property("tensorseg", "3:4:4,3:4:4")              # second segment runs past the buffer
    flow, mems = s.chain(np.zeros(48, np.uint8))
    assert flow != 0 and not mems
    s.close()


@pytest.mark.gpu
def test_split_then_merge_round_trip_on_device(dev):
    x = np.random.default_rng(4).standard_normal(4 * 64 * 48).astype(np.float32)
    s = TensorSplit(tensorseg="4:64:16,4:64:32")
    s.set_caps(caps_for([4, 64, 48, 1]))
    flow, parts = s.chain(x)
    assert flow == 0 and len(parts) == 2
    m = TensorMerge(mode="linear", option="2")
    m.set_caps([s.out_caps(0), s.out_caps(1)])
    flow, out = m.chain(parts)
    assert flow == 0
    assert np.array_equal(dev.download(out.ptr, x.size, np.float32).view(np.uint32), x.view(np.uint32))
    for p in parts:
        p.release()
    out.release()
    s.close()
    m.close()
