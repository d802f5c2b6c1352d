"""tensor_mux / tensor_demux mirrors: memories are handed on as references
(gst_buffer_append_memory of refs, gsttensor_mux.c:467,482; gst_buffer_get_memory,
gsttensor_demux.c:516-522) and a blob of several static tensors is cut into shares
(gst_tensor_buffer_from_config, nnstreamer_plugin_api_impl.c:452-556).

CPU part: properties, caps arithmetic and the reference passing on host memories (no GPU
needed: nothing is computed).  GPU part: the reference's SSAT pattern
`tensor_mux ! tensor_transform apply=0,2` (tests/transform_typecast/runTest.sh:105-119) with
every tensor in HBM, and life-time of shares over pooled device blocks."""
import ctypes as C

import numpy as np
import pytest

from nnstreamer_b200 import TensorDemux, TensorMux
from nnstreamer_b200 import _lib as L


def caps_for(dim, type_name="float32", rate="30/1"):
    return "other/tensor, dimension=(string)%s, type=(string)%s, framerate=(fraction)%s" % (
        ":".join(str(d) for d in dim), type_name, rate)


def tensors_caps(dims, types, rate="30/1"):
    return ('other/tensors, format=(string)static, num_tensors=(int)%d, dimensions=(string)"%s", '
            'types=(string)"%s", framerate=(fraction)%s' % (
                len(dims), ",".join(":".join(str(d) for d in dim) for dim in dims), ",".join(types), rate))


def norm(caps):
    """caps string without blanks, quotes and escapes: `types=(string)"a\\,b"` -> types=(string)a,b"""
    return caps.replace(" ", "").replace("\\", "").replace('"', "")


# ---------------------------------------------------------------------------------------
# CPU
# ---------------------------------------------------------------------------------------
def test_mux_properties_and_merged_caps():
    m = TensorMux()
    assert m.get_property("silent") == "true"
    assert m.get_property("sync-mode") == "nosync"
    assert m.get_property("sync-option") is None
    m.set_property("sync-mode", "slowest")
    m.set_property("sync-option", "0:33333333")
    assert m.get_property("sync-mode") == "slowest"
    assert m.get_property("sync-option") == "0:33333333"
    with pytest.raises(L.NnsbError):
        m.set_property("sync-mode", "whenever")
    out = m.set_caps([caps_for([3, 4, 4, 1], "uint8", "30/1"), caps_for([2, 2, 1, 1], "float32", "10/1"),
                      tensors_caps([[1, 1, 1, 1], [5, 1, 1, 1]], ["int16", "int8"], "30/1")])
    assert out.startswith("other/tensors")
    assert "num_tensors=(int)4" in out
    assert "types=(string)uint8,float32,int16,int8" in norm(out)
    # smallest numerator, smallest denominator (nnstreamer_plugin_api_impl.c:345-348)
    assert "framerate=(fraction)10/1" in out
    m.close()


def test_mux_chain_passes_host_memories_by_reference():
    m = TensorMux()
    a = np.arange(48, dtype=np.uint8)
    b = np.arange(4, dtype=np.float32)
    m.set_caps([caps_for([3, 4, 4, 1], "uint8"), caps_for([2, 2, 1, 1], "float32")])
    flow, outs = m.chain([a, b])
    assert flow == L.FLOW_OK and len(outs) == 2
    assert outs[0].data == a.ctypes.data and outs[0].size == 48 and not outs[0].owned
    assert outs[1].data == b.ctypes.data and outs[1].size == 16
    flow, outs = m.chain([a])                        # one memory short
    assert flow == L.FLOW_ERROR and not outs
    m.close()


def test_demux_tensorpick_grammar_and_pad_caps():
    d = TensorDemux()
    d.set_caps(tensors_caps([[3, 4, 4, 1], [2, 2, 1, 1], [5, 1, 1, 1]], ["uint8", "float32", "int16"]))
    assert d.num_outputs == 3                        # no tensorpick: one pad per tensor
    assert "dimension=(string)2:2:1:1" in d.out_caps(1) and d.out_caps(1).startswith("other/tensor,")
    d.set_property("tensorpick", "0,1:2,2+0")
    assert d.get_property("tensorpick") == "0,1:2,2+0"
    assert d.num_outputs == 3
    assert d.out_caps(0).startswith("other/tensor,") and "type=(string)uint8" in d.out_caps(0)
    c1 = d.out_caps(1)
    assert c1.startswith("other/tensors") and "num_tensors=(int)2" in c1
    assert "types=(string)float32,int16" in norm(c1)
    c2 = norm(d.out_caps(2))
    assert "types=(string)int16,uint8" in c2
    d.set_property("tensorpick", "0,7")
    with pytest.raises(L.NnsbError):
        d.out_caps(1)                                # index beyond the configured tensors
    d.close()


def test_demux_chain_on_host_memories_and_a_blob():
    d = TensorDemux(tensorpick="2,0:1")
    d.set_caps(tensors_caps([[3, 4, 4, 1], [2, 2, 1, 1], [5, 1, 1, 1]], ["uint8", "float32", "int16"]))
    a = np.arange(48, dtype=np.uint8)
    b = np.arange(4, dtype=np.float32)
    c = np.arange(5, dtype=np.int16)
    flow, pads = d.chain([a, b, c])
    assert flow == L.FLOW_OK and [len(p) for p in pads] == [1, 2]
    assert pads[0][0].data == c.ctypes.data and pads[0][0].size == 10
    assert pads[1][0].data == a.ctypes.data and pads[1][1].data == b.ctypes.data
    # one blob holding the three tensors back to back is cut at the configured sizes
    blob = np.concatenate([a.view(np.uint8), b.view(np.uint8), c.view(np.uint8)])
    flow, pads = d.chain([blob])
    assert flow == L.FLOW_OK
    assert pads[0][0].data == blob.ctypes.data + 48 + 16 and pads[0][0].size == 10
    assert pads[1][0].data == blob.ctypes.data and pads[1][0].size == 48
    assert pads[1][1].data == blob.ctypes.data + 48 and pads[1][1].size == 16
    flow, pads = d.chain([a, b])                     # a tensor short
    assert flow == L.FLOW_ERROR and not pads
    d.close()


# ---------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_mux_then_transform_apply_keeps_skipped_tensors_as_references(dev, port):
    """tests/transform_typecast/runTest.sh:105-119: three streams -> tensor_mux ->
    tensor_transform mode=typecast option=float32 apply=0,2.  Tensors 0 and 2 are converted,
    tensor 1 comes out as the very same device block (gst_memory_ref)."""
    from nnstreamer_b200 import TensorTransform
    dims = [[3, 16, 16, 1], [3, 16, 16, 1], [3, 16, 16, 1]]
    rng = np.random.default_rng(3)
    xs = [rng.integers(0, 256, 768, dtype=np.uint8) for _ in range(3)]
    holder = TensorTransform(mode="typecast", option="uint8")      # only to own device memories
    mems = []
    for x in xs:
        m = holder.alloc(x.nbytes, on_device=True)
        assert holder._lib.nnsb_transform_upload(holder.handle, C.byref(m), x.ctypes.data_as(C.c_void_p)) == 0
        from nnstreamer_b200 import DeviceMemory
        mems.append(DeviceMemory(m))
    mux = TensorMux()
    caps = mux.set_caps([caps_for(d, "uint8") for d in dims])
    flow, muxed = mux.chain(mems)
    assert flow == L.FLOW_OK and len(muxed) == 3
    for m in mems:                                   # upstream drops its references: the mux's hold
        m.release()
    t = TensorTransform(mode="typecast", option="float32", apply="0,2")
    out_caps = t.negotiate(caps)
    assert "types=(string)float32,uint8,float32" in norm(out_caps)
    flow, outs = t.transform(muxed)
    assert flow == L.FLOW_OK and len(outs) == 3
    assert outs[1].ptr_async == muxed[1].ptr_async and outs[1].mem.block == muxed[1].mem.block
    for m in muxed:
        m.release()
    for i in (0, 2):
        got = dev.download(outs[i].ptr, 768, np.float32)
        assert got.tobytes() == port.typecast(xs[i], "float32").tobytes()
    assert dev.download(outs[1].ptr, 768, np.uint8).tobytes() == xs[1].tobytes()
    for o in outs:
        o.release()
    for e in (t, mux, holder):
        e.close()


@pytest.mark.gpu
def test_demux_shares_outlive_the_parent_reference(dev):
    """A device blob of three tensors -> demux -> shares.  The producer's reference and the
    element go away first; the shares keep the pooled block alive and intact."""
    from nnstreamer_b200 import DeviceMemory, TensorTransform
    a = np.arange(48, dtype=np.uint8)
    b = np.arange(4, dtype=np.float32) + 0.5
    c = np.arange(5, dtype=np.int16) - 2
    blob = np.concatenate([a, b.view(np.uint8), c.view(np.uint8)])
    holder = TensorTransform(mode="typecast", option="uint8")
    m = holder.alloc(blob.nbytes, on_device=True)
    assert holder._lib.nnsb_transform_upload(holder.handle, C.byref(m), blob.ctypes.data_as(C.c_void_p)) == 0
    parent = DeviceMemory(m)
    d = TensorDemux(tensorpick="1,2+0")
    d.set_caps(tensors_caps([[3, 4, 4, 1], [2, 2, 1, 1], [5, 1, 1, 1]], ["uint8", "float32", "int16"]))
    flow, pads = d.chain([parent])
    assert flow == L.FLOW_OK and [len(p) for p in pads] == [1, 2]
    base = parent.ptr_async
    parent.release()
    holder.close()                                   # the allocating element is gone too
    d.close()
    # churn the allocator: had the block been recycled these would overwrite it
    junk = [dev.alloc(256) for _ in range(8)]
    for p in junk:
        L.check(L.lib().nnsb_dev_memset_async(dev.stream, C.c_void_p(p), 0xEE, 256))
    dev.sync()
    assert pads[0][0].ptr_async == base + 48 and pads[1][0].ptr_async == base + 64
    assert dev.download(pads[0][0].ptr, 4, np.float32).tobytes() == b.tobytes()
    assert dev.download(pads[1][0].ptr, 5, np.int16).tobytes() == c.tobytes()
    assert dev.download(pads[1][1].ptr, 48, np.uint8).tobytes() == a.tobytes()
    for p in pads:
        for mem in p:
            mem.release()
    for p in junk:
        dev.free(p)


def test_branch_rejects_bad_arguments():
    """nnsb_branch_new / _push argument checks (no GPU work)."""
    import ctypes as C
    from nnstreamer_b200 import _lib as L
    lib = L.lib()
    h = C.c_void_p()
    assert lib.nnsb_branch_new(None, None, 0, C.byref(h)) < 0           # no elements at all
    arr = (C.c_void_p * 1)(None)
    assert lib.nnsb_branch_new(None, arr, 1, C.byref(h)) < 0            # a NULL element
    assert lib.nnsb_branch_free(None) == 0
    n = C.c_uint32(7)
    outs = (L.Memory * L.SIZE_LIMIT)()
    ins = (L.Memory * 1)()
    assert lib.nnsb_branch_push(None, ins, 1, outs, C.byref(n)) != 0
