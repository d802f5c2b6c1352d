"""CPU: the element mirrors' host logic -- properties, option revert, caps negotiation,
converter caps parsing.  Follows tests/nnstreamer_plugins/unittest_plugins.cc:208-950 and the
caps-related SSAT cases (transform_typecast/runTest.sh:140-153)."""
import pytest

from nnstreamer_b200 import TensorConverter, TensorTransform
from nnstreamer_b200 import _lib as L

from test_host_core import NEGATIVE


def test_properties01():
    """unittest_plugins.cc:208-268 (tensor_transform mode=1 option=uint32)."""
    t = TensorTransform(mode=1, option="uint32")
    assert t.get_property("silent") is True             # instance default is TRUE (:291)
    t.set_property("silent", False)
    assert t.get_property("silent") is False
    assert t.get_property("acceleration") is False      # DEFAULT_ACCELERATION without ORC
    t.set_property("acceleration", True)
    assert t.get_property("acceleration") is True
    assert t.get_property("mode") == "typecast"
    assert t.get_property("option") == "uint32"
    assert t.get_property("transpose-rank-limit") == 4
    with pytest.raises(L.NnsbError):
        t.set_property("transpose-rank-limit", 5)
    with pytest.raises(L.NnsbError):
        t.set_property("no-such-property", 1)
    t.close()


def test_fresh_element_defaults():
    t = TensorTransform()
    assert t.get_property("mode") == "unknown"
    assert t.get_property("option") is None
    assert t.get_property("apply") == ""
    t.close()


@pytest.mark.parametrize("mode,option", NEGATIVE[:28], ids=["%s|%s" % n for n in NEGATIVE[:28]])
def test_invalid_option_reads_back_null(mode, option):
    """unittest_plugins.cc:290-950: set mode + a bad option, then `option` must be NULL."""
    t = TensorTransform()
    t.set_property("mode", mode)
    assert t.set_property("option", option) is False
    assert t.get_property("option") is None
    t.close()


def test_mode_given_as_garbage_is_rejected_and_option_stays_null():
    """arithmeticProperties4_n (:446-463): g_object_set (mode, "typecast:int8")."""
    t = TensorTransform()
    with pytest.raises(L.NnsbError):
        t.set_property("mode", "typecast:int8")
    assert t.get_property("option") is None
    t.close()


def test_invalid_option_keeps_the_previous_one():
    """gsttensor_transform.c:985-990: revert to the backup option and re-parse it."""
    t = TensorTransform(mode="arithmetic", option="add:.5")
    assert t.set_property("option", "max") is False
    assert t.get_property("option") == "add:.5"
    assert t.set_property("option", "mul:20") is True
    assert t.get_property("option") == "mul:20"
    t.close()


def test_option_before_mode_is_parsed_when_mode_arrives():
    """set_option_data runs on both setters (:973-976): order of mode/option does not matter."""
    t = TensorTransform()
    assert t.set_property("option", "0:2") is True        # mode unknown: skipped, kept
    t.set_property("mode", "dimchg")
    out = t.transform_caps("sink", "other/tensor, dimension=(string)3:4:5:1, type=(string)uint8, framerate=(fraction)0/1")
    assert "dimension=(string)4:5:3:1:1:1:1:1" in out
    t.close()


def test_apply_property_appends():
    t = TensorTransform()
    t.set_property("apply", "0,2")
    assert t.get_property("apply") == "0,2"
    t.set_property("apply", "5")
    assert t.get_property("apply") == "0,2,5"
    t.close()


IN_CAPS = ("other/tensors, format=(string)static, num_tensors=(int)1, dimensions=(string)3:224:224:1, "
           "types=(string)uint8, framerate=(fraction)30/1")


def test_transform_caps_sink_and_src_directions():
    t = TensorTransform(mode="arithmetic", option="typecast:float32,add:-127.5,div:127.5")
    out = t.transform_caps("sink", IN_CAPS)
    assert out == ("other/tensors, format=(string)static, num_tensors=(int)1, framerate=(fraction)30/1, "
                   "dimensions=(string)3:224:224:1:1:1:1:1, types=(string)float32")
    # SRC -> SINK: the incoming type is unknowable, `types` is removed (:2023-2030)
    back = t.transform_caps("src", IN_CAPS)
    assert "types" not in back and "dimensions=(string)3:224:224:1:1:1:1:1" in back
    # without a typecast the type passes through in both directions
    t2 = TensorTransform(mode="arithmetic", option="add:1")
    assert "types=(string)uint8" in t2.transform_caps("src", IN_CAPS)
    t.close()
    t2.close()


def test_transform_caps_other_tensor_and_dimension_modes():
    caps = "other/tensor, dimension=(string)3:4:5:2, type=(string)float32, framerate=(fraction)0/1"
    t = TensorTransform(mode="transpose", option="1:2:0:3")
    assert t.transform_caps("sink", caps) == \
        "other/tensor, framerate=(fraction)0/1, dimension=(string)4:5:3:2:1:1:1:1, type=(string)float32"
    # src direction inverts the permutation
    assert "dimension=(string)5:3:4:2:1:1:1:1" in t.transform_caps("src", caps)
    t = TensorTransform(mode="dimchg", option="0:2")
    assert "dimension=(string)4:5:3:2:1:1:1:1" in t.transform_caps("sink", caps)
    assert "dimension=(string)5:3:4:2:1:1:1:1" in t.transform_caps("src", caps)
    t = TensorTransform(mode="stand", option="default")
    assert "type=(string)float32" in t.transform_caps("sink", caps)
    assert "type=" not in t.transform_caps("src", caps)
    t = TensorTransform(mode="clamp", option="0:1")
    assert "type=(string)float32" in t.transform_caps("src", caps)


def test_transform_caps_filter_adopts_peer_dimension_string():
    """gst_tensor_caps_update_dimension (nnstreamer_plugin_api_impl.c:999-1035): GstCaps compares
    strings, so a numerically equal peer dimension string is adopted before intersecting."""
    t = TensorTransform(mode="typecast", option="float32")
    caps = "other/tensor, dimension=(string)3:4:5:1, type=(string)uint8, framerate=(fraction)0/1"
    peer = "other/tensor, dimension=(string)3:4:5, type=(string)float32, framerate=(fraction)0/1"
    out = t.transform_caps("sink", caps, peer)
    assert out == "other/tensor, framerate=(fraction)0/1, dimension=(string)3:4:5, type=(string)float32"
    # a peer asking for another type cannot intersect
    bad = "other/tensor, dimension=(string)3:4:5, type=(string)uint8, framerate=(fraction)0/1"
    assert t.transform_caps("sink", caps, bad) == "EMPTY"


def test_flexible_caps_stay_flexible():
    t = TensorTransform(mode="typecast", option="float32")
    out = t.transform_caps("sink", "other/tensors, format=(string)flexible, framerate=(fraction)30/1")
    assert out == "other/tensors, format=(string)flexible, framerate=(fraction)30/1"


def test_multi_tensor_caps_with_apply():
    """transform_typecast/runTest.sh:105-119: three tensors, apply=0,2."""
    t = TensorTransform(mode="typecast", option="float32", apply="0,2")
    caps = ('other/tensors, format=(string)static, num_tensors=(int)3, '
            'dimensions=(string)"3:4:4:1\\,2:2:2:1\\,5:1:1:1", types=(string)"uint8\\,uint8\\,uint8", '
            'framerate=(fraction)0/1')
    out = t.negotiate(caps)
    assert 'types=(string)"float32\\,uint8\\,float32"' in out
    cfg = t.config("src")
    assert cfg.num_tensors == 3
    assert [cfg.info[i].type for i in range(3)] == [L.TYPE_ID["float32"], L.TYPE_ID["uint8"], L.TYPE_ID["float32"]]


def test_set_caps_rejects_mismatch():
    """transform_typecast/runTest.sh:140-148 (caps mismatch must fail negotiation)."""
    t = TensorTransform(mode="typecast", option="float32")
    inc = "other/tensor, dimension=(string)3:4:5:1, type=(string)uint8, framerate=(fraction)0/1"
    good = "other/tensor, dimension=(string)3:4:5:1, type=(string)float32, framerate=(fraction)0/1"
    assert t.set_caps(inc, good)
    assert not t.set_caps(inc, good.replace("float32", "uint8"))
    assert not t.set_caps(inc, good.replace("3:4:5:1", "3:4:5:2"))
    assert not t.set_caps(inc, good.replace("0/1", "30/1"))
    assert not t.set_caps("video/x-raw, format=RGB", good)
    # rank-3 declaration of the same dimension is equal (:151-153)
    assert t.set_caps(inc, good.replace("3:4:5:1", "3:4:5"))
    # output may be flexible while the input is static (:2142-2147)
    assert t.set_caps(inc.replace("other/tensor,", "other/tensors, format=static, num_tensors=1,").replace(
        "dimension=", "dimensions=").replace("type=", "types="),
        "other/tensors, format=(string)flexible, framerate=(fraction)0/1")


def test_transform_size_is_always_zero():
    import ctypes as C
    t = TensorTransform(mode="typecast", option="float32")
    sz = C.c_size_t(123)
    assert L.lib().nnsb_transform_transform_size(t.handle, L.PAD_SINK, b"x", 100, b"y", C.byref(sz)) == 0
    assert sz.value == 0


def test_transform_before_configuration_is_a_flow_error():
    import numpy as np
    t = TensorTransform()
    flow, outs = t.transform([np.zeros(4, dtype=np.uint8)])
    assert flow == L.FLOW_ERROR and outs == []


# ---- tensor_converter (video) ------------------------------------------------------------
@pytest.mark.parametrize("fmt,c", [("GRAY8", 1), ("RGB", 3), ("BGR", 3), ("RGBx", 4), ("BGRx", 4),
                                   ("xRGB", 4), ("xBGR", 4), ("RGBA", 4), ("BGRA", 4), ("ARGB", 4),
                                   ("ABGR", 4)])
def test_converter_video_caps(fmt, c):
    """gst_tensor_converter_parse_video (gsttensor_converter.c:1421-1519)."""
    for w, h in ((640, 480), (642, 480), (1918, 1080)):
        conv = TensorConverter()
        out = conv.set_caps("video/x-raw, format=(string)%s, width=(int)%d, height=(int)%d, "
                            "framerate=(fraction)30/1" % (fmt, w, h))
        assert "dimensions=(string)%d:%d:%d:1:1:1:1:1" % (c, w, h) in out
        assert "types=(string)uint8" in out and "framerate=(fraction)30/1" in out
        pad = c in (1, 3) and w % 4 != 0            # gst_tensor_converter_video_stride (:1390-1411)
        assert conv.remove_padding == pad
        assert conv.frame_size == ((c * w + 3) // 4 * 4) * h
        conv.close()


def test_converter_frames_per_tensor_and_errors():
    conv = TensorConverter(frames_per_tensor=32)
    assert conv.get_property("frames-per-tensor") == 32
    out = conv.set_caps("video/x-raw, format=(string)RGB, width=(int)1920, height=(int)1080, framerate=(fraction)30/1")
    assert "dimensions=(string)3:1920:1080:32:1:1:1:1" in out     # :2132,:2204-2206
    with pytest.raises(L.NnsbError):
        conv.set_property("frames-per-tensor", 0)
    with pytest.raises(L.NnsbError):
        conv.set_caps("video/x-raw, format=(string)I420, width=(int)640, height=(int)480")
    with pytest.raises(L.NnsbError):
        conv.set_caps("audio/x-raw, format=(string)S16LE, rate=(int)16000")
    with pytest.raises(L.NnsbError):
        conv.set_caps("video/x-raw, format=(string)RGB, width=(int)[ 1, 100 ], height=(int)480")
    conv.close()
