"""tensor_decoder mode=direct_video on the device and the re-pitch kernels
(K8 strip / K9 pad).  Reference: ext/nnstreamer/tensor_decoder/tensordec-directvideo.c; the
reference's own pipeline test round-trips video -> tensor_converter -> tensor_decoder and
compares with the source frames (tests/nnstreamer_decoder/runTest.sh:36-56,71-77), which is the
property checked here at several widths, including 642x480 (its padded golden size)."""
import numpy as np
import pytest

import nnstreamer_b200 as nb
from nnstreamer_b200 import DirectVideoDecoder, TensorConverter
from oracle import oracle as O


@pytest.fixture(scope="module")
def port():
    return O.port()


@pytest.fixture(scope="module")
def dev():
    d = nb.Device(0)
    yield d
    d.close()


# ---------------------------------------------------------------------------------------
# CPU: oracle properties, option / caps callbacks
# ---------------------------------------------------------------------------------------
def test_oracle_pad_inverts_strip(port):
    rng = np.random.default_rng(1)
    for color, w, h in ((3, 642, 48), (1, 641, 7), (3, 10, 5), (4, 7, 3), (1, 16, 4)):
        dense = rng.integers(0, 256, color * w * h, dtype=np.uint8)
        padded = port.direct_video(dense, color, w, h)
        pitch = (color * w + 3) // 4 * 4
        assert padded.size == pitch * h
        assert np.array_equal(port.stride_strip(padded, color, w, h), dense)
        rows = padded.reshape(h, pitch)
        assert not rows[:, color * w:].any()


def test_options_and_caps():
    dv = DirectVideoDecoder()
    cfg = dv.make_config(3, 640, 480, (30, 1))
    # default formats per channel count, tensordec-directvideo.c:176-243
    assert dv.out_caps(cfg) == ("video/x-raw, format=(string)RGB, width=(int)640, height=(int)480, "
                                "framerate=(fraction)30/1, views=(int)1, "
                                "interlace-mode=(string)progressive")
    assert "format=(string)GRAY8" in dv.out_caps(dv.make_config(1, 64, 48))
    assert "format=(string)BGRx" in dv.out_caps(dv.make_config(4, 64, 48))
    with pytest.raises(nb.NnsbError):
        dv.out_caps(dv.make_config(2, 64, 48))          # "2 channel is not supported"
    assert dv.set_option(0, "BGR") and "format=(string)BGR," in dv.out_caps(cfg)
    with pytest.raises(nb.NnsbError):
        dv.out_caps(dv.make_config(4, 64, 48))          # BGR with 4 channels: invalid format
    assert dv.set_option(0, "rgba")                     # case-insensitive like find_key_strv
    assert "format=(string)RGBA" in dv.out_caps(dv.make_config(4, 64, 48))
    assert not dv.set_option(0, "I420") and not dv.set_option(0, "") and not dv.set_option(0, None)
    # _get_video_xraw_bufsize, :265-270
    assert dv.transform_size(dv.make_config(3, 642, 480)) == 1928 * 480
    assert dv.transform_size(dv.make_config(3, 640, 480)) == 1920 * 480
    assert dv.transform_size(dv.make_config(1, 641, 7)) == 644 * 7
    dv.close()


# ---------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------
SHAPES = [(3, 642, 480), (1, 641, 37), (3, 1918, 1080), (3, 10, 5), (1, 5, 3), (3, 21, 17),
          (4, 7, 3), (3, 640, 48), (1, 1023, 9), (3, 3, 2)]


@pytest.mark.gpu
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "%dx%dx%d" % s)
def test_pad_and_strip_kernels_match_oracle(dev, port, shape):
    color, w, h = shape
    rng = np.random.default_rng(w)
    row, pitch = color * w, (color * w + 3) // 4 * 4
    dense = rng.integers(0, 256, row * h, dtype=np.uint8)
    d_dense = dev.upload(dense)
    guard = 64
    d_pad = dev.alloc(pitch * h + 2 * guard)
    dev.h2d(d_pad, np.full(pitch * h + 2 * guard, 0xA5, dtype=np.uint8))
    dev.pad(d_dense, d_pad + guard, row, pitch, h)
    got = dev.download(d_pad, pitch * h + 2 * guard, np.uint8)
    assert np.all(got[:guard] == 0xA5) and np.all(got[-guard:] == 0xA5), "wrote outside the frame"
    want = port.direct_video(dense, color, w, h)
    assert np.array_equal(got[guard:-guard], want)
    # ... and back through the strip kernel
    d_back = dev.alloc(row * h)
    dev.strip(d_pad + guard, d_back, row, pitch, h)
    assert np.array_equal(dev.download(d_back, row * h, np.uint8), dense)
    for p in (d_dense, d_pad, d_back):
        dev.free(p)


@pytest.mark.gpu
def test_repitch_unaligned_pointers_take_the_byte_kernel(dev, port):
    color, w, h = 3, 35, 9
    row, pitch = color * w, (color * w + 3) // 4 * 4
    dense = np.random.default_rng(3).integers(0, 256, row * h, dtype=np.uint8)
    d_src = dev.alloc(row * h + 16)
    dev.h2d(d_src + 1, dense)
    d_dst = dev.alloc(pitch * h + 16)
    dev.pad(d_src + 1, d_dst + 3, row, pitch, h)
    got = dev.download(d_dst + 3, pitch * h, np.uint8)
    assert np.array_equal(got, port.direct_video(dense, color, w, h))
    dev.free(d_src)
    dev.free(d_dst)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(3, 642, 480), (1, 641, 37), (3, 640, 48), (4, 7, 3)],
                         ids=lambda s: "%dx%dx%d" % s)
def test_decode_all_memory_combinations(dev, port, shape):
    color, w, h = shape
    dv = DirectVideoDecoder()
    cfg = dv.make_config(color, w, h)
    size = dv.transform_size(cfg)
    dense = np.random.default_rng(h).integers(0, 256, color * w * h, dtype=np.uint8)
    want = port.direct_video(dense, color, w, h)
    pitch = (color * w + 3) // 4 * 4
    payload = np.zeros(size, dtype=bool).reshape(h, pitch)
    payload[:, :color * w] = True
    payload = payload.reshape(-1)
    d_dense = dev.upload(dense)
    for src in (dense, nb.device_memory(d_dense, dense.nbytes)):
        # allocate
        flow, mem = dv.decode(cfg, src)
        assert flow == 0 and mem.size == size
        assert np.array_equal(dev.download(mem.ptr, size, np.uint8), want)
        mem.release()
        # fill a given host buffer (padding bytes stay as they were, like the reference)
        host = np.full(size, 0x5A, dtype=np.uint8)
        flow, _ = dv.decode(cfg, src, host)
        assert flow == 0 and np.array_equal(host[payload], want[payload])
        assert np.all(host[~payload] == 0x5A)
        # fill a given device buffer
        d_out = dev.alloc(size)
        flow, _ = dv.decode(cfg, src, nb.device_memory(d_out, size))
        assert flow == 0 and np.array_equal(dev.download(d_out, size, np.uint8), want)
        dev.free(d_out)
    dev.free(d_dense)
    # wrong tensor type: g_assert (type == _NNS_UINT8), :304
    bad = dv.make_config(color, w, h, type_name="float32")
    flow, _ = dv.decode(bad, dense)
    assert flow != 0
    dv.close()


@pytest.mark.gpu
def test_converter_then_decoder_round_trip_on_device(dev):
    """video -> tensor_converter -> tensor_decoder direct_video == video (payload bytes), the
    reference's nnstreamer_decoder SSAT, with every hop in HBM."""
    w, h = 642, 480
    pitch = (3 * w + 3) // 4 * 4
    frame = np.random.default_rng(9).integers(0, 256, pitch * h, dtype=np.uint8)
    conv = TensorConverter()
    conv.set_caps("video/x-raw, format=(string)RGB, width=(int)%d, height=(int)%d, "
                  "framerate=(fraction)30/1" % (w, h))
    flow, tensor = conv.chain(frame)
    assert flow == 0 and tensor is not None and tensor.size == 3 * w * h
    dv = DirectVideoDecoder(option1="RGB")
    cfg = dv.make_config(3, w, h, (30, 1))
    flow, video = dv.decode(cfg, tensor)
    assert flow == 0
    got = dev.download(video.ptr, video.size, np.uint8).reshape(h, pitch)
    assert np.array_equal(got[:, :3 * w], frame.reshape(h, pitch)[:, :3 * w])
    video.release()
    tensor.release()
    conv.close()
    dv.close()
