"""GPU: the BASELINE.json configurations at FULL size, checked through size-independent
properties (round trips, idempotence, batch == per-frame, numpy definitions) plus oracle
comparisons on the frames the CPU can afford (first / last)."""
import numpy as np
import pytest

from helpers import bits_equal

pytestmark = pytest.mark.gpu

W, H, C = 1920, 1080, 3
FRAME = C * W * H
IMAGENET = ("typecast:float32,per-channel:true@0,add:-123.675@0,add:-116.28@1,add:-103.53@2,"
            "div:58.395@0,div:57.12@1,div:57.375@2")
IMAGENET_OPS = [("typecast", "float32"), ("add", -123.675, 0), ("add", -116.28, 1), ("add", -103.53, 2),
                ("div", 58.395, 0), ("div", 57.12, 1), ("div", 57.375, 2)]


def test_cfg2_1080p_batch32_per_channel_arithmetic(dev, port):
    """configs[1]: 3:1920:1080:32 u8 -> f32, 199 065 600 elements in one launch."""
    B = 32
    rng = np.random.default_rng(2)
    x = rng.integers(0, 256, FRAME * B, dtype=np.uint8)
    plan = dev.plan("arithmetic", IMAGENET, "uint8", [C, W, H, 1])
    y = plan.run_host(x, batch=B)
    # (a) oracle on the first and the last frame
    for f in (0, B - 1):
        want = port.arithmetic(x[f * FRAME:(f + 1) * FRAME], [C, W, H, 1], IMAGENET_OPS, True, 0)
        assert bits_equal(y[f * FRAME:(f + 1) * FRAME], want), "frame %d" % f
    # (b) the chain is a function of (channel, byte): every element must equal the table the
    #     oracle produces for the 3 x 256 possible inputs
    probe = np.tile(np.arange(256, dtype=np.uint8).repeat(3), 1)          # value v on all 3 channels
    table = port.arithmetic(probe, [3, 256, 1, 1], IMAGENET_OPS, True, 0).reshape(256, 3)
    ch = np.arange(FRAME * B, dtype=np.int64) % 3
    assert bits_equal(y, table[x, ch])
    # (c) one batched launch == one tensor with the frame dimension folded in
    plan2 = dev.plan("arithmetic", IMAGENET, "uint8", [C, W, H, B])
    assert bits_equal(plan2.run_host(x), y)
    plan.close()
    plan2.close()


def test_cfg3_4k_transpose_and_dimchg_round_trips(dev):
    """configs[2]: 3:3840:2160 f32; HWC->CHW == numpy transpose; CHW->HWC restores the input;
    dimchg 0:2 is the same permutation; the other four orders against numpy."""
    dim = [3, 3840, 2160, 1]
    n = int(np.prod(dim))
    rng = np.random.default_rng(3)
    x = rng.standard_normal(n, dtype=np.float32)
    plan = dev.plan("transpose", "1:2:0:3", "float32", dim)
    y = plan.run_host(x)
    plan.close()
    assert bits_equal(y, x.reshape(2160, 3840, 3).transpose(2, 0, 1).reshape(-1))
    back = dev.plan("transpose", "2:0:1:3", "float32", [3840, 2160, 3, 1])
    assert bits_equal(back.run_host(y), x)
    back.close()
    d = dev.plan("dimchg", "0:2", "float32", dim)
    assert bits_equal(d.run_host(x), y)
    d.close()
    a = x.reshape(2160, 3840, 3)                     # numpy axes (d2, d1, d0)
    for order in ((0, 2, 1), (1, 0, 2), (2, 0, 1), (2, 1, 0)):
        p = dev.plan("transpose", "%d:%d:%d:3" % order, "float32", dim)
        got = p.run_host(x)
        p.close()
        # out dim i = in dim order[i]  ->  numpy axes are reversed dims
        axes = tuple(2 - order[2 - k] for k in range(3))
        assert bits_equal(got, a.transpose(axes).reshape(-1)), order


def test_cfg4_clamp_1080p_batch64(dev):
    """configs[3]: clamp on 3:1920:1080:64 f32 == np.clip, and clamp is idempotent."""
    B = 64
    rng = np.random.default_rng(5)
    x = (rng.standard_normal(FRAME * B, dtype=np.float32) * 50)
    plan = dev.plan("clamp", "-33.3:77.7", "float32", [C, W, H, 1])
    y = plan.run_host(x, batch=B)
    want = np.clip(x.astype(np.float64), -33.3, 77.7).astype(np.float32)
    assert bits_equal(y, want)
    assert bits_equal(plan.run_host(y, batch=B), y)
    plan.close()


@pytest.mark.parametrize("option,per_channel", [("default", False), ("dc-average,per-channel:true", True)])
def test_cfg4_stand_1080p_batch16_per_frame(dev, port, option, per_channel):
    """configs[3]: stand on 1080p f32 frames (16 queued frames; statistics per frame)."""
    B = 16
    rng = np.random.default_rng(4)
    x = rng.random(FRAME * B, dtype=np.float32) * 10
    plan = dev.plan("stand", option, "float32", [C, W, H, 1])
    y = plan.run_host(x, batch=B)
    nch = 3 if per_channel else 1
    mean, std = plan.stand_stats(B, nch)
    plan.close()
    x64 = x.astype(np.float64).reshape(B, -1, nch)
    np.testing.assert_allclose(mean.reshape(B, nch), x64.mean(axis=1), rtol=1e-12)
    if option == "default":
        np.testing.assert_allclose(std.reshape(B, nch), x64.std(axis=1), rtol=1e-10)
        ref = np.abs((x64 - x64.mean(axis=1, keepdims=True)) / x64.std(axis=1, keepdims=True))
    else:
        ref = x64 - x64.mean(axis=1, keepdims=True)
    g = y.astype(np.float64).reshape(B, -1, nch)
    assert np.all(np.abs(g - ref) <= 1e-6 * np.abs(ref) + 1e-6)
    # oracle (sequential running mean) on the first frame
    w = port.stand(x[:FRAME], [C, W, H, 1], option.split(",")[0], None, per_channel).astype(np.float64)
    assert np.all(np.abs(y[:FRAME].astype(np.float64) - w) <= 1e-6 * np.abs(w) + 1e-6)


def _stand_close_chunked(got, want, scale, chunk=1 << 24):
    """|gpu - reference| <= 1e-6 |reference| + 1e-6 scale, without a float64 copy of 400 M values."""
    worst = 0.0
    for a in range(0, got.size, chunk):
        g = got[a:a + chunk].astype(np.float64)
        w = want[a:a + chunk].astype(np.float64)
        err = np.abs(g - w) - (1e-6 * np.abs(w) + 1e-6 * scale)
        worst = max(worst, float(err.max()))
    assert worst <= 0.0, "beyond tolerance by %g" % worst


def _cfg4_frames(batch, seed):
    """`batch` 1080p float32 frames whose offset and spread differ from frame to frame, so
    statistics taken over the wrong range (one frame instead of all, or the reverse) show up."""
    rng = np.random.default_rng(seed)
    x = rng.random(FRAME * batch, dtype=np.float32)
    x = x.reshape(batch, FRAME)
    x *= np.linspace(1.0, 30.0, batch, dtype=np.float32)[:, None]
    x += np.linspace(-5.0, 50.0, batch, dtype=np.float32)[:, None]
    return x.reshape(-1)


@pytest.mark.parametrize("option,mode,per_channel", [("default", "default", False),
                                                     ("dc-average,per-channel:true", "dc-average", True)])
def test_cfg4_stand_1080p_batch64_per_frame_vs_reference(dev, port, option, mode, per_channel):
    """configs[3] at its stated size: 64 queued 1080p f32 frames in one set of launches; every
    frame against the CPU reference's sequential loops (tensor_data.c:314-493)."""
    B = 64
    x = _cfg4_frames(B, 41)
    plan = dev.plan("stand", option, "float32", [C, W, H, 1])
    y = plan.run_host(x, batch=B)
    plan.close()
    for f in range(B):
        want = port.stand(x[f * FRAME:(f + 1) * FRAME], [C, W, H, 1], mode, None, per_channel)
        _stand_close_chunked(y[f * FRAME:(f + 1) * FRAME], want, 60.0 if mode == "dc-average" else 1.0)


@pytest.mark.parametrize("option,mode,per_channel", [
    ("default", "default", False), ("default,per-channel:true", "default", True),
    ("dc-average", "dc-average", False), ("dc-average,per-channel:true", "dc-average", True)])
def test_cfg4_stand_one_tensor_3_1920_1080_64_vs_reference(dev, port, option, mode, per_channel):
    """The reference's own semantics for a batched tensor: ONE tensor 3:1920:1080:64, statistics
    over all 398 M elements (per channel: 133 M each) -- where a one-pass variance would drift."""
    B = 64
    dim = [C, W, H, B]
    x = _cfg4_frames(B, 43)
    plan = dev.plan("stand", option, "float32", dim)
    y = plan.run_host(x)
    nch = 3 if per_channel else 1
    mean, std = plan.stand_stats(1, nch)
    plan.close()
    omean, ostd = port.stand_stats(x, dim, per_channel)
    # the reference's running mean over 4e8 terms carries its own rounding (random walk ~1e-12)
    np.testing.assert_allclose(mean, omean, rtol=1e-9)
    if mode == "default":
        np.testing.assert_allclose(std, ostd, rtol=1e-9)
    want = port.stand(x, dim, mode, None, per_channel)
    _stand_close_chunked(y, want, 100.0 if mode == "dc-average" else 1.0)


def test_cfg5_chain_converter_typecast_arith_transpose(dev, port):
    """configs[4], one stream: 1918x1080 RGB (padded rows) -> converter -> typecast -> arithmetic
    -> transpose, all on the device; equals the oracle chain on the same frame."""
    from nnstreamer_b200 import TensorConverter, TensorTransform
    from nnstreamer_b200 import _lib as L
    w, h = 1918, 1080
    conv = TensorConverter()
    c1 = conv.set_caps("video/x-raw, format=(string)RGB, width=(int)%d, height=(int)%d, framerate=(fraction)30/1" % (w, h))
    assert conv.remove_padding
    t1 = TensorTransform(mode="typecast", option="float32")
    c2 = t1.negotiate(c1)
    t2 = TensorTransform(mode="arithmetic", option="add:-127.5,div:127.5")
    c3 = t2.negotiate(c2)
    t3 = TensorTransform(mode="transpose", option="1:2:0:3")
    c4 = t3.negotiate(c3)
    assert "dimensions=(string)%d:%d:3:1" % (w, h) in c4
    pitch = (3 * w + 3) // 4 * 4
    rng = np.random.default_rng(7)
    frame = rng.integers(0, 256, pitch * h, dtype=np.uint8)
    flow, m0 = conv.chain(frame)
    assert flow == L.FLOW_OK and m0 is not None
    flow, (m1,) = t1.transform([m0])
    flow, (m2,) = t2.transform([m1])
    flow, (m3,) = t3.transform([m2])
    assert flow == L.FLOW_OK
    got = dev.download(m3.ptr, 3 * w * h, np.float32)
    dense = port.stride_strip(frame, 3, w, h)
    f = port.arithmetic(dense, [3, w, h, 1], [("typecast", "float32"), ("add", -127.5), ("div", 127.5)])
    want = port.transpose(f, [3, w, h, 1], (1, 2, 0, 3))
    assert bits_equal(got, want)
    for m in (m0, m1, m2, m3):
        m.release()
    for e in (conv, t1, t2, t3):
        e.close()
