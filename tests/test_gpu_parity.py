"""GPU parity: the CUDA kernels (through the C-ABI) against the CPU oracle, same inputs.

Bars (BASELINE.json north_star / SURVEY.md 8c):
  typecast, arithmetic (integer AND float chains), transpose, dimchg, clamp, stride strip:
      BIT-EXACT against the oracle (float chains: FMA contraction off, IEEE division);
  stand: |gpu - oracle| <= 1e-6 * |oracle| + 1e-6 * scale  (parallel fp64 reduction vs the
      reference's sequential running mean; the reference's own test allows 1e-5 absolute).
"""
import numpy as np
import pytest

from helpers import ORC_TYPES, TYPES, bits_equal, first_mismatch, rand_full, rand_small

pytestmark = pytest.mark.gpu


# ---------------------------------------------------------------------------------------
# typecast
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("itype", TYPES)
def test_typecast_all_pairs_c_path(dev, port, itype):
    n = 10007                        # odd: exercises the scalar tail
    x = rand_full(itype, n, seed=11)
    for otype in TYPES:
        plan = dev.plan("typecast", otype, itype, [n])
        got = plan.run_host(x)
        plan.close()
        want = port.typecast(x, otype)
        assert bits_equal(got, want), "%s->%s %s" % (itype, otype, first_mismatch(got, want, x))


@pytest.mark.parametrize("itype", ORC_TYPES)
def test_typecast_all_pairs_orc_path(dev, port, itype):
    n = 4099
    x = rand_full(itype, n, seed=12)
    for otype in ORC_TYPES:
        plan = dev.plan("typecast", otype, itype, [n], acceleration=True)
        got = plan.run_host(x)
        plan.close()
        want = port.typecast(x, otype, accel=True)
        assert bits_equal(got, want), "%s->%s %s" % (itype, otype, first_mismatch(got, want, x))


def test_typecast_empty_and_tiny(dev, port):
    for n in (1, 2, 3, 7, 8, 9, 31, 33):
        x = rand_full("uint8", n, seed=n)
        plan = dev.plan("typecast", "float32", "uint8", [n])
        got = plan.run_host(x)
        plan.close()
        assert bits_equal(got, port.typecast(x, "float32"))


def test_typecast_unaligned_device_pointers(dev, port):
    """Flexible tensors put the payload 128 bytes into a block; users may pass any pointer."""
    n = 5000
    x = rand_full("uint8", n, seed=5)
    plan = dev.plan("typecast", "float32", "uint8", [n])
    d_in = dev.alloc(n + 64)
    d_out = dev.alloc(4 * n + 64)
    for off_in, off_out in ((1, 0), (0, 4), (3, 12), (128, 128)):
        dev.h2d(d_in + off_in, x)
        plan.launch(d_in + off_in, d_out + off_out, 1)
        got = dev.download(d_out + off_out, n, np.float32)
        assert bits_equal(got, port.typecast(x, "float32")), (off_in, off_out)
    dev.free(d_in)
    dev.free(d_out)
    plan.close()


# ---------------------------------------------------------------------------------------
# arithmetic
# ---------------------------------------------------------------------------------------
def _arith_case(dev, port, x, dim, option, ops, per_channel=False, ch_dim=0, accel=False, batch=1):
    plan = dev.plan("arithmetic", option, x.dtype.name, dim, acceleration=accel)
    got = plan.run_host(x, batch=batch)
    plan.close()
    if batch == 1:
        want = port.arithmetic(x, dim, ops, per_channel, ch_dim, accel)
    else:
        per = x.size // batch
        want = np.concatenate([port.arithmetic(x[i * per:(i + 1) * per], dim, ops, per_channel,
                                               ch_dim, accel) for i in range(batch)])
    assert bits_equal(got, want), "%s %s: %s" % (x.dtype, option, first_mismatch(got, want, x))


def test_arith_mobilenet_chain_exhaustive_u8(dev, port):
    """cfg1: typecast:float32,add:-127.5,div:127.5 on every u8 value, 224x224 RGB."""
    rng = np.random.default_rng(1)
    x = rng.integers(0, 256, 3 * 224 * 224, dtype=np.uint8)
    x[:256] = np.arange(256, dtype=np.uint8)
    _arith_case(dev, port, x, [3, 224, 224, 1], "typecast:float32,add:-127.5,div:127.5",
                [("typecast", "float32"), ("add", -127.5), ("div", 127.5)])


def test_arith_imagenet_per_channel(dev, port):
    """cfg2 chain (per-channel mean/std) on a small frame, batch 1 and 3."""
    opt = ("typecast:float32,per-channel:true@0,add:-123.675@0,add:-116.28@1,add:-103.53@2,"
           "div:58.395@0,div:57.12@1,div:57.375@2")
    ops = [("typecast", "float32"), ("add", -123.675, 0), ("add", -116.28, 1), ("add", -103.53, 2),
           ("div", 58.395, 0), ("div", 57.12, 1), ("div", 57.375, 2)]
    rng = np.random.default_rng(2)
    for (w, h, b) in ((64, 48, 1), (61, 37, 1), (64, 48, 3), (1920, 8, 2)):
        x = rng.integers(0, 256, 3 * w * h * b, dtype=np.uint8)
        _arith_case(dev, port, x, [3, w, h, 1], opt, ops, True, 0, batch=b)


@pytest.mark.parametrize("ch_dim", [0, 1, 2, 3])
def test_arith_per_channel_every_dim(dev, port, ch_dim):
    dim = [3, 5, 4, 2]
    rng = np.random.default_rng(30 + ch_dim)
    x = rng.standard_normal(int(np.prod(dim))).astype(np.float32)
    nch = dim[ch_dim]
    ops = [("add", 1.5, 0), ("mul", 2, nch - 1), ("div", 3.0), ("add", -7, 1 if nch > 1 else 0),
           ("mul", 0.25, 99)]           # channel 99 never matches
    opt = "per-channel:true@%d,add:1.5@0,mul:2@%d,div:3.0,add:-7@%d,mul:0.25@99" % (
        ch_dim, nch - 1, 1 if nch > 1 else 0)
    _arith_case(dev, port, x, dim, opt, ops, True, ch_dim)


@pytest.mark.parametrize("dtype", TYPES)
def test_arith_chains_every_type_c_path(dev, port, dtype):
    n = 3 * 50 * 41
    x = rand_full(dtype, n, seed=40)
    cases = [
        ("mul:2", [("mul", 2)]),
        ("add:-10,mul:3,div:7", [("add", -10), ("mul", 3), ("div", 7)]),
        ("mul:-5.5", [("mul", -5.5)]),
        ("add:9.9e-1", [("add", 9.9e-1)]),
        ("div:0,add:1", [("div", 0), ("add", 1)]),          # zero divisor is skipped
        ("div:-1", [("div", -1)]),
        ("typecast:int32,mul:2,div:2,add:-1",
         [("typecast", "int32"), ("mul", 2), ("div", 2), ("add", -1)]),
        ("typecast:float64,add:0.2,add:0.1", [("typecast", "float64"), ("add", 0.2), ("add", 0.1)]),
        ("typecast:float32,add:.5,mul:0.2", [("typecast", "float32"), ("add", .5), ("mul", 0.2)]),
        ("typecast:uint8,add:300", [("typecast", "uint8"), ("add", 300)]),
        ("typecast:float16,mul:1.5,add:3", [("typecast", "float16"), ("mul", 1.5), ("add", 3)]),
    ]
    for opt, ops in cases:
        _arith_case(dev, port, x, [3, 50, 41], opt, ops)


@pytest.mark.parametrize("dtype", ORC_TYPES)
def test_arith_chains_orc_path(dev, port, dtype):
    """acceleration=true: saturating integer add/mul, ORC conversions."""
    n = 3 * 33 * 17
    x = rand_full(dtype, n, seed=41)
    cases = [
        ("add:100", [("add", 100)]),
        ("add:-100", [("add", -100)]),
        ("mul:3", [("mul", 3)]),
        ("mul:-3,add:7", [("mul", -3), ("add", 7)]),
        ("div:3", [("div", 3)]),
        ("typecast:int8,add:100,mul:2", [("typecast", "int8"), ("add", 100), ("mul", 2)]),
        ("typecast:uint16,mul:300", [("typecast", "uint16"), ("mul", 300)]),
        ("typecast:float32,add:-127.5,div:127.5",
         [("typecast", "float32"), ("add", -127.5), ("div", 127.5)]),
        ("typecast:int16,add:1", [("typecast", "int16"), ("add", 1)]),
    ]
    for opt, ops in cases:
        _arith_case(dev, port, x, [3, 33, 17], opt, ops, accel=True)


@pytest.mark.parametrize("itype", ["uint8", "int8"])
@pytest.mark.parametrize("accel", [False, True])
def test_arith_one_byte_inputs_table_path(dev, port, itype, accel):
    """>= 65536 one-byte elements take the shared-memory table kernel; it must equal the
    oracle for every output type, scalar and per-channel (innermost, middle and odd channel
    counts, i.e. both the loop-invariant and the walking channel mapping)."""
    rng = np.random.default_rng(55)
    info = np.iinfo(itype)
    outs = ["float32", "uint8", "int16", "float64"] if accel else \
        ["float32", "uint8", "int8", "int16", "uint16", "int32", "float64", "float16", "int64"]
    for otype in outs:
        dim = [3, 160, 144, 1]                  # 69120 elements
        x = rng.integers(info.min, info.max, 69120, dtype=itype, endpoint=True)
        _arith_case(dev, port, x, dim, "typecast:%s,add:-12,mul:3,div:7" % otype,
                    [("typecast", otype), ("add", -12), ("mul", 3), ("div", 7)], accel=accel)
        _arith_case(dev, port, x, dim,
                    "typecast:%s,per-channel:true@0,add:-12@0,add:5@1,mul:3@2,div:7@0,div:2" % otype,
                    [("typecast", otype), ("add", -12, 0), ("add", 5, 1), ("mul", 3, 2), ("div", 7, 0),
                     ("div", 2)], True, 0, accel=accel)
    # channel dimension in the middle (run length 160*3) and outermost; 7 channels (odd period)
    for dim, ch_dim in (([3, 160, 144, 1], 1), ([3, 160, 144, 1], 2), ([7, 9877, 1, 1], 0),
                        ([5, 7, 1999, 1], 1)):
        n = int(np.prod(dim))
        x = rng.integers(info.min, info.max, n, dtype=itype, endpoint=True)
        nch = dim[ch_dim]
        _arith_case(dev, port, x, dim,
                    "typecast:float32,per-channel:true@%d,add:1.5@0,mul:2@%d,div:3.0" % (ch_dim, nch - 1),
                    [("typecast", "float32"), ("add", 1.5, 0), ("mul", 2, nch - 1), ("div", 3.0)],
                    True, ch_dim, accel=accel)
    # no typecast: u8 -> u8 chain, batch of 2
    x = rng.integers(info.min, info.max, 2 * 69120, dtype=itype, endpoint=True)
    _arith_case(dev, port, x, [3, 160, 144, 1], "add:100,mul:2", [("add", 100), ("mul", 2)],
                accel=accel, batch=2)


def test_arith_float_chain_random_f32_bitexact(dev, port):
    """Random f32 inputs over many magnitudes incl. denormals: add/mul/div rounding per op."""
    rng = np.random.default_rng(77)
    n = 1 << 16
    x = (rng.standard_normal(n) * 10.0 ** rng.integers(-44, 38, n)).astype(np.float32)
    x[:4] = [np.float32(1e-45), np.float32(-1e-45), np.float32(3.4e38), np.float32(0)]
    for opt, ops in (("div:3.0", [("div", 3.0)]), ("div:127.5", [("div", 127.5)]),
                     ("mul:1e-30,div:1e-20", [("mul", 1e-30), ("div", 1e-20)]),
                     ("add:1e10,add:-1e10", [("add", 1e10), ("add", -1e10)])):
        _arith_case(dev, port, x, [n], opt, ops)
    xd = x.astype(np.float64) * 1.000000001
    _arith_case(dev, port, xd, [n], "div:3.0,mul:1.7", [("div", 3.0), ("mul", 1.7)])


FAST_CHAINS = [
    # option, oracle ops, per-channel dim (None = scalar chain)
    ("div:127.5", [("div", 127.5)], None),
    ("add:-127.5,div:127.5", [("add", -127.5), ("div", 127.5)], None),
    ("mul:0.017,div:3.0,add:1.5", [("mul", 0.017), ("div", 3.0), ("add", 1.5)], None),
    ("div:0.1,div:-7e8", [("div", 0.1), ("div", -7e8)], None),
    ("div:1e-5,mul:1e-5", [("div", 1e-5), ("mul", 1e-5)], None),
    ("div:1.0000001", [("div", 1.0000001)], None),
    ("div:1.9999999", [("div", 1.9999999)], None),          # significand all ones: generic kernel
    ("div:1e12", [("div", 1e12)], None),                    # outside 2^+-30: generic kernel
    ("per-channel:true@0,add:-123.675@0,add:-116.28@1,add:-103.53@2,div:58.395@0,div:57.12@1,div:57.375@2",
     [("add", -123.675, 0), ("add", -116.28, 1), ("add", -103.53, 2), ("div", 58.395, 0),
      ("div", 57.12, 1), ("div", 57.375, 2)], 0),
    ("per-channel:true@1,div:3@0,mul:2@1,div:-0.7@3", [("div", 3, 0), ("mul", 2, 1), ("div", -0.7, 3)], 1),
]


@pytest.mark.parametrize("case", FAST_CHAINS, ids=lambda c: c[0][:40])
def test_float32_fast_chain_is_bit_identical_to_the_careful_path(dev, port, case, monkeypatch):
    """fchain_kernel (division by constant through a correctly rounded reciprocal, NaN check per
    vector) against the generic kernel on 2^23 arbitrary bit patterns (NaNs with payloads,
    infinities, denormals, zeros of both signs) and 2^23 image-range values, and against the
    oracle (gsttensor_transform.c:563-645) on a subset.  Bit-exact."""
    opt, ops, ch_dim = case
    n = 1 << 23
    dim = [4, 1 << 10, n >> 12, 1] if ch_dim is not None else [n]
    rng = np.random.default_rng(len(opt))
    bits = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    bits[:64] = np.array([0, 0x80000000, 1, 0x80000001, 0x7f800000, 0xff800000, 0x7fc00000, 0x7f800001,
                          0x00800000, 0x007fffff, 0x7f7fffff, 0x1f800000, 0x5f000000, 0x5f800000,
                          0x1f7fffff, 0x20000000] * 4, dtype=np.uint32)
    if ch_dim is not None:
        # Channels an operator does not apply to go through an identity operand (x + -0, x / 1):
        # exact for every value except that a SIGNALLING NaN comes out quieted where the reference
        # leaves the element untouched -- keep signalling NaNs out of here.
        snan = ((bits & 0x7f800000) == 0x7f800000) & ((bits & 0x007fffff) != 0) & ((bits & 0x00400000) == 0)
        bits[snan] |= 0x00400000
    for x in (bits.view(np.float32), (rng.random(n, dtype=np.float32) * 255.0)):
        monkeypatch.delenv("NNSB_CHAIN_NO_FAST", raising=False)
        fast = dev.plan("arithmetic", opt, "float32", dim)
        monkeypatch.setenv("NNSB_CHAIN_NO_FAST", "1")
        slow = dev.plan("arithmetic", opt, "float32", dim)
        monkeypatch.delenv("NNSB_CHAIN_NO_FAST", raising=False)
        assert "fast path" not in slow.describe()
        if "1.9999999" in opt or "1e12" in opt:
            assert "fast path" not in fast.describe()
        else:
            assert "fast path" in fast.describe(), fast.describe()
        got = fast.run_host(x)
        ref = slow.run_host(x)
        fast.close()
        slow.close()
        assert bits_equal(got, ref), "%s: %s" % (opt, first_mismatch(got, ref))
        m = 1 << 16
        sub = [m] if ch_dim is None else [4, 1 << 10, m >> 12, 1]
        want = port.arithmetic(x[:m], sub, ops, ch_dim is not None, ch_dim or 0)
        assert bits_equal(got[:m], want), "%s vs oracle: %s" % (opt, first_mismatch(got[:m], want))


@pytest.mark.parametrize("in_type", ["int32", "uint16", "int16", "float64", "int64", "float16"])
def test_float32_fast_chain_other_input_types(dev, port, in_type):
    n = 3 * 640 * 48
    dim = [3, 640, 48, 1]
    x = rand_small(in_type, n, seed=77)
    opt = ("typecast:float32,per-channel:true@0,add:-123.675@0,add:-116.28@1,add:-103.53@2,"
           "div:58.395@0,div:57.12@1,div:57.375@2")
    ops = [("typecast", "float32"), ("add", -123.675, 0), ("add", -116.28, 1), ("add", -103.53, 2),
           ("div", 58.395, 0), ("div", 57.12, 1), ("div", 57.375, 2)]
    plan = dev.plan("arithmetic", opt, in_type, dim)
    assert "fast path" in plan.describe()
    got = plan.run_host(x)
    plan.close()
    want = port.arithmetic(x, dim, ops, True, 0)
    assert bits_equal(got, want), first_mismatch(got, want)


# ---------------------------------------------------------------------------------------
# transpose / dimchg
# ---------------------------------------------------------------------------------------
ORDERS = [(0, 1, 2, 3), (0, 2, 1, 3), (1, 0, 2, 3), (1, 2, 0, 3), (2, 0, 1, 3), (2, 1, 0, 3)]


@pytest.mark.parametrize("order", ORDERS)
@pytest.mark.parametrize("dtype", ["uint8", "uint16", "float32", "float64"])
def test_transpose_all_orders(dev, port, order, dtype):
    for dim in ([3, 50, 100, 1], [100, 50, 3, 1], [7, 1, 5, 3], [33, 65, 17, 2], [1, 1, 9, 1]):
        n = int(np.prod(dim))
        x = rand_small(dtype, n, seed=50)
        opt = "%d:%d:%d:3" % order[:3]
        plan = dev.plan("transpose", opt, dtype, dim)
        got = plan.run_host(x)
        plan.close()
        want = port.transpose(x, dim, order)
        assert bits_equal(got, want), "%s %s %s: %s" % (dtype, dim, opt, first_mismatch(got, want))


def test_transpose_hwc_to_chw_batched(dev, port):
    dim = [3, 64, 36, 1]
    n = int(np.prod(dim))
    rng = np.random.default_rng(3)
    x = rng.standard_normal(n * 4).astype(np.float32)
    plan = dev.plan("transpose", "1:2:0:3", "float32", dim)
    got = plan.run_host(x, batch=4)
    plan.close()
    want = np.concatenate([port.transpose(x[i * n:(i + 1) * n], dim, (1, 2, 0, 3)) for i in range(4)])
    assert bits_equal(got, want)
    # numpy cross-check of the oracle itself: C:W:H (innermost first) -> W:H:C
    np_want = x[:n].reshape(36, 64, 3).transpose(2, 0, 1).reshape(-1)
    assert bits_equal(want[:n], np_want)


@pytest.mark.parametrize("dtype", ["uint8", "uint16", "float32"])
@pytest.mark.parametrize("channels", [2, 3, 4])
def test_transpose_video_shapes_tma_path(dev, port, dtype, channels):
    """Large interleaved<->planar transposes take the TMA bulk-copy pipeline (k_plane.cu):
    full and partial tiles, batches, both directions, plus sizes that must fall back."""
    rng = np.random.default_rng(channels)
    for (w, h, batch) in ((640, 120, 1), (1000, 77, 1), (336, 100, 3), (333, 99, 1)):
        dim = [channels, w, h, 1]
        n = channels * w * h
        x = rand_small(dtype, n * batch, seed=w)
        # HWC -> CHW
        plan = dev.plan("transpose", "1:2:0:3", dtype, dim)
        got = plan.run_host(x, batch=batch)
        plan.close()
        want = np.concatenate([port.transpose(x[i * n:(i + 1) * n], dim, (1, 2, 0, 3))
                               for i in range(batch)])
        assert bits_equal(got, want), "to planar %s %s: %s" % (dtype, dim, first_mismatch(got, want))
        # and back: CHW (W:H:C) -> HWC with 2:0:1:3
        dim2 = [w, h, channels, 1]
        plan = dev.plan("transpose", "2:0:1:3", dtype, dim2)
        back = plan.run_host(got, batch=batch)
        plan.close()
        assert bits_equal(back, x), "round trip %s %s" % (dtype, dim)
        # dimchg 0:2 is the same permutation as transpose 1:2:0:3
        plan = dev.plan("dimchg", "0:2", dtype, dim)
        got2 = plan.run_host(x, batch=batch)
        plan.close()
        assert bits_equal(got2, want)


VEC_CASES = [
    # dtype, dim, batch: every box edge case of the table-driven 16-byte vector kernel
    ("float32", [3, 128, 96, 1], 1),      # whole boxes only
    ("float32", [3, 132, 100, 1], 2),     # partial boxes on both sides, batch
    ("float32", [4, 36, 20, 1], 3),       # tensor smaller than one full box
    ("float32", [8, 12, 52, 1], 1),
    ("float64", [3, 130, 98, 1], 2),      # 2-element vectors
    ("int64", [6, 34, 18, 1], 1),
    ("int32", [64, 64, 64, 1], 1),
    ("uint8", [3, 192, 112, 1], 2),       # 1-byte elements: 16 gathered per output vector
    ("uint8", [3, 160, 80, 1], 1),
    ("int8", [16, 48, 40, 1], 1),
    ("uint16", [3, 64, 48, 1], 3),        # 2-byte elements: 8 per vector
    ("float16", [5, 32, 24, 1], 1),
    ("uint8", [3, 1080, 64, 1], 1),       # 3240-byte rows: 8-byte vectors
    ("float32", [3, 227, 31, 1], 2),      # rows that are only 4-byte multiples: 4-byte vectors
    ("uint16", [3, 226, 34, 1], 1),
    ("uint32", [5, 7, 9, 1], 4),
    ("uint16", [3, 225, 113, 1], 1),      # odd byte counts: must fall back to the scalar kernel
]


@pytest.mark.parametrize("order", [(0, 2, 1, 3), (1, 0, 2, 3), (2, 0, 1, 3), (2, 1, 0, 3),
                                   (1, 2, 0, 3)])
@pytest.mark.parametrize("case", VEC_CASES, ids=lambda c: "%s-%s-b%d" % (c[0], "x".join(map(str, c[1])), c[2]))
def test_transpose_vector_path(dev, port, order, case):
    """k_perm_vec.cu against the oracle (gsttensor_transform.c:1417-1500), bit-exact; the plan
    must report the vector table for the aligned shapes and the scalar kernel for the odd one."""
    dtype, dim, batch = case
    n = int(np.prod(dim))
    x = rand_small(dtype, n * batch, seed=n + batch)
    opt = "%d:%d:%d:3" % order[:3]
    plan = dev.plan("transpose", opt, dtype, dim)
    desc = plan.describe()
    got = plan.run_host(x, batch=batch)
    plan.close()
    want = np.concatenate([port.transpose(x[i * n:(i + 1) * n], dim, order) for i in range(batch)])
    assert bits_equal(got, want), "%s %s %s [%s]: %s" % (dtype, dim, opt, desc,
                                                          first_mismatch(got, want))
    if dim == [3, 225, 113, 1] and order[:3] != (1, 2, 0) and order[:3] != (1, 0, 2):
        assert "scalar" in desc, desc
    elif dim in ([3, 128, 96, 1], [64, 64, 64, 1], [3, 132, 100, 1]):
        assert "vector table" in desc, desc


@pytest.mark.parametrize("order", [(0, 2, 1, 3), (0, 1, 2, 3)])
def test_transpose_innermost_block_folds_into_element(dev, port, order):
    """A block that is innermost on both sides (RGBA bytes, a float4) moves as one wide element."""
    for dtype, dim, batch in (("uint8", [4, 68, 36, 1], 2), ("uint8", [16, 20, 12, 1], 1),
                              ("uint8", [32, 20, 12, 1], 1), ("uint8", [6, 20, 12, 1], 1),
                              ("uint16", [2, 36, 24, 1], 3), ("uint16", [8, 12, 20, 1], 1),
                              ("float32", [4, 36, 20, 1], 2), ("float32", [2, 132, 36, 1], 1),
                              ("float64", [2, 36, 20, 1], 1), ("uint8", [4, 5, 3, 1], 3)):
        n = int(np.prod(dim))
        x = rand_small(dtype, n * batch, seed=n)
        opt = "%d:%d:%d:3" % order[:3]
        plan = dev.plan("transpose", opt, dtype, dim)
        desc = plan.describe()
        got = plan.run_host(x, batch=batch)
        plan.close()
        want = np.concatenate([port.transpose(x[i * n:(i + 1) * n], dim, order)
                               for i in range(batch)])
        assert bits_equal(got, want), "%s %s %s [%s]: %s" % (dtype, dim, opt, desc,
                                                              first_mismatch(got, want))


def test_dimchg_vector_path(dev, port):
    for dtype, dim, f, t in (("float32", [4, 20, 36, 8], 0, 3), ("float32", [16, 12, 40, 1], 1, 2),
                             ("float64", [6, 10, 34, 2], 0, 2), ("float32", [32, 3, 50, 4], 0, 1)):
        n = int(np.prod(dim))
        x = rand_small(dtype, n, seed=n)
        plan = dev.plan("dimchg", "%d:%d" % (f, t), dtype, dim)
        desc = plan.describe()
        got = plan.run_host(x)
        plan.close()
        want = port.dimchg(x, dim, f, t)
        assert bits_equal(got, want), "%s %s %d:%d [%s] %s" % (dtype, dim, f, t, desc,
                                                                first_mismatch(got, want))


@pytest.mark.parametrize("from_to", [(0, 1), (0, 2), (0, 3), (0, 4), (1, 2), (1, 3), (2, 3),
                                     (2, 2), (0, 7)])
def test_dimchg(dev, port, from_to):
    f, t = from_to
    for dtype, dim in (("uint8", [3, 16, 16, 1]), ("uint8", [4, 16, 16, 1]),
                       ("float32", [3, 10, 10, 4, 5, 6]), ("int16", [2, 3, 4, 5, 2, 3, 2, 2])):
        n = int(np.prod(dim))
        x = rand_small(dtype, n, seed=60)
        plan = dev.plan("dimchg", "%d:%d" % (f, t), dtype, dim)
        got = plan.run_host(x)
        plan.close()
        want = port.dimchg(x, dim, f, t)
        assert bits_equal(got, want), "%s %s %d:%d %s" % (dtype, dim, f, t, first_mismatch(got, want))


def test_dimchg_from_gt_to_is_an_error(dev):
    from nnstreamer_b200 import NnsbError
    with pytest.raises(NnsbError) as e:
        dev.plan("dimchg", "2:0", "uint8", [3, 4, 5, 1])
    assert "not permitted" in str(e.value)


# ---------------------------------------------------------------------------------------
# clamp
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", TYPES)
def test_clamp_every_type(dev, port, dtype):
    n = 9001
    x = rand_full(dtype, n, seed=70)
    for lo, hi, opt in ((-50.0, 50.0, "-50:50"), (20.0, 80.0, "+20:80"), (-33.3, 77.7, "-33.3:+77.7"),
                        (0.0, 0.0, "0:0"), (-1e30, 1e30, "-1e30:1e30"), (1e10, 1e18, "1e10:1e18"),
                        (-0.5, 0.5, "-0.5:0.5"), (254.5, 300.0, "254.5:300"),
                        (-200.7, -100.2, "-200.7:-100.2"), (3e9, 4e9, "3e9:4e9"),
                        (-1e18, -1e10, "-1e18:-1e10"), (-32768.5, 65535.5, "-32768.5:65535.5"),
                        (127.0, 128.0, "127:128"), (-2147483648.0, 2147483647.0, "-2147483648:2147483647")):
        plan = dev.plan("clamp", opt, dtype, [n])
        got = plan.run_host(x)
        plan.close()
        want = port.clamp(x, lo, hi)
        assert bits_equal(got, want), "%s %s %s" % (dtype, opt, first_mismatch(got, want, x))


# ---------------------------------------------------------------------------------------
# stand
# ---------------------------------------------------------------------------------------
def _stand_close(got, want, scale):
    g = got.astype(np.float64)
    w = want.astype(np.float64)
    tol = 1e-6 * np.abs(w) + 1e-6 * scale
    if np.dtype(got.dtype).kind != "f":
        tol = tol + 1.0            # integer outputs: a boundary case may truncate differently
    bad = np.abs(g - w) > tol
    assert not bad.any(), "%d of %d beyond tolerance; worst %g" % (bad.sum(), bad.size,
                                                                   np.abs(g - w).max())


@pytest.mark.parametrize("mode", ["default", "dc-average"])
@pytest.mark.parametrize("per_channel", [False, True])
def test_stand_reference_test_shape(dev, port, mode, per_channel):
    """tests/transform_stand: f32 50:100:1:1, U(0,10)."""
    rng = np.random.default_rng(4)
    dim = [50, 100, 1, 1]
    x = (rng.random(5000) * 10).astype(np.float32)
    opt = mode + (",per-channel:true" if per_channel else "")
    plan = dev.plan("stand", opt, "float32", dim)
    got = plan.run_host(x)
    mean, std = plan.stand_stats(1, 50 if per_channel else 1)
    plan.close()
    want = port.stand(x, dim, mode, None, per_channel)
    _stand_close(got, want, 1.0)
    omean, ostd = port.stand_stats(x, dim, per_channel)
    np.testing.assert_allclose(mean, omean, rtol=1e-12, atol=1e-12)
    if mode == "default":
        np.testing.assert_allclose(std, ostd, rtol=1e-10)


@pytest.mark.parametrize("itype", TYPES)
def test_stand_every_input_type(dev, port, itype):
    dim = [3, 40, 30, 1]
    n = int(np.prod(dim))
    x = rand_small(itype, n, seed=80, lo=0, hi=100)
    for mode, otype, pc in (("default", None, False), ("dc-average", "float32", True),
                            ("default", "float64", True), ("dc-average", "int16", False)):
        opt = mode + ((":" + otype) if otype else "") + (",per-channel:true" if pc else "")
        plan = dev.plan("stand", opt, itype, dim)
        got = plan.run_host(x)
        plan.close()
        want = port.stand(x, dim, mode, otype, pc)
        _stand_close(got, want, 100.0 if mode == "dc-average" else 1.0)


@pytest.mark.parametrize("nch", [1, 2, 3, 4, 8, 9, 16, 100, 1000, 70000])
def test_stand_channel_counts_and_paths(dev, port, nch):
    """nch <= 8: vector path; larger: strided path; more channels than threads: column walk."""
    rows = 7 if nch >= 1000 else 301
    dim = [nch, rows, 1, 1]
    rng = np.random.default_rng(nch)
    x = (rng.standard_normal(nch * rows) * 3 + 5).astype(np.float32)
    plan = dev.plan("stand", "default,per-channel:true", "float32", dim)
    got = plan.run_host(x)
    plan.close()
    want = port.stand(x, dim, "default", None, True)
    _stand_close(got, want, 1.0)


def test_stand_batched_is_per_tensor(dev, port):
    """A batch of B tensors in one launch == B reference calls (statistics per tensor)."""
    dim = [3, 32, 24, 1]
    n = int(np.prod(dim))
    rng = np.random.default_rng(9)
    x = np.concatenate([(rng.random(n) * (i + 1) + 10 * i).astype(np.float32) for i in range(5)])
    for opt, mode, pc in (("default", "default", False), ("dc-average,per-channel:true", "dc-average", True)):
        plan = dev.plan("stand", opt, "float32", dim)
        got = plan.run_host(x, batch=5)
        assert plan.last_launch_count == 3
        plan.close()
        want = np.concatenate([port.stand(x[i * n:(i + 1) * n], dim, mode, None, pc) for i in range(5)])
        _stand_close(got, want, 50.0 if mode == "dc-average" else 1.0)


def test_stand_constant_tensor_uses_1e_minus_10(dev, port):
    dim = [4, 8, 8, 1]
    x = np.full(256, 3.25, dtype=np.float32)
    plan = dev.plan("stand", "default", "float32", dim)
    got = plan.run_host(x)
    mean, std = plan.stand_stats(1, 1)
    plan.close()
    assert std[0] == 1e-10 and mean[0] == 3.25
    assert bits_equal(got, port.stand(x, dim, "default"))


def test_stand_large_offset_small_variance(dev, port):
    """mean 1e6, std 1e-2: the shifted one-pass variance must not cancel catastrophically."""
    rng = np.random.default_rng(10)
    dim = [1, 1 << 16, 1, 1]
    x = (1e6 + rng.standard_normal(1 << 16) * 1e-2).astype(np.float64)
    plan = dev.plan("stand", "default", "float64", dim)
    got = plan.run_host(x)
    mean, std = plan.stand_stats(1, 1)
    plan.close()
    omean, ostd = port.stand_stats(x, dim, False)
    assert abs(mean[0] - omean[0]) <= 1e-9 * abs(omean[0])
    assert abs(std[0] - ostd[0]) <= 1e-9 * ostd[0]
    _stand_close(got, port.stand(x, dim, "default"), 1.0)


# ---------------------------------------------------------------------------------------
# converter stride strip
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("cwh", [(3, 642, 480), (1, 642, 480), (3, 1918, 32), (1, 5, 7), (3, 1, 9),
                                 (3, 3, 1), (1, 1918, 16), (3, 21, 2)])
def test_stride_strip(dev, port, cwh):
    c, w, h = cwh
    row = c * w
    pitch = (row + 3) // 4 * 4
    assert pitch != row
    rng = np.random.default_rng(w)
    src = rng.integers(0, 256, pitch * h, dtype=np.uint8)
    d_src = dev.upload(src)
    d_dst = dev.alloc(row * h)
    dev.strip(d_src, d_dst, row, pitch, h)
    got = dev.download(d_dst, row * h, np.uint8)
    dev.free(d_src)
    dev.free(d_dst)
    want = port.stride_strip(src, c, w, h)
    assert bits_equal(got, want), first_mismatch(got, want)
    # numpy cross-check of the oracle
    assert bits_equal(want, src.reshape(h, pitch)[:, :row].reshape(-1))


def test_stride_strip_unaligned_pointers(dev, port):
    c, w, h = 3, 642, 16
    row, pitch = c * w, (c * w + 3) // 4 * 4
    rng = np.random.default_rng(1)
    src = rng.integers(0, 256, pitch * h, dtype=np.uint8)
    d_src = dev.alloc(pitch * h + 16)
    d_dst = dev.alloc(row * h + 16)
    dev.h2d(d_src + 1, src)
    dev.strip(d_src + 1, d_dst + 3, row, pitch, h)
    got = dev.download(d_dst + 3, row * h, np.uint8)
    assert bits_equal(got, port.stride_strip(src, c, w, h))
    dev.free(d_src)
    dev.free(d_dst)


# ---------------------------------------------------------------------------------------
# against the committed outputs of the reference's own code (tests/golden/ref_golden.npz)
# ---------------------------------------------------------------------------------------
def test_gpu_matches_reference_golden_fixture(dev):
    import os
    from test_oracle_golden import ARITH_CASES
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_golden.npz"))
    for it in TYPES:
        x = gold["typecast/%s/in" % it]
        for ot in TYPES:
            plan = dev.plan("typecast", ot, it, [x.size])
            got = plan.run_host(x)
            plan.close()
            assert bits_equal(got, gold["typecast/%s/%s" % (it, ot)]), (it, ot)
    for name, it, dim, opt, ops, pc, cd in ARITH_CASES:
        plan = dev.plan("arithmetic", opt, it, dim)
        got = plan.run_host(gold["arith/%s/in" % name])
        plan.close()
        assert bits_equal(got, gold["arith/%s/out" % name]), name
    x = gold["transpose/in"]
    for order in ORDERS:
        plan = dev.plan("transpose", "%d:%d:%d:3" % order[:3], "float32", [20, 10, 3, 2])
        assert bits_equal(plan.run_host(x), gold["transpose/%d%d%d" % order[:3]])
        plan.close()
    for to in (1, 2, 3):
        plan = dev.plan("dimchg", "0:%d" % to, "uint8", [3, 16, 16, 1])
        assert bits_equal(plan.run_host(gold["dimchg/in"]), gold["dimchg/0_%d" % to])
        plan.close()
    for it, lo, hi, opt in (("int8", -50, 50, "-50:50"), ("uint32", 20, 80, "20:80"),
                            ("float32", -33.3, 77.7, "-33.3:77.7"), ("float64", -1e3, 1e3, "-1e3:1e3"),
                            ("int64", -1e18, 1e18, "-1e18:1e18")):
        x = gold["clamp/%s/in" % it]
        plan = dev.plan("clamp", opt, it, [x.size])
        assert bits_equal(plan.run_host(x), gold["clamp/%s/out" % it]), it
        plan.close()
    xs = gold["stand/in"]
    for mode in ("default", "dc-average"):
        for pc in (False, True):
            plan = dev.plan("stand", mode + (",per-channel:true" if pc else ""), "float32", [50, 100, 1, 1])
            _stand_close(plan.run_host(xs), gold["stand/%s/%d" % (mode, int(pc))], 10.0 if mode == "dc-average" else 1.0)
            plan.close()


# ---------------------------------------------------------------------------------------
# guard bands: no kernel may write outside its output (compute-sanitizer is closed on this pool)
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("case", [
    ("typecast", "float32", "uint8", [10007], 1),
    ("arithmetic", "typecast:float32,add:-127.5,div:127.5", "uint8", [3, 211, 157, 1], 2),
    ("arithmetic", "per-channel:true@0,mul:2@1", "int8", [3, 160, 144, 1], 1),
    ("clamp", "-5:5", "int16", [9001], 1),
    ("transpose", "1:2:0:3", "float32", [3, 1000, 77, 1], 1),      # TMA path, partial last tile
    ("transpose", "2:0:1:3", "uint8", [336, 100, 3, 1], 3),        # TMA path, planar -> interleaved
    ("transpose", "2:1:0:3", "float32", [33, 65, 17, 2], 1),       # box tiles sticking out
    ("transpose", "1:0:2:3", "uint16", [61, 37, 5, 1], 1),
    ("dimchg", "0:3", "float64", [3, 10, 10, 4, 5], 1),
    ("stand", "default,per-channel:true", "float32", [3, 61, 37, 1], 3),
    ("stand", "dc-average", "uint8", [1000, 7, 1, 1], 1),
], ids=lambda c: "%s-%s-%s" % (c[0], c[1], "x".join(str(d) for d in c[3])))
def test_no_writes_outside_the_output(dev, case):
    mode, option, itype, dim, batch = case
    GUARD = 512
    n = int(np.prod(dim)) * batch
    x = rand_small(itype, n, seed=99)
    plan = dev.plan(mode, option, itype, dim)
    out_bytes = plan.out_size * batch
    d_in = dev.upload(x)
    d_buf = dev.alloc(out_bytes + 2 * GUARD)
    fill = np.full(out_bytes + 2 * GUARD, 0xAB, dtype=np.uint8)
    dev.h2d(d_buf, fill)
    dev.sync()
    plan.launch(d_in, d_buf + GUARD, batch)
    got = dev.download(d_buf, out_bytes + 2 * GUARD, np.uint8)
    assert (got[:GUARD] == 0xAB).all(), "wrote before the output"
    assert (got[GUARD + out_bytes:] == 0xAB).all(), "wrote past the output"
    dev.free(d_in)
    dev.free(d_buf)
    plan.close()


def test_strip_does_not_write_outside_the_output(dev):
    for c, w, h in ((3, 642, 480), (1, 1918, 7), (3, 5, 3)):
        row, pitch = c * w, (c * w + 3) // 4 * 4
        rng = np.random.default_rng(w)
        src = rng.integers(0, 256, pitch * (h - 1) + row, dtype=np.uint8)     # last row unpadded
        d_src = dev.upload(src)
        d_buf = dev.alloc(row * h + 1024)
        dev.h2d(d_buf, np.full(row * h + 1024, 0xAB, dtype=np.uint8))
        dev.sync()
        dev.strip(d_src, d_buf + 512, row, pitch, h)
        got = dev.download(d_buf, row * h + 1024, np.uint8)
        assert (got[:512] == 0xAB).all() and (got[512 + row * h:] == 0xAB).all()
        want = np.concatenate([src[r * pitch:r * pitch + row] for r in range(h)])
        assert bits_equal(got[512:512 + row * h], want)
        dev.free(d_src)
        dev.free(d_buf)
