"""CPU: the oracle port against (a) the committed outputs of the reference's own code
(tests/golden/ref_golden.npz, made by tests/golden/make_golden.py) and (b) the known-answer
vectors of the reference's unit tests (tests/nnstreamer_plugins/unittest_plugins.cc)."""
import os

import numpy as np
import pytest

from helpers import TYPES, bits_equal, first_mismatch

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_golden.npz"))

import importlib.util
_spec = importlib.util.spec_from_file_location(
    "make_golden_cases", os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "make_golden.py"))


def _arith_cases():
    # the case table lives in the generator; read it without running the generator
    src = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "make_golden.py")).read()
    start = src.index("ARITH_CASES = [")
    end = src.index("]\n\n\ndef main")
    ns = {}
    exec(src[start:end + 1], ns)
    return ns["ARITH_CASES"]


ARITH_CASES = _arith_cases()


@pytest.mark.parametrize("itype", TYPES)
def test_port_typecast_matches_reference_golden(port, itype):
    x = GOLD["typecast/%s/in" % itype]
    for otype in TYPES:
        got = port.typecast(x, otype)
        want = GOLD["typecast/%s/%s" % (itype, otype)]
        assert bits_equal(got, want), "%s->%s %s" % (itype, otype, first_mismatch(got, want, x))


@pytest.mark.parametrize("case", ARITH_CASES, ids=[c[0] for c in ARITH_CASES])
def test_port_arithmetic_matches_reference_golden(port, case):
    name, it, dim, opt, ops, pc, cd = case
    x = GOLD["arith/%s/in" % name]
    got = port.arithmetic(x, dim, ops, pc, cd)
    want = GOLD["arith/%s/out" % name]
    assert bits_equal(got, want), first_mismatch(got, want, x)


def test_port_transpose_dimchg_clamp_match_reference_golden(port):
    x = GOLD["transpose/in"]
    for order in ((0, 1, 2, 3), (0, 2, 1, 3), (1, 0, 2, 3), (1, 2, 0, 3), (2, 0, 1, 3), (2, 1, 0, 3)):
        assert bits_equal(port.transpose(x, [20, 10, 3, 2], order), GOLD["transpose/%d%d%d" % order[:3]])
    xb = GOLD["dimchg/in"]
    for to in (1, 2, 3):
        assert bits_equal(port.dimchg(xb, [3, 16, 16, 1], 0, to), GOLD["dimchg/0_%d" % to])
    for it, lo, hi in (("int8", -50, 50), ("uint32", 20, 80), ("float32", -33.3, 77.7),
                       ("float64", -1e3, 1e3), ("int64", -1e18, 1e18)):
        assert bits_equal(port.clamp(GOLD["clamp/%s/in" % it], float(lo), float(hi)),
                          GOLD["clamp/%s/out" % it]), it


def test_port_stand_matches_reference_golden_bit_for_bit(port):
    """The port follows the reference's summation order, so even stand is bit-identical."""
    xs = GOLD["stand/in"]
    for mode in ("default", "dc-average"):
        for pc in (False, True):
            assert bits_equal(port.stand(xs, [50, 100, 1, 1], mode, None, pc),
                              GOLD["stand/%s/%d" % (mode, int(pc))])
    assert bits_equal(port.stand(GOLD["stand_u8/in"], [3, 40, 30, 1], "default", "float32", True),
                      GOLD["stand_u8/default_f32"])


def test_golden_agrees_with_numpy_definitions():
    """Sanity of the fixture itself against the generators the reference's SSAT tests use:
    np.transpose (transform_transpose/generateTest.py), np.clip (transform_clamp)."""
    x = GOLD["transpose/in"].reshape(2, 3, 10, 20)      # [d3][d2][d1][d0]
    # option 1:2:0:3 -> out dims (10, 3, 20, 2): out[l][c][h][w] with new d0 = old d1 ...
    want = np.transpose(x, (0, 3, 1, 2)).reshape(-1)    # new order (d3, d0, d2, d1)
    assert bits_equal(GOLD["transpose/120"], want)
    xc = GOLD["clamp/float32/in"]
    finite = np.isfinite(xc)
    assert bits_equal(GOLD["clamp/float32/out"][finite],
                      np.clip(xc[finite].astype(np.float64), -33.3, 77.7).astype(np.float32))
    xs = GOLD["stand/in"].astype(np.float64)
    want = np.abs((xs - xs.mean()) / (xs.std() + 1e-10))
    np.testing.assert_allclose(GOLD["stand/default/0"], want, atol=1e-5)     # checkResult.py bound


# ---- known-answer vectors of unittest_plugins.cc ---------------------------------------
TYPECAST_PAIRS = [  # unittest_plugins.cc:955-1118 (typecast_1 .. typecast_14, each also _accel)
    ("uint8", "uint32"), ("uint32", "float64"), ("int32", "float32"), ("int8", "float32"),
    ("uint8", "float32"), ("int16", "float32"), ("uint16", "float32"), ("uint64", "int64"),
    ("float32", "uint32"), ("uint8", "int8"), ("uint32", "int16"), ("float32", "uint8"),
    ("float64", "uint16"), ("float64", "uint64")]


@pytest.mark.parametrize("accel", [False, True])
@pytest.mark.parametrize("pair", TYPECAST_PAIRS, ids=["%s_%s" % p for p in TYPECAST_PAIRS])
def test_reference_unittest_typecast_vectors(port, pair, accel):
    it, ot = pair
    for b in range(3):
        x = np.array([(i + 1) * (b + 1) for i in range(5)], dtype=it)
        want = np.array([(i + 1) * (b + 1) for i in range(5)], dtype=ot)
        assert bits_equal(port.typecast(x, ot, accel=accel), want)


@pytest.mark.parametrize("accel", [False, True])
def test_reference_unittest_arithmetic_vectors(port, accel):
    ulp = lambda a, b, dt: np.abs(a.view(dt).astype(np.int64) - b.view(dt).astype(np.int64)).max()
    for b in range(3):
        i = np.arange(5)
        # arithmetic1 (:1123): f32 add:.5, EXPECT_FLOAT_EQ (4 ulp)
        x = ((i + 1) * (b + 1) + .2).astype(np.float32)
        got = port.arithmetic(x, [5], [("add", .5)], accel=accel)
        assert ulp(got, ((i + 1) * (b + 1) + .2 + .5).astype(np.float32), np.int32) <= 4
        # arithmetic2 (:1267): f64 mul:.5
        xd = ((i + 1) * (b + 1) + .2).astype(np.float64)
        got = port.arithmetic(xd, [5], [("mul", .5)], accel=accel)
        assert ulp(got, (((i + 1) * (b + 1) + .2) * .5).astype(np.float64), np.int64) <= 4
        # arithmetic3 (:1411): u8 typecast:float32,add:.5,mul:0.2
        xu = ((i + 1) * (b + 1)).astype(np.uint8)
        got = port.arithmetic(xu, [5], [("typecast", "float32"), ("add", .5), ("mul", 0.2)], accel=accel)
        assert ulp(got, (((i + 1) * (b + 1) + .5) * .2).astype(np.float32), np.int32) <= 4
        # arithmetic4 (:1563): typecast:float64,add:0.2,add:0.1 (trailing typecast stripped)
        got = port.arithmetic(xu, [5], [("typecast", "float64"), ("add", 0.2), ("add", 0.1)], accel=accel)
        assert got.dtype == np.float64
        assert ulp(got, ((i + 1) * (b + 1) + .3).astype(np.float64), np.int64) <= 4
        # arithmetic5 (:1717): typecast:int32,mul:2,div:2,add:-1, EXPECT_EQ
        got = port.arithmetic(xu, [5], [("typecast", "int32"), ("mul", 2), ("div", 2), ("add", -1)], accel=accel)
        assert bits_equal(got, ((i + 1) * (b + 1) - 1).astype(np.int32))
    # arithmeticPerChannel (:1869): "1 2 3 4 5 -> 0 2 6 12 20", dim 5, channel = dim 0
    ops = [("typecast", "int32")] + [("mul", k + 1, k) for k in range(5)] + [("add", -(k + 1), k) for k in range(5)]
    for b in range(3):
        xu = np.array([(i + 1) * (b + 1) for i in range(5)], dtype=np.uint8)
        got = port.arithmetic(xu, [5], ops, True, 0, accel=False)
        want = np.array([((i + 1) * (b + 1)) * (i + 1) - (i + 1) for i in range(5)], dtype=np.int32)
        assert bits_equal(got, want)


def test_stride_strip_matches_reference_bmp_layout(port):
    """nnstreamer_converter/runTest.sh:57-59: 642x480 RGB / GRAY8 -- rows padded to 4 bytes."""
    rng = np.random.default_rng(5)
    for c, w, h in ((3, 642, 480), (1, 642, 480)):
        row, pitch = c * w, (c * w + 3) // 4 * 4
        dense = rng.integers(0, 256, (h, row), dtype=np.uint8)
        padded = np.zeros((h, pitch), dtype=np.uint8)
        padded[:, :row] = dense
        padded[:, row:] = 0xEE
        assert bits_equal(port.stride_strip(padded.reshape(-1), c, w, h), dense.reshape(-1))
