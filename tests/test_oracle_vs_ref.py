"""CPU: pins the oracle port against the reference's OWN code compiled here
(oracle/_ref/libnns_ref.so, built by oracle/build_ref.py from /root/reference).
Skipped on machines where that library has not been built."""
import numpy as np
import pytest

from helpers import TYPES, bits_equal, first_mismatch, rand_full, rand_small


@pytest.mark.parametrize("itype", TYPES)
def test_typecast_every_pair_full_range(port, ref, itype):
    x = rand_full(itype, 3001, seed=201)
    for otype in TYPES:
        a, b = port.typecast(x, otype), ref.typecast(x, otype)
        assert bits_equal(a, b), "%s->%s %s" % (itype, otype, first_mismatch(a, b, x))


@pytest.mark.parametrize("dtype", TYPES)
def test_arithmetic_chains(port, ref, dtype):
    x = rand_full(dtype, 3 * 10 * 20, seed=202)
    dim = [3, 10, 20]
    for ops in ([("mul", 2)], [("add", -10), ("mul", 3), ("div", 7)], [("mul", -5.5)],
                [("typecast", "int32"), ("mul", 2), ("div", 2), ("add", -1)],
                [("div", 0), ("add", 1)], [("div", 0.0), ("mul", 3)],
                [("typecast", "float32"), ("add", -127.5), ("div", 127.5)],
                [("typecast", "uint8"), ("add", 300)], [("typecast", "float16"), ("mul", 1.5)],
                [("div", -1)]):
        if ops == [("div", -1)] and dtype in ("int32", "int64"):
            continue    # INT_MIN / -1 raises SIGFPE in the reference (x86 idiv); nothing to pin
        a, b = port.arithmetic(x, dim, ops), ref.arithmetic(x, dim, ops)
        assert bits_equal(a, b), "%s %s %s" % (dtype, ops, first_mismatch(a, b, x))


@pytest.mark.parametrize("ch_dim", [0, 1, 2, 3])
def test_arithmetic_per_channel(port, ref, ch_dim):
    dim = [3, 5, 4, 2]
    x = rand_small("float32", 120, seed=203)
    nch = dim[ch_dim]
    ops = [("add", 1.5, 0), ("mul", 2, nch - 1), ("div", 3.0), ("add", -7, min(1, nch - 1)), ("mul", 0.25, 99)]
    assert bits_equal(port.arithmetic(x, dim, ops, True, ch_dim), ref.arithmetic(x, dim, ops, True, ch_dim))


def test_transpose_dimchg_clamp_stand(port, ref):
    x = rand_small("float32", 100 * 50 * 3, seed=204)
    for order in ((0, 1, 2, 3), (0, 2, 1, 3), (1, 0, 2, 3), (1, 2, 0, 3), (2, 0, 1, 3), (2, 1, 0, 3)):
        assert bits_equal(port.transpose(x, [100, 50, 3, 1], order), ref.transpose(x, [100, 50, 3, 1], order))
    for dtype, dim in (("uint8", [3, 16, 16, 1]), ("float32", [3, 10, 10, 4, 5, 6])):
        xd = rand_small(dtype, int(np.prod(dim)), seed=205)
        for to in (1, 2, 4):
            assert bits_equal(port.dimchg(xd, dim, 0, to), ref.dimchg(xd, dim, 0, to))
    for t in TYPES:
        xc = rand_full(t, 1000, seed=206)
        assert bits_equal(port.clamp(xc, -33.3, 77.7), ref.clamp(xc, -33.3, 77.7)), t
    for t in TYPES:
        xs = rand_small(t, 3 * 40 * 30, seed=207, lo=0, hi=100)
        for mode, ot, pc in (("default", None, False), ("dc-average", "float32", True),
                             ("default", "float64", True), ("dc-average", "int16", False)):
            a = port.stand(xs, [3, 40, 30, 1], mode, ot, pc)
            b = ref.stand(xs, [3, 40, 30, 1], mode, ot, pc)
            assert bits_equal(a, b), (t, mode, ot, pc)


def test_convert_dimension(port, ref):
    from oracle import oracle as O
    dim = [3, 10, 10, 4, 5, 6, 1, 1]
    for f in range(8):
        for t in range(8):
            want = ref.convert_dimension("dimchg", dim, True, from_=f, to=t)
            assert O.convert_dims("dimchg", dim, from_=f, to=t) == want, (f, t)
    for order in ((0, 1, 2, 3), (0, 2, 1, 3), (1, 0, 2, 3), (1, 2, 0, 3), (2, 0, 1, 3), (2, 1, 0, 3)):
        assert O.convert_dims("transpose", dim, order=order) == \
            ref.convert_dimension("transpose", dim, True, order=order)
