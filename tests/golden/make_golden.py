#!/usr/bin/env python3
"""Generates tests/golden/ref_golden.npz from the REFERENCE's own code.

Run in the development container (needs /root/reference):  python tests/golden/make_golden.py
The reference cannot travel to the GPU box, so its outputs on seeded inputs are committed
here as a small fixture; tests compare the oracle port (CPU) and the CUDA kernels (GPU) with
them bit for bit (stand: with tolerance).  Inputs are regenerated from the seeds below, the
fixture stores inputs too so nothing depends on numpy's generator staying stable.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from helpers import TYPES, rand_full, rand_small  # noqa: E402
from oracle import oracle as O  # noqa: E402

ARITH_CASES = [
    # (name, in type, dim, option string, oracle ops, per_channel, ch_dim)
    ("mobilenet", "uint8", [3, 16, 16, 1], "typecast:float32,add:-127.5,div:127.5",
     [("typecast", "float32"), ("add", -127.5), ("div", 127.5)], False, 0),
    ("imagenet", "uint8", [3, 16, 8, 2],
     "typecast:float32,per-channel:true@0,add:-123.675@0,add:-116.28@1,add:-103.53@2,"
     "div:58.395@0,div:57.12@1,div:57.375@2",
     [("typecast", "float32"), ("add", -123.675, 0), ("add", -116.28, 1), ("add", -103.53, 2),
      ("div", 58.395, 0), ("div", 57.12, 1), ("div", 57.375, 2)], True, 0),
    ("f32_add", "float32", [50, 4], "add:-10", [("add", -10)], False, 0),
    ("f32_mul", "float32", [50, 4], "mul:-5.5", [("mul", -5.5)], False, 0),
    ("f64_chain", "float64", [50, 4], "add:9.9e-1,mul:2.0", [("add", 9.9e-1), ("mul", 2.0)], False, 0),
    ("u8_i32", "uint8", [5, 40], "typecast:int32,mul:2,div:2,add:-1",
     [("typecast", "int32"), ("mul", 2), ("div", 2), ("add", -1)], False, 0),
    ("i16_wrap", "int16", [200], "mul:300,add:-7,div:3", [("mul", 300), ("add", -7), ("div", 3)], False, 0),
    ("u8_neg", "uint8", [200], "add:-10,mul:0.5", [("add", -10), ("mul", 0.5)], False, 0),
    ("perch_dim1", "float32", [3, 5, 4, 2], "per-channel:true@1,mul:2@0,add:1.5@4,div:3@2",
     [("mul", 2, 0), ("add", 1.5, 4), ("div", 3, 2)], True, 1),
    ("f16_chain", "float16", [120], "mul:1.5,add:3", [("mul", 1.5), ("add", 3)], False, 0),
    ("i64_div", "int64", [120], "div:-7,mul:3", [("div", -7), ("mul", 3)], False, 0),
]


def main():
    ref = O.ref()
    out = {}
    # typecast: every pair, full-range inputs incl. out-of-range floats
    for it in TYPES:
        x = rand_full(it, 257, seed=101)
        out["typecast/%s/in" % it] = x
        for ot in TYPES:
            out["typecast/%s/%s" % (it, ot)] = ref.typecast(x, ot)
    for name, it, dim, opt, ops, pc, cd in ARITH_CASES:
        n = int(np.prod(dim))
        x = rand_full(it, n, seed=102) if it not in ("float32", "float64", "float16") \
            else rand_small(it, n, seed=102)
        out["arith/%s/in" % name] = x
        out["arith/%s/out" % name] = ref.arithmetic(x, dim, ops, pc, cd)
    # transpose (the reference SSAT shapes, scaled down) and dimchg from=0
    x = rand_small("float32", 20 * 10 * 3 * 2, seed=103)
    out["transpose/in"] = x
    for order in ((0, 1, 2, 3), (0, 2, 1, 3), (1, 0, 2, 3), (1, 2, 0, 3), (2, 0, 1, 3), (2, 1, 0, 3)):
        out["transpose/%d%d%d" % order[:3]] = ref.transpose(x, [20, 10, 3, 2], order)
    xb = rand_full("uint8", 3 * 16 * 16, seed=104)
    out["dimchg/in"] = xb
    for to in (1, 2, 3):
        out["dimchg/0_%d" % to] = ref.dimchg(xb, [3, 16, 16, 1], 0, to)
    # clamp (reference SSAT cases)
    for it, lo, hi in (("int8", -50, 50), ("uint32", 20, 80), ("float32", -33.3, 77.7),
                       ("float64", -1e3, 1e3), ("int64", -1e18, 1e18)):
        x = rand_full(it, 300, seed=105)
        out["clamp/%s/in" % it] = x
        out["clamp/%s/out" % it] = ref.clamp(x, float(lo), float(hi))
    # stand (reference SSAT: f32 50:100:1:1 U(0,10))
    rng = np.random.default_rng(106)
    xs = (rng.random(5000) * 10).astype(np.float32)
    out["stand/in"] = xs
    for mode in ("default", "dc-average"):
        for pc in (False, True):
            out["stand/%s/%d" % (mode, int(pc))] = ref.stand(xs, [50, 100, 1, 1], mode, None, pc)
    xu = rand_full("uint8", 3 * 40 * 30, seed=107)
    out["stand_u8/in"] = xu
    out["stand_u8/default_f32"] = ref.stand(xu, [3, 40, 30, 1], "default", "float32", True)
    np.savez_compressed(os.path.join(HERE, "ref_golden.npz"), **out)
    print("wrote %d arrays" % len(out))


if __name__ == "__main__":
    main()
