#!/usr/bin/env python3
"""Launches the kernels that have no plan -- the converter's stride strip (repitch_kernel) and
tensor_merge's small-block kernel (merge_phase_kernel) -- a few times on the bench's shapes: the
command ncu wraps for their captures (profiles/r2_profile_commands.sh)."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nnstreamer_b200 import Device  # noqa: E402
from nnstreamer_b200 import _lib as L  # noqa: E402


def main():
    rng = np.random.default_rng(0)
    with Device(0) as dev:
        lib = L.lib()
        w, h, frames = 1918, 1080, 32
        row, pitch, rows = 3 * w, (3 * w + 3) // 4 * 4, h * frames
        d_src = dev.upload(rng.integers(0, 256, pitch * rows, dtype=np.uint8))
        d_dst = dev.alloc(row * rows)
        px = 1920 * 1080 * frames
        d_rgb = dev.upload(rng.integers(0, 256, 3 * px, dtype=np.uint8))
        d_a = dev.upload(rng.integers(0, 256, px, dtype=np.uint8))
        d_out = dev.alloc(4 * px)
        ptrs = (C.c_void_p * 2)(d_rgb, d_a)
        blk = (C.c_uint64 * 2)(3, 1)
        e0, e1 = dev.event(), dev.event()
        for it in range(3):
            dev.record(e0)
            dev.strip(d_src, d_dst, row, pitch, rows)
            dev.record(e1)
            ms = dev.elapsed_ms(e0, e1)
            print("strip iter %d: %.4f ms, %.1f GB/s" % (it, ms, 2 * row * rows / ms / 1e6))
            dev.record(e0)
            L.check(lib.nnsb_merge_launch(dev.stream, 2, ptrs, blk, C.c_void_p(d_out), px))
            dev.record(e1)
            ms = dev.elapsed_ms(e0, e1)
            print("merge iter %d: %.4f ms, %.1f GB/s" % (it, ms, 2 * 4 * px / ms / 1e6))


if __name__ == "__main__":
    main()
