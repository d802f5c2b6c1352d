#!/usr/bin/env python3
"""Runs one plan a few times on synthetic data -- the command ncu wraps for the per-kernel
captures under profiles/ (bench.py is the number; this is only the profiling harness).

usage: python profiles/run_kernel.py MODE OPTION IN_TYPE DIM [--batch B] [--iters N]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nnstreamer_b200 import Device  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("mode")
    ap.add_argument("option")
    ap.add_argument("in_type")
    ap.add_argument("dim")
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--iters", type=int, default=3)
    a = ap.parse_args()
    dim = [int(d) for d in a.dim.split(":")]
    n = int(np.prod(dim)) * a.batch
    rng = np.random.default_rng(0)
    dt = np.dtype(a.in_type)
    if dt.kind == "f":
        x = (rng.random(n, dtype=np.float32) * 10).astype(dt)
    else:
        info = np.iinfo(dt)
        x = rng.integers(max(info.min, -1000), min(info.max, 1000), n, endpoint=True).astype(dt)
    with Device(0) as dev:
        plan = dev.plan(a.mode, a.option, a.in_type, dim)
        print("plan:", plan.describe())
        d_in = dev.upload(x)
        d_out = dev.alloc(plan.out_size * a.batch)
        e0, e1 = dev.event(), dev.event()
        for it in range(a.iters):
            dev.record(e0)
            plan.launch(d_in, d_out, a.batch)
            dev.record(e1)
            ms = dev.elapsed_ms(e0, e1)
            es_in, es_out = dt.itemsize, np.dtype(plan.out_type).itemsize
            print("iter %d: %.4f ms, %d launches, %.1f GB/s algorithmic" % (
                it, ms, plan.last_launch_count, n * (es_in + es_out) / ms / 1e6))
        plan.close()


if __name__ == "__main__":
    main()
