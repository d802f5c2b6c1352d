#!/usr/bin/env python3
"""Sweeps the slab-permute knobs (NNSB_SLAB_KB, NNSB_SLAB_CTAS, NNSB_SLAB_STORE, NNSB_PERM_SLAB) on the narrow-row
video shapes of bench.py's extras; every combination runs in a child process (the knobs are read
when a plan is created), inputs rotate over a set larger than the L2.

usage: python profiles/sweep_perm.py            (prints one line per combination)
       python profiles/sweep_perm.py --one MODE OPTION TYPE DIM BATCH   (child)
"""
import itertools
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

KBS = (16, 24, 32, 48, 64)
CASES = [
    ("transpose", "1:2:0:3", "uint16", "3:225:113:1", 256),
    ("transpose", "0:2:1:3", "uint16", "3:225:113:1", 256),
    ("transpose", "2:1:0:3", "float32", "3:227:227:1", 64),
    ("transpose", "0:2:1:3", "uint8", "3:1920:1080:1", 32),
]


def one(mode, option, in_type, dim_s, batch):
    from nnstreamer_b200 import Device
    dim = [int(d) for d in dim_s.split(":")]
    dt = np.dtype(in_type)
    n = int(np.prod(dim)) * batch
    nbytes = n * dt.itemsize
    sets = max(3, int(2.5 * 128e6 / (2 * nbytes)) + 1)
    rng = np.random.default_rng(0)
    x = rng.integers(0, 200, n).astype(dt)
    with Device(0) as dev:
        plan = dev.plan(mode, option, in_type, dim)
        ins = [dev.upload(x) for _ in range(sets)]
        outs = [dev.alloc(nbytes) for _ in range(sets)]
        e0, e1 = dev.event(), dev.event()
        for i in range(sets):
            plan.launch(ins[i], outs[i], batch)
        dev.sync()
        reps = 4
        best, tot = 1e9, 0.0
        for _ in range(reps):
            dev.record(e0)
            for i in range(sets):
                plan.launch(ins[i], outs[i], batch)
            dev.record(e1)
            ms = dev.elapsed_ms(e0, e1) / sets
            best = min(best, ms)
            tot += ms
        print("%.2f us mean, %.2f us best, %.0f GB/s mean | %s" % (
            1e3 * tot / reps, 1e3 * best, 2 * nbytes / (tot / reps) / 1e6, plan.describe()[-60:]))
        plan.close()


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--one":
        return one(sys.argv[2], sys.argv[3], sys.argv[4], sys.argv[5], int(sys.argv[6]))
    quick = "--quick" in sys.argv
    for case in (CASES[:3] if quick else CASES):
        print("==", case, flush=True)
        combos = [({}, "default"), ({"NNSB_PERM_SLAB": "0", "NNSB_PERM_NO_PIX": "1"}, "vector kernel")]
        if case[3].startswith("3:1920:1080") and case[2] == "uint8":
            for stage, swz in (("1", "0"), ("0", "0")):
                combos.append(({"NNSB_PIX_STAGE": stage, "NNSB_PIX_SWIZZLE": swz},
                               "pixel stage=%s swizzle=%s" % (stage, swz)))
        stores = ("lanes",)
        combos.append(({"NNSB_SLAB_NO_FOLD": "1"}, "default, no pixel fold"))
        for store, kb, ctas in itertools.product(stores, (24, 32, 48) if quick else KBS, (4, 6)):
            if (kb + 1) * ctas > 220:
                continue
            combos.append(({"NNSB_PERM_SLAB": "2", "NNSB_PERM_NO_PIX": "1", "NNSB_SLAB_KB": str(kb),
                            "NNSB_SLAB_CTAS": str(ctas), "NNSB_SLAB_STORE": store},
                           "%-6s %2d KB x %d CTAs" % (store, kb, ctas)))
        for env, label in combos:
            e = dict(os.environ)
            e.update(env)
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--one"] + [str(c) for c in case],
                               env=e, capture_output=True, text=True, timeout=120)
            line = r.stdout.strip().splitlines()[-1] if r.returncode == 0 and r.stdout.strip() else \
                "FAILED " + r.stderr.strip()[-200:]
            print("  %-22s %s" % (label, line), flush=True)


if __name__ == "__main__":
    main()
