#!/usr/bin/env python3
"""Reads the NNSB_STAND_TRACE dump of the fused stand kernel (k_stand_fused.cu) and prints, per
tensor, when each role of the CTAs reached each point (microseconds, relative to the tensor's
first event; min / median / max over the CTAs).  Developer aid for profiles/."""
import sys

import numpy as np

NAMES = ["apply start (ready seen)", "apply end", "stats: last tile landed", "stats: warp 0 done",
         "control: stats_done seen", "control: published", "control: all CTAs arrived",
         "control: ready", "producer: first tile issued", "producer: last tile issued"]


def main():
    raw = np.fromfile(sys.argv[1], dtype=np.uint64)
    grid, batch, pts = (int(v) for v in raw[:3])
    t = raw[3:].reshape(grid, batch, pts).astype(np.int64)
    t0 = t[t > 0].min()
    frames = [int(a) for a in sys.argv[2:]] or [batch // 2, batch // 2 + 1]
    for f in frames:
        base = t[:, f, :][t[:, f, :] > 0].min()
        print("tensor %d (starts %.2f us after the kernel's first event)" % (f, (base - t0) / 1e3))
        order = sorted(range(pts), key=lambda p: np.median(t[:, f, p]))
        for p in order:
            v = (t[:, f, p] - base) / 1e3
            print("  %-30s min %7.2f  med %7.2f  max %7.2f" % (NAMES[p], v.min(), np.median(v), v.max()))
    per = np.diff(np.median(t[:, :, 7], axis=0)) / 1e3
    print("period between `ready` of consecutive tensors: median %.2f us (min %.2f, max %.2f)" % (
        np.median(per), per.min(), per.max()))


if __name__ == "__main__":
    main()
