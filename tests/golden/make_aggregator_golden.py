#!/usr/bin/env python3
"""Extracts the reference's own tensor_aggregator known-answer vectors into
tests/golden/aggregator_golden.npz.

Source: /root/reference/tests/nnstreamer_plugins/unittest_plugins.cc, TEST (testTensorAggregator,
aggregate1 .. aggregate10) -- the two input frames `aggr_test_frames` (:2145-2153), the
properties each test sets on the element, the sink dimension string and its `expected[96]`
array.  Run in the development container (needs /root/reference):
    python tests/golden/make_aggregator_golden.py
The reference cannot travel to the GPU box, so the vectors are committed as a fixture; the
oracle (CPU) and the device aggregator (GPU) are both checked against them.
"""
import os
import re

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/tests/nnstreamer_plugins/unittest_plugins.cc"


def ints(text):
    return np.array([int(v) for v in re.findall(r"-?\d+", text)], dtype=np.int32)


def main():
    src = open(SRC).read()
    out = {}
    m = re.search(r"const gint aggr_test_frames\[2\]\[48\]\s*=\s*\{(.*?)\};", src, re.S)
    frames = ints(m.group(1)).reshape(2, 48)
    out["frames"] = frames
    names = []
    for i in range(1, 11):
        m = re.search(r"TEST \(testTensorAggregator, aggregate%d\)\s*\{(.*?)\n\}" % i, src, re.S)
        body = m.group(1)
        props = {"frames-in": 1, "frames-out": 1, "frames-flush": 0, "frames-dim": 7, "concat": 1}
        for pm in re.finditer(r'g_object_set \(h->element,(.*?)NULL\);', body, re.S):
            for key, val in re.findall(r'"([a-z-]+)",\s*(\(gboolean\)\s*\w+|\w+)', pm.group(1)):
                val = val.replace("(gboolean)", "").strip()
                props[key] = {"TRUE": 1, "FALSE": 0}.get(val, None)
                if props[key] is None:
                    props[key] = int(val)
        dim = re.search(r'gst_tensor_parse_dimension \("([0-9:]+)"', body).group(1)
        em = re.search(r"const gint expected\[\d+\]\s*=\s*\{(.*?)\};", body, re.S)
        name = "aggregate%d" % i
        names.append(name)
        out[name + "/props"] = np.array([props["frames-in"], props["frames-out"],
                                         props["frames-flush"], props["frames-dim"],
                                         props["concat"]], dtype=np.int64)
        out[name + "/dim"] = np.array([int(d) for d in dim.split(":")], dtype=np.int64)
        if em:
            out[name + "/expected"] = ints(em.group(1))
        else:
            # aggregate5 (concat=false, frames-in == frames-out): every buffer comes back as it is
            assert "_aggregator_test_check_output (h, aggr_test_frames[i], 48)" in body
            out[name + "/expected"] = frames.reshape(-1)
            out[name + "/per_buffer"] = np.array([1], dtype=np.int64)
    # flushData (:2719-2760): 10 frames in, 6 out, flush 6, flush-stop between two pushes
    out["flush/data"] = np.arange(1, 11, dtype=np.int32)
    out["flush/props"] = np.array([10, 6, 6, 0, 1], dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "aggregator_golden.npz"), **out)
    for n in names:
        print(n, out[n + "/props"], out[n + "/dim"], out[n + "/expected"].size)


if __name__ == "__main__":
    main()
