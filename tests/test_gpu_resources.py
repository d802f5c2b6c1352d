"""Resource hygiene on the GPU: elements, plans and their pools give everything back, and a
long-running stream of frames does not grow device memory (the pools recycle)."""
import numpy as np
import pytest

import nnstreamer_b200 as nb
from nnstreamer_b200 import (DirectVideoDecoder, TensorAggregator, TensorConverter, TensorMerge,
                             TensorSplit, TensorTransform)

pytestmark = pytest.mark.gpu


def free_bytes():
    import torch
    torch.cuda.synchronize()
    return torch.cuda.mem_get_info(0)[0]


def caps(dim, type_name):
    return "other/tensor, dimension=(string)%s, type=(string)%s, framerate=(fraction)30/1" % (
        ":".join(str(d) for d in dim), type_name)


def one_round(seed):
    rng = np.random.default_rng(seed)
    dim = [3, 320, 240, 1]
    n = int(np.prod(dim))
    x = rng.integers(0, 256, n, dtype=np.uint8)
    with nb.Device(0) as dev:
        for mode, opt, ty in (("typecast", "float32", "uint8"),
                              ("arithmetic", "typecast:float32,per-channel:true@0,add:-1@0,div:3@1", "uint8"),
                              ("transpose", "1:2:0:3", "uint8"), ("transpose", "2:1:0:3", "float32"),
                              ("dimchg", "0:2", "float32"), ("clamp", "10:200", "uint8"),
                              ("stand", "default,per-channel:true", "float32")):
            plan = dev.plan(mode, opt, ty, dim)
            plan.run_host(x.astype(ty), batch=1)
            plan.close()
    conv = TensorConverter(frames_per_tensor=2)
    conv.set_caps("video/x-raw, format=(string)RGB, width=(int)322, height=(int)240, framerate=(fraction)30/1")
    frame = rng.integers(0, 256, conv.frame_size, dtype=np.uint8)
    t = TensorTransform(mode="arithmetic", option="typecast:float32,add:-127.5,div:127.5")
    t.negotiate(caps([3, 322, 240, 2], "uint8"))
    agg = TensorAggregator(frames_out=3, frames_flush=1, frames_dim=1)
    agg.set_caps(caps([3, 322, 240, 2], "float32"))
    for _ in range(6):
        flow, tensor = conv.chain(frame)
        assert flow == 0
        if tensor is None:
            continue
        flow, (y,) = t.transform([tensor])
        assert flow == 0
        flow, outs, _ = agg.chain(y)
        assert flow == 0
        for m in outs:
            m.release()
        y.release()
        tensor.release()
    dv = DirectVideoDecoder()
    flow, video = dv.decode(dv.make_config(3, 322, 240), frame[:3 * 322 * 240])
    assert flow == 0
    video.release()
    mg = TensorMerge(mode="linear", option="0")
    mg.set_caps([caps([3, 64, 48, 1], "uint8"), caps([1, 64, 48, 1], "uint8")])
    flow, merged = mg.chain([x[:3 * 64 * 48], x[:64 * 48]])
    assert flow == 0
    sp = TensorSplit(tensorseg="1:64:48,3:64:48")
    sp.set_caps(caps([4, 64, 48, 1], "uint8"))
    flow, parts = sp.chain(merged)
    assert flow == 0 and len(parts) == 2
    for p in parts:
        p.release()
    merged.release()
    for e in (conv, t, agg, dv, mg, sp):
        e.close()


def test_create_use_destroy_gives_device_memory_back():
    one_round(0)                        # contexts, module loading, first-use allocations
    before = free_bytes()
    for i in range(10):
        one_round(i + 1)
    after = free_bytes()
    assert before - after < 64 << 20, "device memory shrank by %.1f MB over 10 rounds" % (
        (before - after) / 1e6)


def test_steady_state_streaming_does_not_allocate():
    """Thousands of frames through one element: free device memory stays flat (pooled blocks)."""
    t = TensorTransform(mode="arithmetic", option="typecast:float32,add:-127.5,div:127.5")
    dim = [3, 640, 480, 1]
    t.negotiate(caps(dim, "uint8"))
    x = np.random.default_rng(1).integers(0, 256, int(np.prod(dim)), dtype=np.uint8)
    for _ in range(5):
        flow, (y,) = t.transform([x])
        assert flow == 0
    before = free_bytes()
    for _ in range(2000):
        flow, (y,) = t.transform([x])
        assert flow == 0
    after = free_bytes()
    t.close()
    assert abs(before - after) < 8 << 20
