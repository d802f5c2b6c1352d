"""Two GPUs driven from ONE process: kernels that opt into more than 48 KB of dynamic shared memory
set that attribute per device (it belongs to the device's context), so the second device must
work as the first does.  Skipped on single-GPU boxes."""
import ctypes as C

import numpy as np
import pytest

from helpers import bits_equal
from nnstreamer_b200 import Device
from nnstreamer_b200 import _lib as L

pytestmark = pytest.mark.gpu


def n_devices():
    n = C.c_int(0)
    L.check(L.lib().nnsb_device_count(C.byref(n)))
    return n.value


def np_transpose(x, dim, order):
    a = x.reshape(list(reversed(dim[:3])))
    axes = tuple(2 - order[2 - k] for k in range(3))
    return a.transpose(axes).reshape(-1)


def test_large_shared_memory_kernels_on_a_second_device(dev):
    if n_devices() < 2:
        pytest.skip("one GPU")
    cases = [("float32", [3, 640, 360, 1], (1, 2, 0)),       # plane kernel
             ("float32", [4, 256, 130, 1], (2, 1, 0)),       # vector kernel
             ("uint16", [3, 225, 113, 1], (1, 2, 0)),        # slab kernel
             ("float32", [3, 227, 61, 1], (2, 0, 1))]        # narrow rows
    rng = np.random.default_rng(0)
    with Device(1) as dev1:
        for d in (dev, dev1, dev):                          # first device, second, first again
            for dtype, dim, order in cases:
                x = rng.integers(0, 1000, int(np.prod(dim))).astype(dtype)
                plan = d.plan("transpose", "%d:%d:%d:3" % order, dtype, dim)
                got = plan.run_host(x)
                plan.close()
                assert bits_equal(got, np_transpose(x, dim, order)), (dtype, dim, order)
        # stand's fused kernel (cooperative, > 200 KB of shared memory) on the second device
        dim = [3, 1920, 1080, 1]
        x = rng.standard_normal(int(np.prod(dim)) * 2, dtype=np.float32)
        plan = dev1.plan("stand", "default", "float32", dim)
        got = plan.run_host(x, batch=2).view(np.float32)
        plan.close()
        for b in range(2):
            f = x[b * x.size // 2:(b + 1) * x.size // 2].astype(np.float64)
            want = np.abs((f - f.mean()) / (f.std() + 1e-10))
            assert np.allclose(got[b * x.size // 2:(b + 1) * x.size // 2], want, rtol=1e-5, atol=1e-6)
