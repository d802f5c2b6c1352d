"""The slab permute kernel (k_perm_slab.cu): rows that are not multiples of 16 bytes, frames at
odd byte offsets in a batch.  Every order x element size against numpy's transpose (the
reference's nested loops, gsttensor_transform.c:1417-1500, are a plain index permutation), with
the kernel forced wherever it fits (NNSB_PERM_SLAB=2) and in its default role, batches included."""
import itertools

import numpy as np
import pytest

from helpers import bits_equal

pytestmark = pytest.mark.gpu

ORDERS = [o for o in itertools.permutations((0, 1, 2)) if o != (0, 1, 2)]
SHAPES = [("uint16", [3, 225, 113, 1], 5), ("float32", [3, 227, 227, 1], 3), ("uint8", [3, 321, 77, 1], 4),
          ("float64", [5, 33, 117, 1], 3), ("uint8", [1, 1001, 333, 1], 2), ("int16", [7, 64, 129, 1], 6),
          # small frames: a slab is a stack of whole frames and all three dims sit inside a run
          ("uint16", [6, 10, 14, 1], 40), ("float32", [3, 20, 9, 1], 50), ("uint8", [5, 12, 7, 1], 333)]
STORES = ["lanes", "chunks", None]


def np_transpose(x, dim, order, batch):
    """out dim i = in dim order[i] (innermost first); numpy axes are the reversed dims."""
    a = x.reshape([batch] + list(reversed(dim[:3])))
    axes = (0,) + tuple(1 + (2 - order[2 - k]) for k in range(3))
    return a.transpose(axes).reshape(-1)


@pytest.mark.parametrize("store", STORES + ["fold", "window"], ids=["lanes", "chunks", "auto", "fold", "window"])
@pytest.mark.parametrize("force", [True, False], ids=["forced", "default"])
@pytest.mark.parametrize("dtype,dim,batch", SHAPES, ids=lambda v: str(v).replace(" ", ""))
def test_slab_permute_matches_numpy(dev, monkeypatch, dtype, dim, batch, force, store):
    """store: the two store phases of the kernel (one element per lane / 16-byte chunks per
    thread, NNSB_SLAB_STORE) and the default choice between them."""
    monkeypatch.delenv("NNSB_SLAB_FOLD", raising=False)
    monkeypatch.delenv("NNSB_SLAB_WINDOW", raising=False)
    if store == "window":                          # taller slabs of row segments (k_perm_slab.cu, window)
        monkeypatch.setenv("NNSB_SLAB_WINDOW", "1")
        store = "lanes"
    if store == "fold":                            # RGB pixels as 3 / 6 / 12-byte elements
        monkeypatch.setenv("NNSB_SLAB_FOLD", "1")
        store = "lanes"
    if store:
        monkeypatch.setenv("NNSB_SLAB_STORE", store)
    else:
        monkeypatch.delenv("NNSB_SLAB_STORE", raising=False)
    if force:
        monkeypatch.setenv("NNSB_PERM_SLAB", "2")
    else:
        monkeypatch.delenv("NNSB_PERM_SLAB", raising=False)
    n = int(np.prod(dim))
    rng = np.random.default_rng(n)
    x = rng.integers(0, 250, n * batch).astype(dtype)
    used = 0
    for order in ORDERS:
        plan = dev.plan("transpose", "%d:%d:%d:3" % order, dtype, dim)
        used += "slabs" in plan.describe()
        # an odd base offset inside the allocation: the batch does not start 16-byte aligned
        pad = np.dtype(dtype).itemsize * 3
        d_in = dev.alloc(x.nbytes + 64)
        d_out = dev.alloc(x.nbytes + 64)
        dev.h2d(d_in + pad, x)
        dev.sync()
        plan.launch(d_in + pad, d_out + pad, batch)
        got = dev.download(d_out + pad, x.size, np.dtype(dtype))
        plan.close()
        dev.free(d_in)
        dev.free(d_out)
        want = np_transpose(x, dim, order, batch)
        assert bits_equal(got, want), (order, dtype, dim)
    if force and dim[0] > 1:                       # (a plain 2-D byte transpose has no long runs)
        assert used > 0, "the slab kernel was never selected for %s %s" % (dtype, dim)


@pytest.mark.parametrize("order", [(1, 2, 0), (0, 2, 1), (2, 1, 0)])
@pytest.mark.parametrize("store", ["lanes", "chunks", "window"])
def test_slab_leaves_guard_bytes_alone(dev, monkeypatch, store, order):
    """Partial head / tail chunks of a run are written element by element: nothing outside."""
    monkeypatch.setenv("NNSB_PERM_SLAB", "2")
    monkeypatch.setenv("NNSB_SLAB_WINDOW", "1" if store == "window" else "0")
    monkeypatch.setenv("NNSB_SLAB_STORE", "lanes" if store == "window" else store)
    dim, batch, dtype = [3, 225, 113, 1], 3, "uint16"
    n = int(np.prod(dim))
    x = np.random.default_rng(1).integers(0, 65535, n * batch).astype(dtype)
    plan = dev.plan("transpose", "%d:%d:%d:3" % order, dtype, dim)
    guard = 256
    d_in = dev.upload(x)
    host = np.full(x.nbytes + 2 * guard + 6, 0xA5, dtype=np.uint8)
    d_out = dev.upload(host)
    plan.launch(d_in, d_out + guard + 6, batch)
    back = dev.download(d_out, host.size, np.uint8)
    plan.close()
    assert np.all(back[:guard + 6] == 0xA5) and np.all(back[guard + 6 + x.nbytes:] == 0xA5)
    assert back[guard + 6:guard + 6 + x.nbytes].tobytes() == np_transpose(x, dim, order, batch).tobytes()
    dev.free(d_in)
    dev.free(d_out)


@pytest.mark.parametrize("w,h,batch", [(1920, 1080, 3), (64, 64, 1), (80, 8, 5), (16, 200, 2), (1936, 1096, 1)])
def test_rgb_pixel_transpose(dev, w, h, batch):
    """transpose 0:2:1:3 of packed RGB bytes (k_perm_pix.cu): whole tiles, edge tiles in both
    directions, batches; and widths the kernel must refuse (W % 16 != 0) still come out right."""
    for ww, hh in ((w, h), (w + 2, h)):
        dim = [3, ww, hh, 1]
        n = 3 * ww * hh
        x = np.random.default_rng(ww * hh).integers(0, 256, n * batch, dtype=np.uint8)
        plan = dev.plan("transpose", "0:2:1:3", "uint8", dim)
        got = plan.run_host(x, batch=batch)
        plan.close()
        assert bits_equal(got, np_transpose(x, dim, (0, 2, 1), batch)), (ww, hh, batch)
