"""Shared test-data generators (seeded, deterministic)."""
import numpy as np

TYPES = ["int32", "uint32", "int16", "uint16", "int8", "uint8", "float64", "float32",
         "int64", "uint64", "float16"]
ORC_TYPES = ["int32", "uint32", "int16", "uint16", "int8", "uint8", "float64", "float32"]


def rand_full(dtype, n, seed):
    """Full-range values incl. edge cases; floats span many magnitudes, +-0, inf, nan."""
    rng = np.random.default_rng(seed)
    dt = np.dtype(dtype)
    if dt.kind == "f":
        with np.errstate(over="ignore"):
            x = (rng.standard_normal(n) * np.float64(10) ** rng.integers(-4, 13, n)).astype(dt)
        special = np.array([0.0, -0.0, 1.0, -1.0, 0.5, -0.5, 127.5, 128.0, 255.0, 256.0, 32767.0,
                            32768.0, 65535.0, 65536.0, 2147483647.0, 2147483648.0, -2147483648.0,
                            -2147483904.0, 4294967296.0, 9.3e18, -9.3e18, np.inf, -np.inf, np.nan],
                           dtype=np.float64)
        with np.errstate(over="ignore"):
            sp = special.astype(dt)
        k = min(len(sp), n)
        x[:k] = sp[:k]
        # NaNs with payloads (quiet/signalling, both signs): x86 propagates them, so must we
        ui = {2: np.uint16, 4: np.uint32, 8: np.uint64}[dt.itemsize]
        pats = {2: [0x7e01, 0xfe00, 0x7d55, 0xfc01, 0x7fff],
                4: [0x7fc00001, 0xffc00000, 0x7fa55555, 0xff800001, 0x7fffffff],
                8: [0x7ff8000000000001, 0xfff8000000000000, 0x7ff5555555555555,
                    0xfff0000000000001, 0x7fffffffffffffff]}[dt.itemsize]
        if n >= k + len(pats):
            x.view(ui)[k:k + len(pats)] = np.array(pats, dtype=ui)
        return x
    info = np.iinfo(dt)
    x = rng.integers(info.min, info.max, n, dtype=dt, endpoint=True)
    edge = np.array([info.min, info.max, 0, 1, info.max // 2, info.min // 2 if info.min else 2],
                    dtype=dt)
    k = min(len(edge), n)
    x[:k] = edge[:k]
    return x


def rand_small(dtype, n, seed, lo=-100, hi=100):
    """In-range values (what the reference's own tests use)."""
    rng = np.random.default_rng(seed)
    dt = np.dtype(dtype)
    if dt.kind == "f":
        return (rng.random(n) * (hi - lo) + lo).astype(dt)
    info = np.iinfo(dt)
    return rng.integers(max(lo, info.min), min(hi, info.max), n, endpoint=True).astype(dt)


def bits_equal(a, b):
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    return a.dtype == b.dtype and a.shape == b.shape and a.tobytes() == b.tobytes()


def first_mismatch(a, b, x=None):
    au = np.ascontiguousarray(a).view(np.uint8).reshape(a.size, -1)
    bu = np.ascontiguousarray(b).view(np.uint8).reshape(b.size, -1)
    idx = np.nonzero((au != bu).any(axis=1))[0]
    if len(idx) == 0:
        return "no mismatch"
    i = int(idx[0])
    s = "%d mismatches; first at %d: got %r want %r" % (len(idx), i, a.reshape(-1)[i], b.reshape(-1)[i])
    if x is not None:
        s += " (input %r)" % (x.reshape(-1)[i],)
    return s
