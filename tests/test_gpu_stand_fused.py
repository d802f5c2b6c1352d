"""The single-read stand kernel (k_stand_fused.cu): batches of float32 tensors whose per-SM
slice fits on the SM, normalised out of shared memory / tensor memory in alternation.  Checked
against the CPU reference (tensor_data.c:314-493, gsttensor_transform.c:1512-1619) within stand's
tolerance, on shapes that exercise: odd and even batch counts (the last tensor lives in TMEM or
in shared memory), a short last tile, 1..4 channels, both modes, repeated launches (scratch and
TMEM reuse), and the three-launch path on the same data as a cross-check."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _close(got, want, scale):
    g = got.astype(np.float64)
    w = want.astype(np.float64)
    bad = np.abs(g - w) > 1e-6 * np.abs(w) + 1e-6 * scale
    assert not bad.any(), "%d of %d beyond tolerance; worst %g" % (bad.sum(), bad.size, np.abs(g - w).max())


CASES = [
    # dim, batch, option, mode, per_channel
    ([3, 1920, 1080, 1], 5, "default", "default", False),
    ([3, 1920, 1080, 1], 4, "dc-average,per-channel:true", "dc-average", True),
    ([3, 1920, 1080, 1], 3, "default,per-channel:true", "default", True),
    ([4, 1001, 500, 1], 7, "default,per-channel:true", "default", True),     # short last tile, 4 channels
    ([2, 1284, 731, 1], 6, "dc-average,per-channel:true", "dc-average", True),
    ([1, 2500, 1001, 1], 2, "default", "default", False),
    ([3, 1280, 720, 1], 9, "default", "default", False),
]


@pytest.mark.parametrize("dim,batch,option,mode,per_channel", CASES,
                         ids=lambda v: "x".join(str(i) for i in v) if isinstance(v, list) else str(v))
def test_fused_stand_matches_the_reference(dev, port, monkeypatch, dim, batch, option, mode, per_channel):
    monkeypatch.delenv("NNSB_STAND_NO_FUSED", raising=False)
    n = int(np.prod(dim))
    rng = np.random.default_rng(n % 1000 + batch)
    x = rng.standard_normal(n * batch, dtype=np.float32).reshape(batch, n)
    x *= np.linspace(0.5, 20.0, batch, dtype=np.float32)[:, None]
    x += np.linspace(-30.0, 300.0, batch, dtype=np.float32)[:, None]
    x = x.reshape(-1)
    nch = dim[0] if per_channel else 1
    plan = dev.plan("stand", option, "float32", dim)
    for rep in range(2):                             # second launch: scratch, TMEM and pools reused
        got = plan.run_host(x, batch=batch)
        assert "single-read fused kernel" in plan.describe(), plan.describe()
        assert plan.last_launch_count == 1
        mean, std = plan.stand_stats(batch, nch)
        for f in range(batch):
            xf = x[f * n:(f + 1) * n]
            want = port.stand(xf, dim, mode, None, per_channel)
            _close(got[f * n:(f + 1) * n], want, 400.0 if mode == "dc-average" else 1.0)
            omean, ostd = port.stand_stats(xf, dim, per_channel)
            np.testing.assert_allclose(mean[f * nch:(f + 1) * nch], omean, rtol=1e-11, atol=1e-11)
            if mode == "default":
                np.testing.assert_allclose(std[f * nch:(f + 1) * nch], ostd, rtol=1e-10)
    plan.close()


def test_fused_and_three_launch_paths_agree(dev, monkeypatch):
    """Same data through both schemes: equal within stand's tolerance (the summation orders differ)."""
    dim, batch = [3, 1920, 1080, 1], 6
    n = int(np.prod(dim))
    x = (np.random.default_rng(8).random(n * batch, dtype=np.float32) * 255.0)
    plan = dev.plan("stand", "default,per-channel:true", "float32", dim)
    a = plan.run_host(x, batch=batch)
    assert "fused" in plan.describe()
    plan.close()
    monkeypatch.setenv("NNSB_STAND_NO_FUSED", "1")
    plan = dev.plan("stand", "default,per-channel:true", "float32", dim)
    b = plan.run_host(x, batch=batch)
    assert "fused" not in plan.describe() and plan.last_launch_count == 3
    plan.close()
    _close(a, b, 1.0)
    # small or huge tensors and other types never take the fused path
    monkeypatch.delenv("NNSB_STAND_NO_FUSED")
    for d, ty, bt in (([3, 64, 48, 1], "float32", 8), ([3, 1920, 1080, 1], "uint8", 4),
                      ([3, 3840, 2160, 1], "float32", 2), ([3, 1920, 1080, 1], "float32", 1)):
        plan = dev.plan("stand", "default", ty, d)
        plan.run_host(np.zeros(int(np.prod(d)) * bt, dtype=ty) + 1, batch=bt)
        assert "fused" not in plan.describe()
        plan.close()
