"""The GStreamer glue (nnstreamer_b200/gst/*.c) cannot be built in this image (no GLib /
GStreamer), but it must at least be valid C against the GStreamer 1.x API it uses: `make syntax`
runs gcc -fsyntax-only over the three files against declaration-only stubs (tests/gst_stub,
never shipped) -- or against the real headers where pkg-config finds them."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("gcc") is None or shutil.which("make") is None, reason="needs gcc and make")
def test_glue_sources_pass_the_syntax_check():
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "nnstreamer_b200", "gst"), "syntax"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    for name in ("gsttensor_transform_b200.c", "gsttensor_converter_b200.c", "gstnnsbcudaallocator.c"):
        assert name in r.stdout
    assert "warning" not in (r.stdout + r.stderr)
