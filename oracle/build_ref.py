#!/usr/bin/env python3
"""oracle/build_ref.py -- TEST INFRASTRUCTURE ONLY.

Compiles the REFERENCE's own CPU code for the tensor_transform hot path into
``oracle/_ref/libnns_ref.so`` (git-ignored; it travels to the GPU box like our own .so).

Nothing from /root/reference is copied into the repository: the reference sources are read
where they lie, the needed line ranges are streamed into a temporary translation unit under
a fresh /tmp directory (deleted afterwards) and compiled with the same language/FP settings
the reference uses (meson.build:9-12 ``c_std=gnu89``; SURVEY.md section 8c:
``-O2 -ffp-contract=off``).

  TU 1: gst/nnstreamer/tensor_data.c                       (unmodified, whole file)
  TU 2: oracle/ref_prelude.h                               (our stubs for GLib/GStreamer names)
        gsttensor_transform.h:54-168                        (mode/option structs)
        gsttensor_transform.c:82-86                         (transpose rank limit)
        gsttensor_transform.c:333-457                       (float16 helpers / stubs)
        gsttensor_transform.c:560-645                       (handle_operator, do_operator)
        gsttensor_transform.c:1111-1658                     (dimchg, typecast, arithmetic,
                                                             transpose, stand, clamp)
        gsttensor_transform.c:1841-1959                     (convert_dimension)
        oracle/ref_wrap.c                                   (our plain-C entry points)

The ORC path (HAVE_ORC) is NOT built: liborc/orcc are un-vendored third-party dependencies
(meson.build:174-175) absent from this image.  Its semantics are restated in
oracle/nns_oracle.c ("parity unpinned" for overflow cases).

If /root/reference is absent (the GPU box) this script does nothing and keeps the prebuilt file.
"""
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("NNS_REFERENCE_ROOT", "/root/reference")
OUT_DIR = os.path.join(HERE, "_ref")
OUT = os.path.join(OUT_DIR, "libnns_ref.so")

RANGES = [
    ("gst/nnstreamer/elements/gsttensor_transform.h", 54, 168),
    ("gst/nnstreamer/elements/gsttensor_transform.c", 82, 86),
    ("gst/nnstreamer/elements/gsttensor_transform.c", 333, 457),
    ("gst/nnstreamer/elements/gsttensor_transform.c", 560, 645),
    ("gst/nnstreamer/elements/gsttensor_transform.c", 1111, 1658),
    ("gst/nnstreamer/elements/gsttensor_transform.c", 1841, 1959),
]

CFLAGS = ["-O2", "-std=gnu89", "-ffp-contract=off", "-fPIC", "-DFLOAT16_SUPPORT",
          "-Wno-unused-function", "-Wno-unused-variable"]


def _lines(path, lo, hi):
    with open(os.path.join(REF, path), "r", encoding="utf-8") as fh:
        all_lines = fh.readlines()
    return "".join(all_lines[lo - 1:hi])


def reference_available():
    return os.path.isfile(os.path.join(REF, "gst/nnstreamer/tensor_data.c"))


def build(force=False, verbose=False):
    """Build oracle/_ref/libnns_ref.so. Returns its path, or None if it cannot be built here."""
    if not reference_available():
        return OUT if os.path.isfile(OUT) else None
    srcs = [os.path.join(REF, r[0]) for r in RANGES] + [
        os.path.join(REF, "gst/nnstreamer/tensor_data.c"),
        os.path.join(HERE, "ref_wrap.c"), os.path.join(HERE, "ref_prelude.h"),
        os.path.join(HERE, "shim", "glib.h"), os.path.abspath(__file__)]
    if (not force and os.path.isfile(OUT)
            and all(os.path.getmtime(OUT) >= os.path.getmtime(s) for s in srcs)):
        return OUT
    os.makedirs(OUT_DIR, exist_ok=True)
    tmp = tempfile.mkdtemp(prefix="nns_ref_build_")
    try:
        tu = ['#include "ref_prelude.h"\n', '#include "tensor_data.h"\n']
        for path, lo, hi in RANGES:
            tu.append('#line %d "%s"\n' % (lo, os.path.join(REF, path)))
            tu.append(_lines(path, lo, hi))
        tu.append('#line 1 "%s"\n' % os.path.join(HERE, "ref_wrap.c"))
        with open(os.path.join(HERE, "ref_wrap.c"), "r", encoding="utf-8") as fh:
            tu.append(fh.read())
        tu_path = os.path.join(tmp, "ref_transform_tu.c")
        with open(tu_path, "w", encoding="utf-8") as fh:
            fh.write("".join(tu))
        inc = ["-I" + os.path.join(HERE, "shim"), "-I" + HERE,
               "-I" + os.path.join(REF, "gst/nnstreamer/include"),
               "-I" + os.path.join(REF, "gst/nnstreamer")]
        objs = []
        for src, obj in ((os.path.join(REF, "gst/nnstreamer/tensor_data.c"), "tensor_data.o"),
                         (tu_path, "ref_transform_tu.o")):
            o = os.path.join(tmp, obj)
            cmd = ["gcc"] + CFLAGS + inc + ["-c", src, "-o", o]
            if verbose:
                print(" ".join(cmd))
            subprocess.check_call(cmd)
            objs.append(o)
        cmd = ["gcc", "-shared", "-o", OUT] + objs + ["-lm"]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return OUT


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose=True)
    print("built" if p else "reference not available and no prebuilt library", p or "")
