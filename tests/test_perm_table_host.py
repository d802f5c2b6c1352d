"""Host-only check of the vector-permute plan builder (nnstreamer_b200/csrc/k_perm_vec.cu).

tests/native/perm_table_check.cu includes the kernel source to reach its host-side builder,
builds the box table for a permutation, replays on the CPU what the kernel does with the table
(tile decode, edge predicates, swizzled shared-memory slots) and compares with the definition of
gst_tensor_transform_transpose (gsttensor_transform.c:1417-1500).  No GPU needed: nvcc only
cross-compiles.  The GPU parity of the same shapes is in tests/test_gpu_parity.py.
"""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "native", "perm_table_check.cu")
NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    if not os.path.exists(NVCC):
        pytest.skip("nvcc not available")
    exe = str(tmp_path_factory.mktemp("perm") / "perm_table_check")
    subprocess.run([NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-std=c++17",
                    "-I" + os.path.join(ROOT, "include"),
                    "-I" + os.path.join(ROOT, "nnstreamer_b200", "csrc"), "-o", exe, SRC],
                   check=True, capture_output=True, timeout=600)
    return exe


def run(exe, es, batch, dim, order):
    args = [exe, str(es), str(batch)] + [str(d) for d in dim] + [str(o) for o in order]
    out = subprocess.run(args, capture_output=True, text=True, timeout=300)
    return out.returncode, out.stdout.strip()


ORDERS = [(0, 2, 1, 3), (1, 0, 2, 3), (1, 2, 0, 3), (2, 0, 1, 3), (2, 1, 0, 3)]
CASES = [
    # es, batch, dim
    (4, 1, (3, 384, 216, 1)),       # video-shaped float32, whole boxes
    (4, 2, (3, 132, 100, 1)),       # partial boxes on every side, batch
    (4, 3, (4, 36, 20, 1)),         # tensor smaller than a box; float4 folds to 16-byte elements
    (4, 1, (8, 12, 52, 1)),
    (8, 2, (3, 130, 98, 1)),        # 2-element vectors
    (8, 1, (6, 34, 18, 1)),
    (4, 1, (64, 64, 64, 1)),
    (1, 2, (4, 68, 36, 1)),         # RGBA bytes fold into one 4-byte element
    (1, 1, (16, 20, 12, 1)),
    (2, 3, (8, 12, 20, 1)),
    (1, 1, (3, 192, 112, 1)),       # RGB bytes: 16 one-byte elements per vector
    (1, 2, (3, 160, 80, 1)),
    (2, 1, (3, 64, 48, 1)),         # 8 two-byte elements per vector
    (2, 2, (5, 32, 24, 1)),
    (1, 1, (3, 1080, 64, 1)),       # 3240-byte rows: 8-byte vectors
    (4, 2, (3, 227, 31, 1)),        # 4-byte vectors
    (2, 1, (3, 226, 34, 1)),        # 2-byte elements in 4-byte vectors
    (4, 1, (3, 3840, 16, 1)),       # long rows
    (4, 1, (1, 2160, 512, 1)),      # plain 2-D transposes
]


@pytest.mark.parametrize("order", ORDERS)
@pytest.mark.parametrize("case", CASES, ids=lambda c: "es%d-b%d-%s" % (c[0], c[1], "x".join(map(str, c[2]))))
def test_table_replay_matches_definition(checker, case, order):
    es, batch, dim = case
    rc, out = run(checker, es, batch, dim, order)
    assert rc == 0, out
    assert out.startswith("ok") or out == "ineligible", out


def test_expected_boxes(checker):
    # 2-D transpose of 4-byte elements: a whole 64x64 box, gather at most 2-way conflicted
    rc, out = run(checker, 4, 1, (3, 3840, 2160, 1), (2, 0, 1, 3))
    assert rc == 0 and "box=64x64" in out and "n_vec=1024" in out, out
    waves = int(out.split("wavefronts=")[1].split()[0])
    assert waves <= 256, out
    # rows that are 8- but not 16-byte multiples (1080 RGB bytes = 3240) take 8-byte vectors,
    # float rows that are only 4-byte multiples take 4-byte ones
    rc, out = run(checker, 1, 1, (3, 1080, 1920, 1), (0, 2, 1, 3))
    assert rc == 0 and out.startswith("ok vb=8 es=1"), out
    rc, out = run(checker, 4, 4, (5, 7, 9, 1), (2, 1, 0, 3))
    assert rc == 0 and out.startswith("ok vb=4 es=4"), out
    rc, out = run(checker, 1, 1, (3, 1920, 1088, 1), (0, 2, 1, 3))
    assert rc == 0 and out.startswith("ok vb=16 es=1"), out
    # odd byte counts cannot be vectorised at all: the scalar kernel takes them
    rc, out = run(checker, 2, 1, (3, 225, 113, 1), (0, 2, 1, 3))
    assert rc == 0 and out == "ineligible", out
    rc, out = run(checker, 1, 3, (3, 50, 100, 1), (2, 0, 1, 3))
    assert rc == 0 and out == "ineligible", out
