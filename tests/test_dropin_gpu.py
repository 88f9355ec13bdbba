"""The drop-in C++ API (include/ivf-hnsw/*.h + libivfhnsw_dropin.so) driven by a C++ program that is written
like the reference's own mains (tests/cpp/dropin_smoke.cpp): files on disk, train_pq, assign, add_batch /
add_group from an OpenMP loop, write, read into a fresh index, one search() call per query, recall."""
import os
import subprocess
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "ivf-hnsw_b200", "dropin_smoke")


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [("adc",), ("grouping",), ("adc", "opq"), ("grouping", "opq")])
def test_cpp_dropin_end_to_end(mode):
    assert os.path.exists(BIN), "build() did not produce dropin_smoke"
    with tempfile.TemporaryDirectory() as tmp:
        r = subprocess.run([BIN, tmp] + list(mode), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "dropin_smoke OK" in r.stdout


def test_reference_mains_compile_against_the_dropin_headers():
    """Source compatibility of the boundary: the reference's UNMODIFIED tests/*.cpp compile and link against
    include/ivf-hnsw, include/faiss, include/hnswlib (only where the reference tree is mounted)."""
    if not os.path.isdir("/root/reference/tests"):
        pytest.skip("reference tree not mounted on this machine")
    if not os.path.exists(os.path.join(ROOT, "ivf-hnsw_b200", "libivfhnsw_dropin.so")):
        pytest.skip("drop-in library not built (run __graft_entry__.build())")
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "ivf-hnsw_b200", "csrc"), "refcheck"], capture_output=True,
                       text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("refcheck OK") == 5
