"""CPU tests of the drop-in boundary: the C-ABI library builds, loads, exports every symbol that
include/ivfhnsw_b200.h declares, and refuses to compute without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT, load_pkg


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "ivfhnsw_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ivfhnsw_b200_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    pkg = load_pkg()
    assert os.path.exists(pkg.LIB_PATH), "run __graft_entry__.build() first"
    L = C.CDLL(pkg.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert L.ivfhnsw_b200_abi_version() == 1
    # the python mirror binds exactly the declared set
    pkg.lib()


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    pkg = load_pkg()
    with pytest.raises(pkg.B200Error) as e:
        pkg.IndexIVF_HNSW(32, 64, 8, 8)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)


def test_bad_shapes_are_rejected_before_touching_the_gpu():
    pkg = load_pkg()
    for args in [(30, 64, 8, 8), (32, 64, 12, 8), (32, 64, 8, 4), (32, 0, 8, 8)]:
        with pytest.raises(pkg.B200Error):
            pkg.IndexIVF_HNSW(*args)


def test_product_never_imports_the_oracle():
    """The package must not reference oracle/ (a product path through the oracle voids parity)."""
    pkgdir = os.path.join(ROOT, "ivf-hnsw_b200")
    for dp, _, files in os.walk(pkgdir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle" not in txt.lower().replace("# oracle-free", ""), os.path.join(dp, f)
