"""The C-ABI library loads and exports every symbol include/mocob200.h declares (no compute calls: CPU only)."""
import ctypes
import os
import re

from conftest import ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "mocob200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mocob200_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(pkg):
    lib = ctypes.CDLL(pkg.LIB_PATH)
    syms = declared_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(lib, s), "libmocob200.so does not export %s" % s
    assert sorted(pkg.ABI_SYMBOLS) == syms, "harness symbol list out of date with the header"


def test_no_device_means_loud_failure(pkg):
    """Creating an evaluator without a GPU must fail with an error, never fall back to a CPU path."""
    import torch

    if torch.cuda.is_available():
        return
    import pytest
    from conftest import golden

    p = pkg.Problem.load(golden("sliding_mass.mb20"))
    with pytest.raises(RuntimeError, match="no CUDA device|no CPU path|CUDA"):
        pkg.Evaluator(p)


def test_product_does_not_reference_oracle():
    """The product tree never includes/links/imports anything under oracle/."""
    pk = os.path.join(ROOT, "opensim-moco_b200")
    for dirpath, _, files in os.walk(pk):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".cu", ".cuh", ".cpp", ".h", ".py")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle/" not in src and "liborc" not in src and "oracle_py" not in src, f
