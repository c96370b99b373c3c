"""The C-ABI library loads and exports every symbol include/*.h declares (no compute calls: runs without a GPU)."""

import ctypes
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for header in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = open(header).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names.update(re.findall(r"\b(rpvg_b200_[a-z0-9_]+)\s*\(", text))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    from rpvg_b200 import capi
    lib = ctypes.CDLL(capi.library_path())
    names = declared_symbols()
    assert len(names) >= 10
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/ but not exported: {missing}"


def test_no_device_is_a_loud_error():
    """Without a CUDA device the product path refuses to run; it never falls back to a CPU implementation."""
    import torch
    from rpvg_b200 import capi
    if torch.cuda.is_available():
        return
    try:
        capi.Context(0)
    except capi.RpvgB200Error as e:
        assert "no CPU fallback" in str(e)
    else:
        raise AssertionError("Context() must fail without a GPU")


def test_product_does_not_reference_the_oracle():
    """Nothing under rpvg_b200/ or include/ may import, link or name the oracle."""
    offenders = []
    for base in ("rpvg_b200", "include"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h", "Makefile")):
                    text = open(os.path.join(dirpath, f), errors="ignore").read()
                    if re.search(r"\boracle\b", text):
                        offenders.append(os.path.join(dirpath, f))
    assert not offenders, offenders
