"""CPU: the C-ABI library loads without a GPU and exports exactly what include/rptx.h declares.
No compute call is made here (there is no device in this container)."""
import ctypes as C
import os
import re

import pytest

from rpt_b200 import _abi as A

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "rptx.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rptx_[a-z0-9_]+)\s*\(", text)))


def test_header_and_prototypes_agree():
    syms = _header_symbols()
    assert len(syms) >= 25
    assert sorted(A.PROTOTYPES) == syms


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(A.LIB_PATH)
    for name in _header_symbols():
        assert hasattr(lib, name), f"librptx.so does not export {name}"
    assert A.load().rptx_abi_version() == 2


def test_struct_layouts_match_the_header():
    # sizes follow from the C declarations in include/rptx.h (natural alignment, LP64)
    assert C.sizeof(A.ShapeDesc) == 4 + 4 + 128 + 4 + 4 + 24 + 8 + 16 + 8
    assert C.sizeof(A.MaterialDesc) == 8 + 24 + 24
    assert C.sizeof(A.LightDesc) == 8 + 48 + C.sizeof(A.ShapeDesc) + C.sizeof(A.MaterialDesc)
    assert C.sizeof(A.CameraDesc) == 96
    assert C.sizeof(A.Params) == 8 + 8 + 16 + 16 + 8 + 8 + 8 + 8


def test_no_device_is_an_error_not_a_fallback():
    """Without a CUDA device the product must fail loudly (no CPU path exists)."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a device is present")
    from rpt_b200 import Context, RptxError

    with pytest.raises(RptxError, match="no CUDA device"):
        Context(0)


def test_product_does_not_import_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may touch oracle/."""
    pkg = os.path.join(ROOT, "rpt_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h")) or f == "Makefile":
                text = open(os.path.join(dirpath, f)).read()
                for needle in ("liborc", "import orc", "oracle/", "orc_"):
                    assert needle not in text, f"rpt_b200/{f} references the oracle ({needle})"
