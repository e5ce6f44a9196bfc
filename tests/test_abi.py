"""The C-ABI library loads on a CPU-only box and exports every symbol include/*.h declares
(no compute calls here); the product path fails loudly without a GPU instead of falling back."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = "".join(open(os.path.join(ROOT, "include", f)).read() for f in sorted(os.listdir(os.path.join(ROOT, "include")))
                  if f.endswith(".h"))  # gsr.h (the boundary) + gsr_measure.h (bench.py only)
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gsr_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from gaussiansplattingviewer_b200 import _gsr, build
    build.build()
    names = _declared()
    assert len(names) >= 24 and "gsr_render" in names and "gsr_sort_pairs" in names
    lib = ctypes.CDLL(_gsr.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(_gsr.SIGNATURES) == names  # the ctypes table binds exactly the header
    _gsr.load()
    assert _gsr.load().gsr_version() >= 100


def test_no_cpu_fallback():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
    with pytest.raises(RuntimeError):
        B200Renderer(64, 64)
    # the C ABI itself also refuses
    from gaussiansplattingviewer_b200 import _gsr
    h = ctypes.c_void_p()
    assert _gsr.load().gsr_create(0, ctypes.byref(h)) != 0
    assert b"CUDA" in _gsr.load().gsr_last_error() or b"device" in _gsr.load().gsr_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "gaussiansplattingviewer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.replace("the oracle", "").replace("CPU oracle", "").replace("oracle/", "ORACLE_DOC/") \
                    or f.endswith((".cu", ".h")), (f, "imports or mentions oracle")
                assert "import oracle" not in txt and "from oracle" not in txt, f
