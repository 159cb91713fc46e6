"""The C-ABI library loads on a CPU-only machine, exports every symbol include/ternary_b200.h
declares, and refuses to compute without an sm_100 GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest
import torch

from ternary_vllm_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "ternary_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tb200_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(_lib.SIGNATURES)


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "build with `make -C ternary_vllm_b200/csrc`"
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(handle, name), name
    assert _lib.lib().tb200_abi_version() == 1


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only behaviour")
def test_no_cpu_fallback():
    l = _lib.lib()
    rc = l.tb200_clip_latent(ctypes.c_void_p(64), 0, 4, None)
    assert rc == -3 and b"no CPU fallback" in l.tb200_last_error()
    rc = l.tb200_gemv(ctypes.c_void_p(64), 1, 1, 64, 64, ctypes.c_void_p(64), 16, 4, None, None, 100.0,
                      ctypes.c_void_p(64), 1, 4, None, None, None, 0, None)
    assert rc == -3
    import ternary_vllm_b200 as tb
    m = tb.ScaledTernaryLinear(8, 4)
    with pytest.raises(tb.TernaryB200Error):
        with torch.no_grad():
            m(torch.randn(1, 8))
    with pytest.raises(tb.TernaryB200Error):
        tb.ops.quantize_pack(torch.randn(4, 8))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "ternary_vllm_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("test oracle", ""), f"{f} must not reference oracle/"
