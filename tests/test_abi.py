"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol include/dfq_b200.h declares."""
import ctypes
import os
import re

from conftest import ROOT


def header_symbols():
    text = open(os.path.join(ROOT, "include", "dfq_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dfq_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from data_free_quantization_b200 import _build, _lib
    so = _build.build()
    assert os.path.exists(so)
    h = ctypes.CDLL(so)
    syms = header_symbols()
    assert len(syms) >= 19
    for s in syms:
        assert hasattr(h, s), f"{s} declared in the header but not exported"
    assert set(_lib.PROTOTYPES) == set(syms), "ctypes prototypes and header disagree"
    lib = _lib.lib()
    assert lib.dfq_abi_version() == _lib.ABI_VERSION


def test_struct_sizes_match_header_layout():
    from data_free_quantization_b200 import _lib as L
    assert ctypes.sizeof(L.DfqQuantParams) == 20
    assert ctypes.sizeof(L.DfqQuantDesc) == 72
    assert ctypes.sizeof(L.DfqClePair) == 80
    assert ctypes.sizeof(L.DfqCleLayer) == 32
    assert ctypes.sizeof(L.DfqAbsorbPair) == 80
    assert ctypes.sizeof(L.DfqCorrDesc) == 104
    assert ctypes.sizeof(L.DfqPwI8Desc) == 144
    assert ctypes.sizeof(L.DfqConvI8Desc) == 144


def test_host_quant_params_match_oracle(oracle):
    """dfq_quant_params_host is pure host code: check it against the oracle without a GPU."""
    from data_free_quantization_b200 import _lib as L
    for bits in (4, 6, 8, 16):
        for sym in (0, 1):
            for rf in (0, 1):
                for lo, hi in ((-0.731, 1.207), (0.0, 6.0), (-1e-12, 1e-12), (-3.5, -0.25)):
                    p = L.DfqQuantParams()
                    L.lib().dfq_quant_params_host(lo, hi, bits, sym | (2 if rf else 0), ctypes.byref(p))
                    want = oracle.quant_params(lo, hi, bits, bool(sym), bool(rf))
                    got = (p.neg_min, p.scale, p.qmin, p.qmax, p.add_back)
                    assert got == want, (bits, sym, rf, lo, hi, got, want)


def test_no_cpu_fallback():
    import pytest
    import torch
    from data_free_quantization_b200.utils.quantize import quantize
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        quantize(torch.randn(8, 8), 8, -1.0, 1.0)
