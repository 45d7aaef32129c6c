"""ctypes front end of the CPU checker (oracle/dfq_oracle.c) -- TEST INFRASTRUCTURE ONLY.

Works on plain numpy float32 arrays.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libdfq_oracle.so")

_f32p = ctypes.POINTER(ctypes.c_float)
_i32p = ctypes.POINTER(ctypes.c_int32)
_i64 = ctypes.c_int64


def build(force=False):
    src = os.path.join(_HERE, "dfq_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
        _lib.dfqo_mean_abs_diff.restype = ctypes.c_float
    return _lib


def _p(a):
    if a is None:
        return None
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_f32p)


def minmax(x):
    mn, mx = ctypes.c_float(), ctypes.c_float()
    lib().dfqo_minmax_f32(_p(x), _i64(x.size), ctypes.byref(mn), ctypes.byref(mx))
    return mn.value, mx.value


def quant_params(min_value, max_value, bits, symmetric=False, range_f32=False):
    o = [ctypes.c_float() for _ in range(5)]
    lib().dfqo_quant_params(ctypes.c_double(min_value), ctypes.c_double(max_value), int(bits),
                            int(symmetric), int(range_f32), *[ctypes.byref(v) for v in o])
    return tuple(v.value for v in o)          # neg_min, scale, qmin, qmax, add_back


def fake_quant(x, bits=8, min_value=None, max_value=None, symmetric=False, want_codes=False):
    """utils/quantize.py:88 quantize(); min/max None => the global 0-dim fp32 path."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    range_f32 = min_value is None or max_value is None
    if range_f32:
        lo, hi = minmax(x)
        min_value = lo if min_value is None else min_value
        max_value = hi if max_value is None else max_value
    prm = quant_params(float(min_value), float(max_value), bits, symmetric, range_f32)
    y = np.empty_like(x)
    codes = np.empty(x.shape, dtype=np.int32) if want_codes else None
    lib().dfqo_fake_quant_f32(_p(x), _p(y), codes.ctypes.data_as(_i32p) if want_codes else None,
                              _i64(x.size), *[ctypes.c_float(v) for v in prm])
    return (y, codes) if want_codes else y


def quantize_per_channel(x, bits=8, want_codes=False):
    x = np.ascontiguousarray(x, dtype=np.float32)
    rows = x.shape[0]
    y = np.empty_like(x)
    codes = np.empty(x.shape, dtype=np.int32) if want_codes else None
    lib().dfqo_quantize_per_channel(_p(x), _p(y), codes.ctypes.data_as(_i32p) if want_codes else None,
                                    _i64(rows), _i64(x.size // rows), int(bits))
    return (y, codes) if want_codes else y


def merge_bn(W, b, gamma, beta, mean, var, eps):
    """In place on W [O,...] and b [O]."""
    O = W.shape[0]
    lib().dfqo_merge_bn(_p(W), _p(b), _i64(O), _i64(W.size // O), _p(gamma), _p(beta), _p(mean),
                        _p(var), ctypes.c_float(eps))


def _w2_dims(W2):
    O2, I2 = W2.shape[0], W2.shape[1]
    return O2, I2, W2.size // (O2 * I2)


def cle_pair(W1, W2, b1, bn_w, bn_b, smin=1e-8, smax=1e8, signed=False, eps=0.0):
    """In place; returns S [C1].  Cross_layer_equal.py:11-59."""
    C1 = W1.shape[0]
    O2, I2, KK = _w2_dims(W2)
    S = np.zeros(C1, dtype=np.float32)
    lib().dfqo_cle_pair(_p(W1), _i64(C1), _i64(W1.size // C1), _p(W2), _i64(O2), _i64(I2), _i64(KK),
                        _p(b1), _p(bn_w), _p(bn_b), _p(S), ctypes.c_double(smin),
                        ctypes.c_double(smax), int(signed), ctypes.c_float(eps))
    return S


def mean_abs_diff(a, b):
    return float(lib().dfqo_mean_abs_diff(_p(a), _p(b), _i64(a.size)))


def absorb_pair(W2, C1, b1, b2, bn_w, bn_b, N=3.0):
    O2, I2, KK = _w2_dims(W2)
    c = np.zeros(C1, dtype=np.float32)
    lib().dfqo_absorb_pair(_p(W2), _i64(O2), _i64(I2), _i64(KK), _i64(C1), _p(b1), _p(b2), _p(bn_w),
                           _p(bn_b), ctypes.c_float(N), _p(c))
    return c


def clip(w, lo, hi):
    lib().dfqo_clip(_p(w), _i64(w.size), ctypes.c_float(lo), ctypes.c_float(hi))


def bias_correct(W, v, bits=8, signed=False, want_eps=False):
    O2, I2, KK = _w2_dims(W)
    out = np.zeros(O2, dtype=np.float32)
    eps = np.zeros((O2, I2), dtype=np.float32) if want_eps else None
    lib().dfqo_bias_correct(_p(W), _i64(O2), _i64(I2), _i64(KK), int(bits), int(signed), _p(v), _p(out),
                            _p(eps))
    return (out, eps) if want_eps else out


def sample_minmax_mean(x):
    N = x.shape[0]
    mn, mx = ctypes.c_float(), ctypes.c_float()
    lib().dfqo_sample_minmax_mean(_p(x), _i64(N), _i64(x.size // N), ctypes.byref(mn), ctypes.byref(mx))
    return mn.value, mx.value
