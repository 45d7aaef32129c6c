"""GPU: the CUDA kernels of the quantised eval forward (csrc/conv.cu, csrc/pw_i8.cu) one by one against a plain
PyTorch fp32 restatement of utils/quantize.py:225-256 (fake-quantise the input with the oracle's recipe, then
F.conv2d in fp32 on the CPU).  Floating-point tolerance: 2e-5 of the output's largest magnitude (accumulation
order differs); per-sample statistics exact for the values actually written."""
import ctypes

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 2e-5


def _slot(lo, hi, bits=8):
    from data_free_quantization_b200 import _lib as L
    p = L.DfqQuantParams()
    L.lib().dfq_quant_params_host(float(lo), float(hi), bits, 0, ctypes.byref(p))
    return torch.tensor([lo, hi, p.scale, 0.0], dtype=torch.float32, device=DEV)


def _decode_stats(stats):
    """[N,2] encoded {max(-y), max(y)} -> float (min, max)."""
    e = stats.cpu().numpy().astype(np.uint32)
    u = np.where(e & 0x80000000, e & 0x7fffffff, ~e).astype(np.uint32)
    f = u.view(np.float32)
    return -f[:, 0], f[:, 1]


def _conv(x, w, b, stride, pad, dil, groups, relu, rng, oracle, want_stats=True):
    from data_free_quantization_b200 import _lib as L
    N, C, H, W = x.shape
    O, _, kh, kw = w.shape
    OH = (H + 2 * pad - dil * (kh - 1) - 1) // stride + 1
    OW = (W + 2 * pad - dil * (kw - 1) - 1) // stride + 1
    xq = x if rng is None else oracle.fake_quant(x, 8, rng[0], rng[1])
    want = F.conv2d(torch.from_numpy(xq), torch.from_numpy(w), None if b is None else torch.from_numpy(b), stride, pad,
                    dil, groups)
    if relu:
        want = torch.relu(want)
    xd, wd = torch.from_numpy(x).to(DEV), torch.from_numpy(w).to(DEV)
    bd = None if b is None else torch.from_numpy(b).to(DEV)
    y = torch.full((N, O, OH, OW), float("nan"), device=DEV)
    stats = torch.zeros(N, 2, dtype=torch.int32, device=DEV)
    slot = None if rng is None else _slot(*rng)
    d = L.DfqConvDesc()
    d.x, d.y, d.w, d.bias = xd.data_ptr(), y.data_ptr(), wd.data_ptr(), None if bd is None else bd.data_ptr()
    d.in_slot = None if slot is None else slot.data_ptr()
    d.out_stats = stats.data_ptr() if want_stats else None
    d.N, d.C_in, d.H, d.W, d.C_out, d.OH, d.OW = N, C, H, W, O, OH, OW
    d.kh, d.kw, d.stride, d.pad, d.dil, d.groups, d.relu, d.bits = kh, kw, stride, pad, dil, groups, int(relu), 8
    L.check(L.lib().dfq_qconv2d_f32(ctypes.byref(d), L.stream()), "dfq_qconv2d_f32")
    got = y.cpu()
    err = float((got - want).abs().max() / want.abs().max())
    assert err <= TOL, err
    if want_stats:
        lo, hi = _decode_stats(stats)
        g = got.view(N, -1).numpy()
        assert np.array_equal(lo, g.min(1)) and np.array_equal(hi, g.max(1))


@pytest.mark.parametrize("C,H,stride,N", [(32, 112, 1, 2), (96, 112, 2, 2), (144, 56, 1, 3), (192, 28, 2, 2),
                                          (384, 14, 1, 2), (960, 7, 1, 3), (5, 9, 1, 1), (7, 10, 2, 2)])
@pytest.mark.parametrize("quant", [False, True])
def test_depthwise3x3(oracle, C, H, stride, N, quant):
    rng = np.random.default_rng(C + H)
    x = np.maximum(rng.standard_normal((N, C, H, H)), 0).astype(np.float32)
    w = rng.standard_normal((C, 1, 3, 3)).astype(np.float32)
    b = rng.standard_normal(C).astype(np.float32)
    _conv(x, w, b, stride, 1, 1, C, True, (0.0, 2.5) if quant else None, oracle)


@pytest.mark.parametrize("C,H,W,stride,dil,N", [(32, 67, 67, 1, 1, 2), (96, 67, 67, 2, 1, 2), (144, 35, 35, 1, 1, 3),
                                                (192, 37, 37, 1, 2, 2), (7, 9, 12, 1, 1, 1), (5, 10, 7, 2, 1, 2),
                                                (3, 5, 5, 1, 2, 1), (16, 131, 131, 2, 1, 1), (8, 259, 259, 1, 1, 1),
                                                (4, 300, 40, 1, 3, 1),
                                                # one plane per warp (C >= 8, <= 2048 outputs), ragged last block
                                                (16, 21, 21, 2, 1, 2), (20, 19, 23, 1, 1, 2), (12, 15, 15, 1, 2, 1)])
@pytest.mark.parametrize("quant", [False, True])
def test_depthwise3x3_valid(oracle, C, H, W, stride, dil, N, quant):
    """pad 0 (DeepLab pads with F.pad first, modeling/segmentation/backbone/mobilenet.py:16-22): the column-per-lane
    kernel for (stride, dilation) in {(1,1), (2,1), (1,2)}, the generic kernel for the rest (dilation 3 here)."""
    rng = np.random.default_rng(C + H + W)
    x = np.maximum(rng.standard_normal((N, C, H, W)), 0).astype(np.float32)
    w = rng.standard_normal((C, 1, 3, 3)).astype(np.float32)
    b = rng.standard_normal(C).astype(np.float32)
    _conv(x, w, b if C != 7 else None, stride, 0, dil, C, C % 2 == 0, (0.0, 2.5) if quant else None, oracle)


@pytest.mark.parametrize("quant", [False, True])
def test_first_layer_dense3x3(oracle, quant):
    rng = np.random.default_rng(1)
    x = rng.standard_normal((2, 3, 224, 224)).astype(np.float32)
    w = (rng.standard_normal((32, 3, 3, 3)) * 0.2).astype(np.float32)
    b = rng.standard_normal(32).astype(np.float32)
    _conv(x, w, b, 2, 1, 1, 1, True, (-2.1179, 2.64) if quant else None, oracle)
    _conv(x[:, :, :37, :41].copy(), w[:5].copy(), None, 1, 1, 1, 1, False, (-2.0, 2.0) if quant else None, oracle)


@pytest.mark.parametrize("K,O,H,N,relu", [(32, 16, 112, 2, False), (16, 96, 56, 2, True), (960, 160, 7, 3, False),
                                          (320, 1280, 7, 2, True), (24, 144, 9, 1, True), (3, 5, 5, 2, False)])
@pytest.mark.parametrize("quant", [False, True])
def test_pointwise_fp32(oracle, K, O, H, N, relu, quant):
    rng = np.random.default_rng(K + O)
    x = np.maximum(rng.standard_normal((N, K, H, H)), 0).astype(np.float32)
    w = (rng.standard_normal((O, K, 1, 1)) / np.sqrt(K)).astype(np.float32)
    b = rng.standard_normal(O).astype(np.float32)
    _conv(x, w, b, 1, 0, 1, 1, relu, (0.0, 3.0) if quant else None, oracle)


def test_linear_add_mean(oracle):
    from data_free_quantization_b200 import _lib as L
    rng = np.random.default_rng(5)
    # linear
    N, K, O = 37, 1280, 1000
    x = np.maximum(rng.standard_normal((N, K)), 0).astype(np.float32)
    w = (rng.standard_normal((O, K)) * 0.03).astype(np.float32)
    b = rng.standard_normal(O).astype(np.float32)
    want = F.linear(torch.from_numpy(oracle.fake_quant(x, 8, 0.0, 2.0)), torch.from_numpy(w), torch.from_numpy(b))
    xd, wd, bd = (torch.from_numpy(t).to(DEV) for t in (x, w, b))
    y = torch.empty(N, O, device=DEV)
    stats = torch.zeros(N, 2, dtype=torch.int32, device=DEV)
    slot = _slot(0.0, 2.0)
    L.check(L.lib().dfq_qlinear_f32(xd.data_ptr(), y.data_ptr(), wd.data_ptr(), bd.data_ptr(), N, K, O, slot.data_ptr(), 8,
                                    0, stats.data_ptr(), L.stream()))
    assert float((y.cpu() - want).abs().max() / want.abs().max()) <= TOL
    lo, hi = _decode_stats(stats)
    assert np.array_equal(lo, y.cpu().numpy().min(1)) and np.array_equal(hi, y.cpu().numpy().max(1))
    # add: bit-exact (two fake-quants and one fp32 add)
    a = rng.standard_normal((3, 24, 56, 56)).astype(np.float32)
    c = rng.standard_normal((3, 24, 56, 56)).astype(np.float32)
    want = oracle.fake_quant(a, 8, -2.0, 2.5) + oracle.fake_quant(c, 8, -1.5, 1.0)
    ad, cd = torch.from_numpy(a).to(DEV), torch.from_numpy(c).to(DEV)
    out = torch.empty_like(ad)
    stats = torch.zeros(3, 2, dtype=torch.int32, device=DEV)
    sa, sb = _slot(-2.0, 2.5), _slot(-1.5, 1.0)
    L.check(L.lib().dfq_qadd_f32(ad.data_ptr(), cd.data_ptr(), out.data_ptr(), 3, a[0].size, sa.data_ptr(), sb.data_ptr(),
                                 8, 8, stats.data_ptr(), L.stream()))
    assert out.cpu().numpy().tobytes() == want.astype(np.float32).tobytes()
    lo, hi = _decode_stats(stats)
    assert np.array_equal(lo, want.reshape(3, -1).min(1)) and np.array_equal(hi, want.reshape(3, -1).max(1))
    # mean
    m = np.maximum(rng.standard_normal((5, 1280, 7, 7)), 0).astype(np.float32)
    want = oracle.fake_quant(m, 8, 0.0, 3.0).reshape(5, 1280, 49).mean(-1)
    md = torch.from_numpy(m).to(DEV)
    out = torch.empty(5, 1280, device=DEV)
    stats = torch.zeros(5, 2, dtype=torch.int32, device=DEV)
    sm = _slot(0.0, 3.0)
    L.check(L.lib().dfq_qmean_f32(md.data_ptr(), out.data_ptr(), 5, 1280, 49, sm.data_ptr(), 8, stats.data_ptr(), L.stream()))
    assert float(np.abs(out.cpu().numpy() - want).max() / np.abs(want).max()) <= 1e-6
