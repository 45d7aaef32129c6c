"""GPU: every CUDA kernel of libdfq_b200 against the CPU oracle (bit-exact for quantise / CLE / clip / BN-merge,
1e-5 relative for the GEMV-like reductions), called through the C ABI via the reference-shaped Python API."""
import numpy as np
import pytest
import torch

from helpers import rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


@pytest.mark.parametrize("n", [1, 7, 4096, 4097, 65537, 1 << 20])
@pytest.mark.parametrize("bits,symmetric", [(8, False), (4, False), (6, True), (16, False), (2, True)])
def test_fake_quant_given_range_bit_exact(oracle, n, bits, symmetric):
    from data_free_quantization_b200.utils.quantize import quantize
    rng = np.random.default_rng(n + bits)
    x = (rng.standard_normal(n) ** 3 * 0.2).astype(np.float32)
    lo, hi = float(np.quantile(x, 0.02)), float(np.quantile(x, 0.98))      # exercises the clamp
    if n == 1:
        lo, hi = -1.0, 1.0
    want = oracle.fake_quant(x, bits, lo, hi, symmetric)
    got = quantize(_t(x), bits, lo, hi, symmetric=symmetric)
    assert got.cpu().numpy().tobytes() == want.tobytes()


@pytest.mark.parametrize("shape", [(3,), (32, 3, 3, 3), (1000, 1280), (17, 5), (96, 1, 3, 3)])
@pytest.mark.parametrize("bits", [4, 8, 16])
def test_fake_quant_none_range_path_bit_exact(oracle, shape, bits):
    """min/max=None: global fp32 statistics and fp32 scale (utils/quantize.py:26-37,66)."""
    from data_free_quantization_b200.utils.quantize import quantize
    rng = np.random.default_rng(sum(shape) + bits)
    x = (rng.standard_normal(shape) * 0.05).astype(np.float32)
    want = oracle.fake_quant(x, bits)
    got = quantize(_t(x), bits)
    assert got.cpu().numpy().tobytes() == want.tobytes()


def test_fake_quant_inplace_codes_and_unaligned(oracle):
    from data_free_quantization_b200.utils.quantize import quantize, quantize_codes
    rng = np.random.default_rng(5)
    x = (rng.standard_normal(10007) * 0.1).astype(np.float32)
    base = _t(np.concatenate([np.zeros(1, np.float32), x]))
    view = base[1:]                                             # 4-byte aligned only
    lo, hi = float(x.min()), float(x.max())
    want, wcodes = oracle.fake_quant(x, 8, lo, hi, want_codes=True)
    out = quantize(view, 8, lo, hi, inplace=True)
    assert out.data_ptr() == view.data_ptr()
    assert view.cpu().numpy().tobytes() == want.tobytes()
    y, codes, (scale, mn) = quantize_codes(_t(x), 8)
    assert y.cpu().numpy().tobytes() == want.tobytes()
    assert np.array_equal(codes.cpu().numpy().astype(np.int32), wcodes)
    # codes dequantise to the same values: q*scale + min
    deq = codes.cpu().numpy().astype(np.float32) * np.float32(scale) + np.float32(mn)
    assert deq.tobytes() == want.tobytes()


def test_fake_quant_empty_and_constant(oracle):
    from data_free_quantization_b200.utils.quantize import quantize
    assert quantize(torch.empty(0, device=DEV), 8, 0.0, 1.0).numel() == 0
    x = np.full(1000, 0.25, np.float32)
    assert quantize(_t(x), 8, 0.25, 0.25).cpu().numpy().tobytes() == oracle.fake_quant(x, 8, 0.25, 0.25).tobytes()
    assert quantize(_t(x), 8).cpu().numpy().tobytes() == oracle.fake_quant(x, 8).tobytes()


def test_minmax_and_rows(oracle):
    from data_free_quantization_b200.utils.quantize import minmax, minmax_rows, sample_minmax_mean
    rng = np.random.default_rng(11)
    for n in (1, 33, 4096, 100003, 3 << 20):
        x = rng.standard_normal(n).astype(np.float32)
        assert tuple(minmax(_t(x)).tolist()) == oracle.minmax(x)
    for rows, cols in ((8, 37), (16, 1024), (5, 1025), (3, 40000), (1000, 1280)):
        x = rng.standard_normal((rows, cols)).astype(np.float32)
        got = minmax_rows(_t(x), rows).cpu().numpy()
        assert np.array_equal(got[:, 0], x.min(1)) and np.array_equal(got[:, 1], x.max(1))
    x = rng.standard_normal((6, 3, 20, 20)).astype(np.float32)
    lo, hi = oracle.sample_minmax_mean(x)
    got = sample_minmax_mean(_t(x)).tolist()
    assert abs(got[0] - lo) <= 1e-6 * abs(lo) and abs(got[1] - hi) <= 1e-6 * abs(hi)


@pytest.mark.parametrize("bits", [4, 8])
def test_per_channel_quantisation(oracle, bits):
    from data_free_quantization_b200 import _lib as L
    from data_free_quantization_b200.utils.quantize import fake_quant_tensors
    rng = np.random.default_rng(3)
    for shape in ((64, 27), (96, 1, 3, 3), (10, 5000), (1000, 1280)):
        x = (rng.standard_normal(shape) * rng.uniform(0.01, 1.0, (shape[0],) + (1,) * (len(shape) - 1))).astype(np.float32)
        want, wcodes = oracle.quantize_per_channel(x, bits, want_codes=True)
        t = _t(x)
        codes = torch.empty(shape, dtype=torch.uint8, device=DEV)
        fake_quant_tensors([t], bits, L.Q_PER_CHANNEL, codes=[codes], rows=shape[0])
        assert t.cpu().numpy().tobytes() == want.tobytes()
        assert np.array_equal(codes.cpu().numpy().astype(np.int32), wcodes)


def test_clip_and_merge_bn(oracle):
    from data_free_quantization_b200 import _lib as L
    rng = np.random.default_rng(9)
    w = (rng.standard_normal(100001) * 10).astype(np.float32)
    want = w.copy()
    oracle.clip(want, -15.0, 15.0)
    t = _t(w)
    L.check(L.lib().dfq_clip_f32(t.data_ptr(), t.numel(), -15.0, 15.0, L.stream()))
    assert t.cpu().numpy().tobytes() == want.tobytes()
    O, K = 96, 27
    W = rng.standard_normal((O, K)).astype(np.float32)
    b = rng.standard_normal(O).astype(np.float32)
    g, be = (rng.random(O) + 0.5).astype(np.float32), rng.standard_normal(O).astype(np.float32)
    mu, var = rng.standard_normal(O).astype(np.float32), (rng.random(O) + 0.5).astype(np.float32)
    Wd, bd = _t(W), _t(b)
    fw, fb = torch.empty(O, device=DEV), torch.empty(O, device=DEV)
    gd, bed, mud, vard = _t(g), _t(be), _t(mu), _t(var)
    L.check(L.lib().dfq_merge_bn(Wd.data_ptr(), bd.data_ptr(), O, K, gd.data_ptr(), bed.data_ptr(),
                                 mud.data_ptr(), vard.data_ptr(), 1e-5, fw.data_ptr(), fb.data_ptr(), L.stream()))
    oracle.merge_bn(W, b, g, be, mu, var, 1e-5)
    assert Wd.cpu().numpy().tobytes() == W.tobytes() and bd.cpu().numpy().tobytes() == b.tobytes()
    assert np.array_equal(fw.cpu().numpy(), np.abs(g)) and np.array_equal(fb.cpu().numpy(), be)


@pytest.mark.parametrize("bits,signed", [(8, False), (4, False), (8, True)])
def test_bias_correction_eps_dot(oracle, bits, signed):
    import torch.nn as nn
    from data_free_quantization_b200.bias_correction import eps_dot
    rng = np.random.default_rng(21)
    layers, vecs, want = [], [], []
    for shape in ((64, 32, 1, 1), (32, 16, 3, 3), (1000, 1280), (24, 1, 3, 3), (7, 5000)):
        W = (rng.standard_normal(shape) * 0.1).astype(np.float32)
        v = rng.standard_normal(shape[1]).astype(np.float32)
        m = nn.Linear(2, 2) if len(shape) == 2 else nn.Conv2d(2, 2, 1)
        m.weight = nn.Parameter(_t(W), requires_grad=False)
        layers.append(m)
        vecs.append(torch.from_numpy(v))
        want.append(oracle.bias_correct(W, v, bits, signed))
    outs = eps_dot(layers, vecs, bits, signed)
    for o, w in zip(outs, want):
        assert rel_err(w, o.cpu().numpy()) <= 1e-5
    # v=None: row mean of the window-summed error
    outs = eps_dot(layers, [None] * len(layers), bits, signed)
    for m, o in zip(layers, outs):
        W = m.weight.detach().cpu().numpy()
        _, eps = oracle.bias_correct(W, np.ones(W.shape[1], np.float32), bits, signed, want_eps=True)
        assert rel_err(eps.astype(np.float64).mean(1).astype(np.float32), o.cpu().numpy()) <= 1e-5


@pytest.mark.parametrize("update_stat", [False, True, 8])
def test_quant_measure_eval_device_range(oracle, update_stat):
    """QuantMeasure.forward in eval mode (utils/quantize.py:106-123): running-range update with Python max/min
    semantics, then quantize() with the Python-float range -- all on the device."""
    from data_free_quantization_b200.utils.quantize import QuantMeasure
    rng = np.random.default_rng(17)
    qm = QuantMeasure(update_stat).to(DEV).eval()
    qm.running_min.fill_(-0.5)
    qm.running_max.fill_(0.75)
    rmin, rmax = np.float32(-0.5), np.float32(0.75)
    for step, scale in enumerate((0.1, 1.0, 0.3)):
        x = (rng.standard_normal((6, 3, 17, 19)) * scale).astype(np.float32)
        if update_stat:
            lo, hi = oracle.sample_minmax_mean(x)
            rmax = np.float32(hi) if np.float32(hi) > rmax else rmax
            rmin = np.float32(lo) if np.float32(lo) < rmin else rmin
        want = oracle.fake_quant(x, 8, float(rmin), float(rmax))
        got = qm(_t(x))
        assert float(qm.running_min) == float(rmin) and float(qm.running_max) == float(rmax), step
        assert got.cpu().numpy().tobytes() == want.tobytes(), step
