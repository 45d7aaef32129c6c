"""GPU: dense k x k convolutions on the tcgen05 pipeline (csrc/pw_i8.cu kConv, dfq_conv_i8) -- DeepLab's atrous and
decoder shapes plus ragged ones -- against (a) an exact restatement: integer codes from the oracle's fake-quant
recipe, the reduction as an exact integer convolution in float64, the affine terms in float64, zero padding as an
exact 0.0 (the reference quantises, then F.conv2d pads); and (b) the reference's own formulation, F.conv2d in fp32
on the fake-quantised tensors (utils/quantize.py:250-256).  Tolerances as for the pointwise kernel: 2e-6 of the
largest output against (a), 2e-5 against (b)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from test_gpu_eval_kernels import _decode_stats, _slot

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _case(oracle, N, C, O, H, W, k=3, stride=1, pad=1, dil=1, relu=True, bias=True, wbits=8, abits=8, seed=0,
          once=False, o_cap=0):
    from data_free_quantization_b200 import pw_i8
    rs = np.random.default_rng(seed)
    x = (rs.standard_normal((N, C, H, W)) * 1.5 + 0.3).astype(np.float32)
    lo, hi = float(np.float32(-2.1)), float(np.float32(3.4))
    w0 = (rs.standard_normal((O, C, k, k)) * 0.2).astype(np.float32)
    w = oracle.fake_quant(w0, wbits, float(w0.min()), float(w0.max()))
    b = (rs.standard_normal(O) * 0.5).astype(np.float32) if bias else None
    xq = oracle.fake_quant(x, abits, lo, hi)
    tb = None if b is None else torch.from_numpy(b)
    want32 = F.conv2d(torch.from_numpy(xq), torch.from_numpy(w), tb, stride, pad, dil)
    exact = F.conv2d(torch.from_numpy(xq).double(), torch.from_numpy(w).double(),
                     None if tb is None else tb.double(), stride, pad, dil).numpy()
    if relu:
        want32, exact = torch.relu(want32), np.maximum(exact, 0)
    pk = pw_i8.pack_weight_conv(torch.from_numpy(w).to(DEV), wbits, o_cap=o_cap)
    assert pk is not None, "quantised weight not recognised as on-grid"
    xd = torch.from_numpy(x).to(DEV)
    y = torch.full(tuple(want32.shape), float("nan"), device=DEV)
    stats = torch.zeros(N, 2, dtype=torch.int32, device=DEV)
    slot = _slot(lo, hi, abits)
    bd = None if b is None else torch.from_numpy(b).to(DEV)
    scratch = pw_i8.conv_scratch(xd.shape, DEV) if once else None     # quantise-once form (s8 NHWC codes)
    d = pw_i8.make_conv_desc(xd, y, pk, bd, slot.data_ptr(), stats.data_ptr(), stride, pad, dil, relu, abits,
                             scratch=scratch)
    pw_i8.launch_conv(d)
    torch.cuda.synchronize()
    got = y.cpu().numpy()
    assert not np.isnan(got).any()
    scale = np.abs(exact).max()
    assert np.abs(got - exact).max() / scale <= 2e-6, ("exact", np.abs(got - exact).max() / scale)
    assert np.abs(got - want32.numpy()).max() / scale <= 2e-5
    mn, mx = _decode_stats(stats)
    g = got.reshape(N, -1)
    assert np.array_equal(mn, g.min(1)) and np.array_equal(mx, g.max(1))


@pytest.mark.parametrize("N,C,O,H,W,k,stride,pad,dil", [
    (2, 16, 32, 12, 12, 3, 1, 1, 1),        # one channel chunk per tap, ragged pixel tile
    (1, 320, 256, 33, 33, 3, 1, 6, 6),      # ASPP branch 2 (aspp.py:50-53)
    (1, 320, 256, 33, 33, 3, 1, 18, 18),    # ASPP branch 4: most taps of most pixels fall in the padding
    (1, 304, 256, 33, 33, 3, 1, 1, 1),      # decoder.last_conv.0 (C not a multiple of 16 -> 19 chunks per tap)
    (2, 24, 40, 17, 15, 3, 2, 1, 1),        # stride 2, odd sizes, N not a multiple of 16
    (2, 8, 16, 9, 11, 5, 1, 2, 1),          # 5 x 5
    (3, 3, 32, 32, 32, 3, 2, 1, 1),         # stem-shaped: 3 channels padded to 16
    (1, 40, 272, 10, 10, 3, 1, 0, 1),       # no padding, two output passes
])
@pytest.mark.parametrize("once", [False, True])
def test_dense_conv(oracle, N, C, O, H, W, k, stride, pad, dil, once):
    _case(oracle, N, C, O, H, W, k, stride, pad, dil, once=once)


@pytest.mark.parametrize("o_cap", [32, 96, 128])
def test_dense_conv_narrow_passes(oracle, o_cap):
    """ASPP branch with fewer output channels per pass (137 pixel tiles at 16 x 33 x 33 do not fill 148 SMs)."""
    _case(oracle, 1, 320, 256, 33, 33, 3, 1, 12, 12, once=True, o_cap=o_cap)
    _case(oracle, 2, 16, 40, 12, 12, 3, 1, 1, 1, once=False, o_cap=o_cap, seed=o_cap)


def test_no_relu_no_bias_low_bits(oracle):
    _case(oracle, 2, 32, 48, 14, 14, relu=False, bias=False, wbits=4, abits=8, seed=3)
    _case(oracle, 1, 32, 48, 14, 14, wbits=6, abits=6, seed=4)
    _case(oracle, 2, 32, 48, 14, 14, relu=False, bias=False, wbits=4, abits=8, seed=3, once=True)
    _case(oracle, 1, 32, 48, 14, 14, wbits=6, abits=6, seed=4, once=True)
