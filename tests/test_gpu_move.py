"""GPU: the one-pass data-movement kernels of the quantised eval forward (csrc/move.cu) against the op sequence the
reference runs -- QuantMeasure's fake-quant (the oracle's recipe, utils/quantize.py:72-76) followed by the torch
op: F.interpolate(bilinear) (utils/layer_transform.py:98-110), torch.cat(dim=1) (:113-125) and F.pad
(modeling/segmentation/backbone/mobilenet.py:16-22).  cat and pad are pure data movement: bit-exact.  Bilinear
interpolation is four products and three sums per output whose contraction into FMAs differs between ATen's CPU
and CUDA builds: tolerance 1e-6 of the operand's range.  Per-sample statistics exact for the values written."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from test_gpu_eval_kernels import _decode_stats, _slot

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _check_stats(stats, got):
    lo, hi = _decode_stats(stats)
    g = got.reshape(got.shape[0], -1).numpy()
    assert np.array_equal(lo, g.min(1)) and np.array_equal(hi, g.max(1))


@pytest.mark.parametrize("shape,size,align", [
    ((2, 5, 33, 33), (129, 129), True),        # decoder.py:39 (4x up, three source columns per 4 outputs)
    ((2, 21, 17, 17), (65, 65), True),         # deeplab.py:32
    ((3, 4, 1, 1), (33, 33), True),            # aspp.py:71: the pooled branch broadcast
    ((2, 3, 9, 7), (20, 31), True),
    ((2, 3, 9, 7), (20, 31), False),
    ((1, 2, 20, 24), (7, 9), True),            # downsampling: the per-element path
    ((2, 3, 8, 8), (16, 16), False),
    ((1, 1, 5, 6), (5, 6), True),
])
@pytest.mark.parametrize("quant", [False, True])
def test_qinterp_bilinear(oracle, shape, size, align, quant):
    from data_free_quantization_b200 import _lib as L
    rng = np.random.default_rng(sum(shape) + size[0])
    x = (rng.standard_normal(shape) * 1.5).astype(np.float32)
    lo_hi = (-2.0, 3.0)
    xq = oracle.fake_quant(x, 8, *lo_hi) if quant else x
    want = F.interpolate(torch.from_numpy(xq), size=size, mode="bilinear", align_corners=align)
    N, C, H, W = shape
    xd = torch.from_numpy(x).to(DEV)
    # destination at every 16-byte phase: the flat-run driver must cope with unaligned sample starts
    for phase in range(4):
        buf = torch.full((N * C * size[0] * size[1] + 8,), float("nan"), device=DEV)
        y = buf[phase:phase + N * C * size[0] * size[1]].view(N, C, *size)
        stats = torch.zeros(N, 2, dtype=torch.int32, device=DEV)
        slot = _slot(*lo_hi) if quant else None
        L.check(L.lib().dfq_qinterp_bilinear_f32(xd.data_ptr(), y.data_ptr(), N, C, H, W, size[0], size[1], int(align),
                                                 None if slot is None else slot.data_ptr(), 8, stats.data_ptr(),
                                                 L.stream()), "dfq_qinterp_bilinear_f32")
        got = y.cpu()
        assert torch.isnan(buf[:phase]).all() and torch.isnan(buf[phase + y.numel():]).all()
        err = float((got - want).abs().max())
        assert err <= 1e-6 * 5.0, (phase, err)
        _check_stats(stats, got)
        # the same arithmetic on the GPU build of ATen: also within the tolerance
        ref = F.interpolate(torch.from_numpy(xq).to(DEV), size=size, mode="bilinear", align_corners=align).cpu()
        assert float((got - ref).abs().max()) <= 1e-6 * 5.0


@pytest.mark.parametrize("hw", [(33, 33), (5, 7), (16, 16)])
def test_qcopy_as_cat(oracle, hw):
    """torch.cat((Q1(a), Q2(b), c), dim=1): one dfq_qcopy_f32 launch per operand into the channel slice."""
    from data_free_quantization_b200 import _lib as L
    rng = np.random.default_rng(hw[0])
    N = 3
    chans = [5, 3, 2]
    ranges = [(-1.0, 2.0), (0.0, 4.0), None]
    xs = [(rng.standard_normal((N, c) + hw) * 2).astype(np.float32) for c in chans]
    want = np.concatenate([x if r is None else oracle.fake_quant(x, 8, *r) for x, r in zip(xs, ranges)], axis=1)
    Ct, HW = sum(chans), hw[0] * hw[1]
    y = torch.full((N, Ct) + hw, float("nan"), device=DEV)
    stats = torch.zeros(N, 2, dtype=torch.int32, device=DEV)
    c0 = 0
    keep = []
    for x, c, r in zip(xs, chans, ranges):
        xd = torch.from_numpy(x).to(DEV)
        slot = None if r is None else _slot(*r)
        keep += [xd, slot]
        L.check(L.lib().dfq_qcopy_f32(xd.data_ptr(), y.data_ptr() + 4 * c0 * HW, N, c * HW, c * HW, Ct * HW,
                                      None if slot is None else slot.data_ptr(), 8, stats.data_ptr(), L.stream()),
                "dfq_qcopy_f32")
        c0 += c
    got = y.cpu()
    assert np.array_equal(got.numpy(), want)
    _check_stats(stats, got)


def test_qcopy_large_and_empty(oracle):
    from data_free_quantization_b200 import _lib as L
    x = torch.randn(2, 3 * 4096 + 5, device=DEV)
    y = torch.empty_like(x)
    slot = _slot(-1.5, 1.5)
    L.check(L.lib().dfq_qcopy_f32(x.data_ptr(), y.data_ptr(), 2, x.shape[1], x.shape[1], x.shape[1], slot.data_ptr(), 8,
                                  None, L.stream()), "dfq_qcopy_f32")
    assert np.array_equal(y.cpu().numpy(), oracle.fake_quant(x.cpu().numpy(), 8, -1.5, 1.5))
    assert L.lib().dfq_qcopy_f32(x.data_ptr(), y.data_ptr(), 0, 0, 0, 0, None, 8, None, L.stream()) == 0


@pytest.mark.parametrize("shape,pad", [((2, 3, 9, 9), (1, 1, 1, 1)), ((2, 4, 17, 13), (2, 2, 2, 2)),
                                       ((1, 5, 8, 8), (0, 1, 0, 1)), ((3, 2, 33, 33), (4, 4, 4, 4)),
                                       ((2, 2, 6, 5), (3, 0, 0, 2)), ((1, 1, 4, 4), (0, 0, 0, 0))])
def test_pad2d(shape, pad):
    from data_free_quantization_b200 import _lib as L
    x = torch.randn(shape, device=DEV) - 2.0          # all-negative interiors: the zero border must reach the maximum
    want = F.pad(x, pad)
    N, C, H, W = shape
    y = torch.full(want.shape, float("nan"), device=DEV)
    stats = torch.zeros(N, 2, dtype=torch.int32, device=DEV)
    L.check(L.lib().dfq_pad2d_f32(x.data_ptr(), y.data_ptr(), N, C, H, W, *pad, stats.data_ptr(), L.stream()),
            "dfq_pad2d_f32")
    assert torch.equal(y, want)
    _check_stats(stats, y.cpu())
    assert L.lib().dfq_pad2d_f32(x.data_ptr(), y.data_ptr(), N, C, H, W, -1, 0, 0, 0, None, L.stream()) != 0
