"""GPU: QConv2d / QLinear (utils/quantize.py:129-210, 260-323): the `scale` / `scale_prev` forms on top of the quantised
layers.  The forward with pending scales must equal (a) the restatement of the reference's forward built from the
library's own quantiser and (b) the forward after merge_scale_to_weight() folded the scales into the stored tensors."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _restate(layer, x, conv):
    from data_free_quantization_b200.utils.quantize import quantize
    w, b = layer.weight.detach(), None if layer.bias is None else layer.bias.detach()
    sp, sc = getattr(layer, "scale_prev", None), getattr(layer, "scale", None)
    if sp is not None:
        if conv:
            step, step_s = w.shape[0] // layer.groups, w.shape[1]
            spv = sp[:, 0, 0, 0].view(1, -1, 1, 1)
            w = torch.cat([w[g * step:(g + 1) * step] / spv[:, g * step_s:(g + 1) * step_s]
                           for g in range(layer.groups)], 0)
        else:
            w = w * sp.view(1, -1)
    if sc is not None:
        w = w * sc.detach()
        b = None if b is None else b * sc.detach().view(-1)
    w = w.contiguous()
    qw = quantize(w, layer.num_bits, float(w.min()), float(w.max()))
    qb = None if b is None else quantize(b.contiguous(), num_bits=layer.num_bits_bias)
    xq = layer.quant(x)
    if conv:
        with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
            return F.conv2d(xq, qw, qb, layer.stride, layer.padding, layer.dilation, layer.groups)
    return F.linear(xq, qw, qb)


@pytest.mark.parametrize("groups", [1, 2])
def test_qconv2d_scale_forms(groups):
    from data_free_quantization_b200.utils import quantize as Q
    torch.manual_seed(3)
    layer = Q.QConv2d(8, 12, 3, padding=1, groups=groups, num_bits=6, num_bits_act=8, num_bits_bias=12).to(DEV).eval()
    assert Q.QConv2d is not Q.QuantConv2d
    x = torch.randn(2, 8, 9, 9, device=DEV)
    with torch.no_grad():
        base = layer(x)
        scale = torch.rand(12, 1, 1, 1, device=DEV) + 0.5
        scale_prev = torch.rand(8, 1, 1, 1, device=DEV) + 0.5
        layer.set_scale(scale, scale_prev)
        assert isinstance(layer.scale, torch.nn.Parameter) and tuple(layer.scale.shape) == (12, 1, 1, 1)
        want = _restate(layer, x, True)
        got = layer(x)
        assert torch.equal(got, want)
        assert not torch.equal(got, base)
        layer.merge_scale_to_weight()
        assert layer.scale is None and layer.scale_prev is None
        assert torch.equal(layer(x), want)


def test_qlinear_scale_forms_and_straight_through_gradient():
    from data_free_quantization_b200.utils import quantize as Q
    torch.manual_seed(4)
    layer = Q.QLinear(16, 10, num_bits=8, num_bits_act=8, num_bits_bias=16).to(DEV).eval()
    x = torch.randn(5, 16, device=DEV)
    scale, scale_prev = torch.rand(10, device=DEV) + 0.5, torch.rand(16, device=DEV) + 0.5
    layer.set_scale(scale, scale_prev)
    assert tuple(layer.scale.shape) == (10, 1)
    with torch.no_grad():
        want = _restate(layer, x, False)
        assert torch.equal(layer(x), want)
    # the scale is a Parameter: the weight quantiser passes gradients straight through (utils/quantize.py:81-85)
    layer(x).sum().backward()
    assert layer.scale.grad is not None and float(layer.scale.grad.abs().sum()) > 0
    with torch.no_grad():
        layer.merge_scale_to_weight()
        assert torch.equal(layer(x), want)
