"""Host mirror of the reference's utils/quantize.py on top of libdfq_b200 (sm_100a).

Same names, signatures and error behaviour as the reference module; the arithmetic runs in the
hand-written CUDA kernels of csrc/quant.cu and csrc/act.cu through the C ABI (include/dfq_b200.h).
Inputs must be contiguous fp32 CUDA tensors: there is no CPU path.
"""
import ctypes

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.autograd.function import InplaceFunction

from .. import _lib as L


def _params(min_value, max_value, num_bits, flags):
    p = L.DfqQuantParams()
    L.lib().dfq_quant_params_host(float(min_value), float(max_value), int(num_bits), int(flags), ctypes.byref(p))
    return p


def _apply(x, out, num_bits, params, codes=None):
    n = x.numel()
    if n == 0:
        return out
    L.check(L.lib().dfq_fake_quant_apply_f32(L.dev_f32(x, "input"), L.dev_f32(out, "output"),
                                             None if codes is None else codes.data_ptr(), n, int(num_bits),
                                             ctypes.byref(params), L.stream()), "dfq_fake_quant_apply_f32")
    return out


def minmax(x):
    """[min, max] of x as a 2-element fp32 CUDA tensor (float(W.min()), float(W.max()))."""
    out = torch.empty(2, dtype=torch.float32, device=x.device)
    L.check(L.lib().dfq_minmax_f32(L.dev_f32(x, "input"), x.numel(), out.data_ptr(), L.stream()), "dfq_minmax_f32")
    return out


def minmax_rows(x, rows):
    """Per-row [min, max] of x.view(rows, -1) -> [rows, 2] fp32 CUDA tensor."""
    out = torch.empty(rows, 2, dtype=torch.float32, device=x.device)
    L.check(L.lib().dfq_minmax_rows_f32(L.dev_f32(x, "input"), rows, x.numel() // rows, out.data_ptr(), L.stream()),
            "dfq_minmax_rows_f32")
    return out


def fake_quant_tensors(tensors, bits, flags=0, outs=None, codes=None, clip=None, rows=None):
    """One multi-tensor launch (range reduce + apply) over a list of tensors.

    bits / flags / rows may be scalars or per-tensor lists.  Returns the [slots, 2] fp32 ranges tensor
    (min, max per tensor, or per row for DFQ_Q_PER_CHANNEL).  utils/layer_transform.py:288-305.
    """
    T = len(tensors)
    if T == 0:
        return None
    as_list = lambda v: list(v) if isinstance(v, (list, tuple)) else [v] * T
    bits, flags, rows = as_list(bits), as_list(flags), as_list(rows if rows is not None else 1)
    dev = tensors[0].device
    # The descriptor table only holds addresses, sizes and flags: it is planned and uploaded once per distinct call
    # signature and reused (a 1 GiB tensor spent 0.1 of its 0.6 ms waiting for the host to rebuild and upload it).
    key = (dev.index, tuple(int(b) for b in bits), tuple(int(f) for f in flags), tuple(int(r) for r in rows),
           None if clip is None else (float(clip[0]), float(clip[1])),
           tuple(t.data_ptr() for t in tensors), tuple(t.numel() for t in tensors),
           None if outs is None else tuple(o.data_ptr() for o in outs),
           None if codes is None else tuple(0 if c is None else c.data_ptr() for c in codes))
    hit = None if L._capture is not None else _table_cache.get(key)   # a captured graph keeps its own tables alive
    if hit is None:
        descs = (L.DfqQuantDesc * T)()
        for i, t in enumerate(tensors):
            d = descs[i]
            d.x = L.dev_f32(t, f"tensor[{i}]")
            d.y = L.dev_f32(outs[i] if outs is not None else t, f"out[{i}]")
            d.codes = codes[i].data_ptr() if codes is not None and codes[i] is not None else None
            d.n, d.rows, d.bits = t.numel(), int(rows[i]), int(bits[i])
            d.flags = int(flags[i]) | (L.Q_CLAMP_OUT if clip is not None else 0)
            if clip is not None:
                d.clip_lo, d.clip_hi = float(clip[0]), float(clip[1])
        tb, ts = L.i64(), L.i64()
        L.check(L.lib().dfq_quant_plan_host(descs, T, ctypes.byref(tb), ctypes.byref(ts)), "dfq_quant_plan_host")
        table = L.table_to_device(descs, dev)
        if L._capture is None:                      # (tables uploaded during a graph capture live in the plan's arena)
            if len(_table_cache) >= 64:
                _table_cache.clear()
            _table_cache[key] = (table, tb.value, ts.value)
        hit = (table, tb.value, ts.value)
    else:
        L.bump_weights_epoch()                      # a weight-side pass is about to run (see L.weights_epoch)
    table, tbv, tsv = hit
    ranges = torch.empty(tsv, 4, dtype=torch.float32, device=dev)   # {min, max, scale, 0} per slot
    L.check(L.lib().dfq_fake_quant_multi_f32(table.data_ptr(), T, tbv, tsv, ranges.data_ptr(), L.stream()),
            "dfq_fake_quant_multi_f32")
    return ranges


_table_cache = {}


class UniformQuantize(InplaceFunction):
    """utils/quantize.py:16-85 -- uniform fake-quantisation, straight-through backward."""

    @staticmethod
    def forward(ctx, input, num_bits=8, min_value=None, max_value=None, inplace=False, symmetric=False,
                num_chunks=None):
        x = input.detach()
        L.dev_f32(x, "input")
        ctx.inplace, ctx.num_bits = inplace, num_bits
        if inplace:
            ctx.mark_dirty(input)
            output = input
        else:
            output = torch.empty_like(x)
        if x.numel() == 0:
            return output
        flags = L.Q_SYMMETRIC if symmetric else 0
        if min_value is not None and max_value is not None:
            # Python-float range: scale formed in double (utils/quantize.py:51-68)
            return _apply(x, output, num_bits, _params(min_value, max_value, num_bits, flags))
        # min/max=None: 0-dim fp32 tensor statistics, scale formed in fp32 (:26-37)
        B = x.shape[0]
        num_chunks = B if num_chunks is None else num_chunks
        rows = B // num_chunks
        if rows == 1 and min_value is None and max_value is None:
            fake_quant_tensors([x], num_bits, flags | L.Q_RANGE_F32, outs=[output])
            return output
        mm = minmax_rows(x, rows).mean(0)                   # y.min(-1)[0].mean(-1), y.max(-1)[0].mean(-1)
        lo, hi = (float(v) for v in mm.tolist())
        lo = lo if min_value is None else float(min_value)
        hi = hi if max_value is None else float(max_value)
        return _apply(x, output, num_bits, _params(lo, hi, num_bits, flags | L.Q_RANGE_F32))

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output, None, None, None, None, None, None


def quantize(x, num_bits=8, min_value=None, max_value=None, inplace=False, symmetric=False, num_chunks=None):
    """utils/quantize.py:88-89."""
    return UniformQuantize.apply(x, num_bits, min_value, max_value, inplace, symmetric, num_chunks)


def quantize_codes(x, num_bits=8, min_value=None, max_value=None, symmetric=False):
    """Fake-quantise and also return the integer codes q - qmin (uint8 for <=8 bits, else int16 storage).

    Returns (dequantised fp32, codes, (scale, min)) -- the packed form SURVEY.md 8(f).4 asks for.
    """
    L.dev_f32(x, "input")
    flags = L.Q_SYMMETRIC if symmetric else 0
    if min_value is None or max_value is None:
        lo, hi = (float(v) for v in minmax(x).tolist())
        min_value = lo if min_value is None else min_value
        max_value = hi if max_value is None else max_value
    p = _params(min_value, max_value, num_bits, flags)
    out = torch.empty_like(x)
    codes = torch.empty(x.shape, dtype=torch.uint8 if num_bits <= 8 else torch.int16, device=x.device)
    _apply(x, out, num_bits, p, codes)
    return out, codes, (p.scale, p.add_back)


def sample_minmax_mean(x):
    """(mean_n min_chw x, mean_n max_chw x) as a [2] fp32 CUDA tensor -- utils/quantize.py:110-111."""
    n = x.size(0)
    return minmax_rows(x.detach(), n).mean(0)


class QuantMeasure(nn.Module):
    """utils/quantize.py:94-126 -- activation range tracker + fake-quant of the activation."""

    def __init__(self, update_stat=False, num_bits=8, momentum=0.1, *, _storage=None):
        """`_storage` (internal): a zeroed [2] tensor whose halves become running_min / running_max, so that a whole
        model's measures can live in one device allocation (set_layer_bits)."""
        super(QuantMeasure, self).__init__()
        if _storage is None:
            _storage = torch.zeros(2)
        self.register_buffer('running_min', _storage[0:1])
        self.register_buffer('running_max', _storage[1:2])
        self.momentum = momentum
        self.num_bits = num_bits
        self.update_stat = update_stat

    def forward(self, input):
        if not self.training:
            return self._forward_eval(input)
        mm = sample_minmax_mean(input)
        if self.update_stat:
            # Python max()/min() on tensors: the running value is kept unless strictly improved (:110-111)
            self.running_max = torch.where(mm[1] > self.running_max, mm[1], self.running_max)
            self.running_min = torch.where(mm[0] < self.running_min, mm[0], self.running_min)
        min_value, max_value = mm[0], mm[1]
        self.running_min.mul_(1 - self.momentum).add_(min_value * self.momentum)
        self.running_max.mul_(1 - self.momentum).add_(max_value * self.momentum)
        return quantize(input, self.num_bits, min_value=float(min_value), max_value=float(max_value), num_chunks=16)

    def measure(self, input):
        """Eval mode (:107-111): per-sample extrema, their batch mean, the running-range update and the scale, all
        on the device (dfq_measure_update; no float() round trip).  Returns the range slot {min, max, scale, 0} a
        consumer kernel quantises with."""
        x = input.detach()
        L.dev_f32(x, "input")
        n = x.size(0)
        slot = torch.empty(4, dtype=torch.float32, device=x.device)
        stats = torch.empty(2 * n, dtype=torch.float32, device=x.device) if self.update_stat else None
        L.check(L.lib().dfq_measure_update(
            x.data_ptr(), n, x.numel() // n, L.dev_f32(self.running_min, "running_min"),
            L.dev_f32(self.running_max, "running_max"), slot.data_ptr(), int(self.num_bits), 0,
            1 if self.update_stat else 0, None if stats is None else stats.data_ptr(), L.stream()),
            "dfq_measure_update")
        return slot

    def _forward_eval(self, input):
        """Eval mode (:107-111, :119-123) with the range resident on the device."""
        x = input.detach()
        L.dev_f32(x, "input")
        out = torch.empty_like(x)
        if x.numel() == 0:
            return out
        slot = self.measure(x)
        L.check(L.lib().dfq_fake_quant_apply_slot_f32(x.data_ptr(), out.data_ptr(), None, x.numel(),
                                                      int(self.num_bits), 0, slot.data_ptr(), L.stream()),
                "dfq_fake_quant_apply_slot_f32")
        return out

    def set_update_stat(self, update_stat):
        self.update_stat = update_stat


def _conv2d_fp32(input, weight, bias, stride, padding, dilation, groups):
    """F.conv2d in true fp32: cuDNN would otherwise pick TF32 tensor-core kernels (10-bit mantissa), which breaks
    the 1e-3 logits tolerance against the reference's CPU fp32 convolution."""
    with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        return F.conv2d(input, weight, bias, stride, padding, dilation, groups)


def _quant_weight(weight, num_bits):
    """quantize(W, bits, float(W.min()), float(W.max())) without the two host round trips.  When a gradient is being
    recorded for the weight the call goes through UniformQuantize (straight-through estimator, utils/quantize.py:81-85)
    with the range read back to the host as the reference does."""
    if torch.is_grad_enabled() and weight.requires_grad:
        lo, hi = (float(v) for v in minmax(weight.detach()).tolist())
        return quantize(weight, num_bits, lo, hi)
    out = torch.empty_like(weight)
    fake_quant_tensors([weight.detach()], num_bits, 0, outs=[out])
    return out


# ---- fused eval forward of the QuantN* layers: quantise-on-load kernels instead of quant() + F.conv2d ------------
# QuantNConv2d / QuantNLinear.forward (utils/quantize.py:250-256, 358-363) are x_hat = quant(x); F.conv2d(x_hat, W, b).
# In eval mode on CUDA, with W on an integer grid of <= 8 bits (what quantize_targ_layer leaves), the same values
# come from ONE kernel that reads x, quantises it on load with the QuantMeasure's device-resident range and runs the
# reduction as exact s8 x s8 -> s32 sums on tcgen05 (dense k x k: dfq_conv_i8; 1 x 1 / linear: dfq_pw_i8); depthwise
# 3 x 3 goes to the CUDA-core kernel (dfq_qconv2d_f32).  FUSED_EVAL = False restores the two-step form (quant on
# our kernel, arithmetic through cuDNN fp32) -- bench.py's `module_path` comparator.
FUSED_EVAL = True
fused_calls = {"conv_i8": 0, "pw_i8": 0, "depthwise": 0, "direct": 0, "linear_i8": 0}     # launches per kind (tests, tools)


def invalidate_packed(model=None):
    """Forget the packed integer codes cached on the layers (or everywhere, by moving the epoch).  Needed after a
    write to `weight.data` that neither goes through this package nor bumps `weight._version` (`.data` detaches the
    version counter): e.g. `layer.weight.data.mul_(2)`."""
    L.bump_weights_epoch()
    if model is not None:
        for m in model.modules():
            m.__dict__.pop("_dfq_packed", None)


def _packed(layer, conv):
    """Integer codes of the layer's weight packed for the tensor-core kernels, cached until any weight-side pass
    runs again (L.weights_epoch() moves with every descriptor-table upload, checkpoint unpack, sharded dequantise,
    pipeline replay) or the tensor is replaced.  In-place writes through `weight.data` from outside the package are
    invisible to the key: call invalidate_packed()."""
    from .. import pw_i8
    w = layer.weight
    key = (w.data_ptr(), w._version, tuple(w.shape), L.weights_epoch())
    hit = getattr(layer, "_dfq_packed", None)
    if hit is not None and hit[0] == key:
        return hit[1]
    pk = pw_i8.pack_weight_conv(w.detach()) if conv else pw_i8.pack_weight(w.detach())
    layer._dfq_packed = (key, pk)
    return pk


def _fused_conv_forward(layer, x):
    """QuantNConv2d.forward in one launch, or None when this layer / input is not covered."""
    from .. import pw_i8
    if not (FUSED_EVAL and x.is_cuda and not layer.quant.training and x.dim() == 4 and x.numel() > 0
            and x.dtype == torch.float32 and layer.padding_mode == "zeros" and isinstance(layer.padding, tuple)):
        return None
    kh, kw = layer.kernel_size
    (sh, sw), (ph, pw_), (dh, dw) = layer.stride, layer.padding, layer.dilation
    if sh != sw or ph != pw_ or dh != dw:
        return None
    N, C, H, W = x.shape
    OH = (H + 2 * ph - dh * (kh - 1) - 1) // sh + 1
    OW = (W + 2 * ph - dh * (kw - 1) - 1) // sh + 1
    O = layer.out_channels
    depthwise = layer.groups == C == O and kh == kw == 3
    if not depthwise and layer.groups != 1:
        return None
    # few input channels (the 3 -> 32 stem): the CUDA-core direct form beats an implicit GEMM whose 16-byte K chunks
    # would be mostly padding
    direct = layer.groups == 1 and kh * kw > 1 and C < 16 and C * kh * kw <= 64
    x = x.detach().contiguous()
    bias = None if layer.bias is None else layer.bias.detach()
    bits = int(layer.quant.num_bits)
    if not 1 <= bits <= 8 and not (depthwise or direct):
        return None                      # the s8 tensor-core kernels cover activation grids of <= 8 bits
    if depthwise or direct:
        slot = layer.quant.measure(x)
        y = torch.empty(N, O, OH, OW, dtype=torch.float32, device=x.device)
        d = L.DfqConvDesc()
        d.x, d.y, d.w = x.data_ptr(), y.data_ptr(), L.dev_f32(layer.weight.detach(), "weight")
        d.bias = None if bias is None else L.dev_f32(bias, "bias")
        d.in_slot, d.out_stats = slot.data_ptr(), None
        d.N, d.C_in, d.H, d.W, d.C_out, d.OH, d.OW = N, C, H, W, O, OH, OW
        d.kh, d.kw, d.stride, d.pad, d.dil, d.groups, d.relu, d.bits = kh, kw, sh, ph, dh, layer.groups, 0, bits
        L.check(L.lib().dfq_qconv2d_f32(ctypes.byref(d), L.stream()), "dfq_qconv2d_f32")
        fused_calls["depthwise" if depthwise else "direct"] += 1
        return y
    pointwise = kh == kw == 1 and sh == 1 and ph == 0
    pk = _packed(layer, conv=not pointwise)
    if pk is None:
        return None
    slot = layer.quant.measure(x)
    y = torch.empty(N, O, OH, OW, dtype=torch.float32, device=x.device)
    if pointwise:
        d = pw_i8.make_desc(x.data_ptr(), y.data_ptr(), pk, bias, slot.data_ptr(), None, N, H * W, False, bits)
        pw_i8.launch(d)
        fused_calls["pw_i8"] += 1
    else:
        scratch = pw_i8.conv_scratch(x.shape, x.device) if kh * kw > 1 else None
        d = pw_i8.make_conv_desc(x, y, pk, bias, slot.data_ptr(), None, sh, ph, dh, False, bits, scratch=scratch)
        pw_i8.launch_conv(d)
        fused_calls["conv_i8"] += 1
    return y


def _fused_linear_forward(layer, x):
    from .. import pw_i8
    if not (FUSED_EVAL and x.is_cuda and not layer.quant.training and x.dim() == 2 and x.numel() > 0
            and x.dtype == torch.float32):
        return None
    if not 1 <= int(layer.quant.num_bits) <= 8:
        return None                      # the s8 tensor-core kernel covers activation grids of <= 8 bits
    pk = _packed(layer, conv=False)
    if pk is None:
        return None
    x = x.detach().contiguous()
    slot = layer.quant.measure(x)
    y = torch.empty(x.shape[0], layer.out_features, dtype=torch.float32, device=x.device)
    bias = None if layer.bias is None else layer.bias.detach()
    d = pw_i8.make_desc(x.data_ptr(), y.data_ptr(), pk, bias, slot.data_ptr(), None, x.shape[0], 1, False,
                        int(layer.quant.num_bits), linear=True)
    pw_i8.launch(d)
    fused_calls["linear_i8"] += 1
    return y


class QuantConv2d(nn.Conv2d):
    """utils/quantize.py:213-238 -- activation, weight and bias re-quantised on every forward."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, bias=True,
                 num_bits=8, num_bits_act=8, num_bits_bias=16, momentum=0.1):
        super(QuantConv2d, self).__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups,
                                          bias)
        self.num_bits = num_bits
        self.num_bits_bias = num_bits_bias
        self.quant = QuantMeasure(num_bits=num_bits_act, momentum=momentum)

    def forward(self, input):
        input = self.quant(input)
        qweight = _quant_weight(self.weight, self.num_bits)
        qbias = quantize(self.bias, num_bits=self.num_bits_bias) if self.bias is not None else None
        return _conv2d_fp32(input, qweight, qbias, self.stride, self.padding, self.dilation, self.groups)


class QuantNConv2d(nn.Conv2d):
    """utils/quantize.py:240-256 -- only the activation is quantised (weights were quantised offline)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, bias=True,
                 num_bits=8, num_bits_act=8, momentum=0.1):
        super(QuantNConv2d, self).__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups,
                                           bias)
        self.quant = QuantMeasure(num_bits=num_bits_act, momentum=momentum)

    def forward(self, input):
        fused = _fused_conv_forward(self, input)
        if fused is not None:
            return fused
        input = self.quant(input)
        return _conv2d_fp32(input, self.weight, self.bias, self.stride, self.padding, self.dilation, self.groups)


class QuantLinear(nn.Linear):
    """utils/quantize.py:326-348."""

    def __init__(self, in_features, out_features, bias=True, num_bits=8, num_bits_act=8, num_bits_bias=16,
                 momentum=0.1):
        super(QuantLinear, self).__init__(in_features, out_features, bias)
        self.num_bits = num_bits
        self.num_bits_bias = num_bits_bias
        self.quant = QuantMeasure(num_bits=num_bits_act, momentum=momentum)

    def forward(self, input):
        input = self.quant(input)
        qweight = _quant_weight(self.weight, self.num_bits)
        qbias = quantize(self.bias, num_bits=self.num_bits_bias) if self.bias is not None else None
        return F.linear(input, qweight, qbias)


class QuantNLinear(nn.Linear):
    """utils/quantize.py:350-363."""

    def __init__(self, in_features, out_features, bias=True, num_bits=8, num_bits_act=8, momentum=0.1):
        super(QuantNLinear, self).__init__(in_features, out_features, bias)
        self.quant = QuantMeasure(num_bits=num_bits_act, momentum=momentum)

    def forward(self, input):
        fused = _fused_linear_forward(self, input)
        if fused is not None:
            return fused
        input = self.quant(input)
        return F.linear(input, self.weight, self.bias)


class _ScaleMixin:
    """set_scale / merge_scale_to_weight of the reference's QConv2d / QLinear (utils/quantize.py:142-166, 270-291): an
    output-channel `scale` (a Parameter) and the previous layer's `scale_prev`, applied to the weight (and bias) ahead of
    the per-forward weight quantisation, or folded into the stored weight once."""

    def set_scale(self, scale=None, scale_prev=None):
        if scale is not None:
            self.register_parameter("scale", nn.Parameter(scale.view(*self._scale_shape)))
        if scale_prev is not None:
            self.scale_prev = scale_prev

    def merge_scale_to_weight(self):
        if getattr(self, "scale_prev", None) is not None:
            self.weight.data.copy_(self.merge_scale_prev(self.weight.detach(), self.scale_prev))
            self.scale_prev = None
        if getattr(self, "scale", None) is not None:
            weight, bias = self.merge_scale(self.weight.detach(), self.bias.detach() if self.bias is not None else None,
                                            self.scale)
            self.weight.data.copy_(weight)
            if self.bias is not None:
                self.bias.data.copy_(bias)
            self.scale = None

    def merge_scale(self, weight, bias, scale):
        weight = weight * scale
        if bias is not None:
            bias = bias * scale.view(-1)
        return weight, bias

    def _scaled(self):
        """(weight, bias) with the pending scales applied -- utils/quantize.py:183-198, 300-311."""
        sweight, sbias = self.weight, self.bias
        if getattr(self, "scale_prev", None) is not None:
            sweight = self.merge_scale_prev(sweight, self.scale_prev)
        if getattr(self, "scale", None) is not None:
            sweight, sbias = self.merge_scale(sweight, sbias, self.scale)
        return sweight, sbias


class QConv2d(_ScaleMixin, nn.Conv2d):
    """utils/quantize.py:129-210.  The weight range is reduced and the weight fake-quantised per forward by the
    library's kernels (quantize with the range from dfq_minmax_f32); the convolution arithmetic itself is F.conv2d,
    as in QuantConv2d."""
    _scale_shape = (-1, 1, 1, 1)

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, bias=True,
                 num_bits=8, num_bits_act=8, num_bits_bias=16, momentum=0.1):
        super(QConv2d, self).__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, bias)
        self.num_bits = num_bits
        self.num_bits_bias = num_bits_bias
        self.quant = QuantMeasure(num_bits=num_bits_act, momentum=momentum)

    def merge_scale_prev(self, weight, scale_prev):
        # :168-177: the input channels of every group are divided by that group's slice of the previous layer's scale
        step_s = weight.shape[1]
        sp = scale_prev[:, 0, 0, 0].view(1, -1, 1, 1)
        step = weight.shape[0] // self.groups
        parts = [weight[g * step:(g + 1) * step] / sp[:, g * step_s:(g + 1) * step_s] for g in range(self.groups)]
        return torch.cat(parts, 0)

    def forward(self, input):
        input = self.quant(input)
        sweight, sbias = self._scaled()
        qweight = _quant_weight(sweight.contiguous(), self.num_bits)
        qbias = quantize(sbias.contiguous(), num_bits=self.num_bits_bias) if sbias is not None else None
        return _conv2d_fp32(input, qweight, qbias, self.stride, self.padding, self.dilation, self.groups)


class QLinear(_ScaleMixin, nn.Linear):
    """utils/quantize.py:260-323."""
    _scale_shape = (-1, 1)

    def __init__(self, in_features, out_features, bias=True, num_bits=8, num_bits_act=8, num_bits_bias=16,
                 momentum=0.1):
        super(QLinear, self).__init__(in_features, out_features, bias)
        self.num_bits = num_bits
        self.num_bits_bias = num_bits_bias
        self.quant = QuantMeasure(num_bits=num_bits_act, momentum=momentum)

    def merge_scale_prev(self, weight, scale_prev):
        return weight * scale_prev.view(1, -1)                     # :293-294

    def forward(self, input):
        input = self.quant(input)
        sweight, sbias = self._scaled()
        qweight = _quant_weight(sweight.contiguous(), self.num_bits)
        qbias = quantize(sbias.contiguous(), num_bits=self.num_bits_bias) if sbias is not None else None
        return F.linear(input, qweight, qbias)


def set_layer_bits(graph, bits_weight=8, bits_activation=8, bits_bias=16, targ_type=None):
    """utils/quantize.py:366-379, bug-compatible: QuantMeasure(bits_activation) sets `update_stat`
    (first positional argument), not `num_bits`."""
    print("Setting num_bits for targ layers...")
    assert targ_type != None, "targ_type cannot be None"
    with_quant = [graph[idx] for idx in graph if type(graph[idx]) in targ_type and hasattr(graph[idx], 'quant')]
    pool = None
    if with_quant:          # one zeroed device table for all the new measures instead of 2 tiny copies per layer
        pool = torch.zeros(len(with_quant), 2, dtype=torch.float32, device=with_quant[0].quant.running_min.device)
    slot = 0
    for idx in graph:
        if type(graph[idx]) in targ_type:
            if hasattr(graph[idx], 'quant'):
                graph[idx].quant = QuantMeasure(bits_activation, _storage=pool[slot])
                slot += 1
            if hasattr(graph[idx], 'num_bits'):
                graph[idx].num_bits = bits_weight
            if hasattr(graph[idx], 'num_bits_bias'):
                graph[idx].num_bits_bias = bits_bias
