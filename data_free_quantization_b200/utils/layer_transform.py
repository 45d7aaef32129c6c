"""Host mirror of the reference's utils/layer_transform.py for the DFQ hot path.

merge_batchnorm / quantize_targ_layer run on libdfq_b200 kernels; set_quant_minmax and find_prev_bn are
scalar graph logic and live in utils/minmax.py (re-exported here); switch_layers / replace_op /
restore_op are provided by the trace-free graph builder in graph.py (re-exported here).
"""
import torch
import torch.nn as nn

from .. import _lib as L
from .quantize import (QConv2d, QLinear, QuantConv2d, QuantLinear, QuantMeasure, QuantNConv2d, QuantNLinear,  # noqa
                       fake_quant_tensors, quantize)

# torch >= 2 refuses batch_norm(eps=0); the reference sets eps = 0 (utils/layer_transform.py:281).  1e-12 is
# bit-identical in fp32 (sqrt(1 + 1e-12) == 1.0f) and keeps the identity BatchNorm callable.
IDENTITY_BN_EPS = 1e-12


def merge_batchnorm(model, graph, bottoms, targ_type=[QConv2d]):
    """utils/layer_transform.py:240-285: fold BatchNorm into the preceding target layer (in place, on the GPU),
    stash |gamma| / beta as the BN's fake_weight / fake_bias buffers, turn the BN into an identity."""
    with torch.no_grad():
        for layer_idx in graph:
            if bottoms[layer_idx] is None:
                continue
            for bot_idx in bottoms[layer_idx]:
                bn, conv = graph[layer_idx], graph[bot_idx]
                if type(bn) == nn.BatchNorm2d and type(conv) in targ_type:
                    W = conv.weight.data
                    O = W.size(0)
                    if conv.bias is None:           # :262-263
                        conv.bias = nn.Parameter(data=torch.zeros(O, dtype=torch.float32, device=W.device),
                                                 requires_grad=False)
                    fake_w = torch.empty(O, dtype=torch.float32, device=W.device)
                    fake_b = torch.empty(O, dtype=torch.float32, device=W.device)
                    L.check(L.lib().dfq_merge_bn(
                        L.dev_f32(W, "weight"), L.dev_f32(conv.bias.data, "bias"), O, W.numel() // O,
                        L.dev_f32(bn.weight.data, "bn.weight"), L.dev_f32(bn.bias.data, "bn.bias"),
                        L.dev_f32(bn.running_mean, "bn.running_mean"), L.dev_f32(bn.running_var, "bn.running_var"),
                        float(bn.eps), fake_w.data_ptr(), fake_b.data_ptr(), L.stream()), "dfq_merge_bn")
                    bn.register_buffer('fake_weight', fake_w)    # :273-274
                    bn.register_buffer('fake_bias', fake_b)
                    bn.weight.fill_(1)                           # :277-281
                    bn.running_var.fill_(1)
                    bn.bias.fill_(0)
                    bn.running_mean.fill_(0)
                    bn.eps = IDENTITY_BN_EPS
                    break
    return model


def quantize_targ_layer(graph, bit_weight=8, bits_bias=16, targ_type=None, per_channel=False, clip=None,
                        want_codes=False):
    """utils/layer_transform.py:288-305: fake-quantise every target layer's weight (and bias if bits_bias < 32)
    in place -- ONE multi-tensor launch (range reduce + apply) instead of 2 x 53 reductions with host syncs.

    Extensions (off by default): per_channel=True quantises each output channel with its own range;
    clip=(lo, hi) fuses clip_weight's clamp behind the quantisation (main_dfq.py:214,227 order);
    want_codes=True also returns {graph key: (uint8/int16 codes, [rows, 2] fp32 (min, max))}."""
    print("Quantizing Layer parameters")
    if bits_bias == 32:
        print("Skipping bias quantization (32 bits)")
    assert targ_type != None, "targ_type cannot be None!"
    tensors, bits, flags, rows, codes, keys = [], [], [], [], [], []
    with torch.no_grad():
        for layer_idx in graph:
            layer = graph[layer_idx]
            if type(layer) in targ_type:
                w = layer.weight.data
                tensors.append(w)
                bits.append(bit_weight)
                flags.append(L.Q_PER_CHANNEL if per_channel else 0)
                rows.append(w.size(0) if per_channel else 1)
                keys.append(layer_idx)
                codes.append(torch.empty(w.shape, dtype=torch.uint8 if bit_weight <= 8 else torch.int16,
                                         device=w.device) if want_codes else None)
                if layer.bias is not None and bits_bias < 32:
                    tensors.append(layer.bias.data)
                    bits.append(bits_bias)
                    flags.append(0)
                    rows.append(1)
                    keys.append(None)
                    codes.append(None)
        if not tensors:
            return {} if want_codes else graph
        if clip is not None:
            # the clamp belongs to the weights only: two launches when biases are present
            wi = [i for i, k in enumerate(keys) if k is not None]
            bi = [i for i, k in enumerate(keys) if k is None]
            pick = lambda lst, idx: [lst[i] for i in idx]
            rw = fake_quant_tensors(pick(tensors, wi), pick(bits, wi), pick(flags, wi), codes=pick(codes, wi),
                                    clip=clip, rows=pick(rows, wi))
            if bi:
                fake_quant_tensors(pick(tensors, bi), pick(bits, bi), pick(flags, bi), rows=pick(rows, bi))
            ranges, order = rw, wi
        else:
            ranges = fake_quant_tensors(tensors, bits, flags, codes=codes, rows=rows)
            order = list(range(len(tensors)))
    if not want_codes:
        return graph
    out, slot = {}, 0
    for i in order:
        if keys[i] is not None:
            out[keys[i]] = (codes[i], ranges[slot:slot + rows[i]])
        slot += rows[i]
    return out


from .minmax import find_prev_bn, set_quant_minmax                    # noqa: E402,F401
from ..graph import (CustomTensorOP, TorchTransformer, replace_op, restore_op, switch_layers)   # noqa: E402,F401
