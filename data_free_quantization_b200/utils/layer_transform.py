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
    stash |gamma| / beta as the BN's fake_weight / fake_bias buffers, turn the BN into an identity.
    Every (conv, BN) pair of the model goes through ONE launch (dfq_merge_bn_multi)."""
    import ctypes
    pairs = []
    for layer_idx in graph:
        if bottoms[layer_idx] is None:
            continue
        for bot_idx in bottoms[layer_idx]:
            if type(graph[layer_idx]) == nn.BatchNorm2d and type(graph[bot_idx]) in targ_type:
                pairs.append((graph[bot_idx], graph[layer_idx]))
                break
    if not pairs:
        return model
    with torch.no_grad():
        dev = pairs[0][0].weight.device
        total = sum(conv.weight.size(0) for conv, _ in pairs)
        stash = torch.empty(2, total, dtype=torch.float32, device=dev)     # all fake_weight | all fake_bias
        missing = sum(conv.weight.size(0) for conv, _ in pairs if conv.bias is None)
        zeros = torch.zeros(missing, dtype=torch.float32, device=dev) if missing else None   # one fill for all
        zoff = 0
        descs = (L.DfqMergeDesc * len(pairs))()
        off = 0
        for d, (conv, bn) in zip(descs, pairs):
            W = conv.weight.data
            O = W.size(0)
            if conv.bias is None:                   # :262-263 ('bias' is a registered, empty parameter slot)
                conv._parameters['bias'] = nn.Parameter(data=zeros[zoff:zoff + O], requires_grad=False)
                zoff += O
            fake_w, fake_b = stash[0, off:off + O], stash[1, off:off + O]
            off += O
            bp, bb = bn._parameters, bn._buffers
            d.W, d.b = L.dev_f32(W, "weight"), L.dev_f32(conv._parameters['bias'].data, "bias")
            d.gamma, d.beta = L.dev_f32(bp['weight'].data, "bn.weight"), L.dev_f32(bp['bias'].data, "bn.bias")
            d.mean, d.var = L.dev_f32(bb['running_mean'], "bn.running_mean"), L.dev_f32(bb['running_var'], "bn.running_var")
            d.fake_weight, d.fake_bias = fake_w.data_ptr(), fake_b.data_ptr()
            d.O, d.K, d.eps = O, W.numel() // O, float(bn.eps)
            if 'fake_weight' in bn._buffers:                      # :273-274; re-registration: plain replacement
                bn._buffers['fake_weight'], bn._buffers['fake_bias'] = fake_w, fake_b
            else:
                bn.register_buffer('fake_weight', fake_w)
                bn.register_buffer('fake_bias', fake_b)
            bn.__dict__['eps'] = IDENTITY_BN_EPS                 # :281 (see IDENTITY_BN_EPS); plain attribute
        rows = ctypes.c_int()
        L.check(L.lib().dfq_merge_plan_host(descs, len(pairs), ctypes.byref(rows)), "dfq_merge_plan_host")
        table = L.table_to_device(descs, dev)
        # reset_bn=1: gamma=1, var=1, beta=0, mean=0 (:277-280) written by the same kernel
        L.check(L.lib().dfq_merge_bn_multi(table.data_ptr(), len(pairs), rows.value, 1, L.stream()),
                "dfq_merge_bn_multi")
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
