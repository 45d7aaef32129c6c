"""High-bias absorption on B200 -- host mirror of the reference's bias_absorption.py.

One multi-relation launch pair (csrc/absorb.cu, dfq_absorb_clip): a pass over every W2 forming
wc = (sum_hw W2) . c per group with c = max(0, beta - N*gamma), then the three bias updates in relation
order.  `clip` optionally fuses clip_weight's clamp of W2 into the same pass.
"""
import ctypes

import torch
import torch.nn as nn

from . import _lib as L


def _is_relu_between(layer_second, layer_first, graph, bottoms):
    idx = layer_second                              # bias_absorption.py:10-17
    while idx != layer_first:
        if isinstance(graph[bottoms[idx][0]], torch.nn.ReLU):
            return True
        idx = bottoms[idx][0]
    return False


def _ensure_bias(layer):
    if layer.bias is None:                          # bias_absorption.py:72-74
        layer.bias = nn.Parameter(data=torch.zeros((layer.weight.size(0)), dtype=torch.float32,
                                                   device=layer.weight.device), requires_grad=False)


def bias_absorption(graph, relations, bottoms, N=3, visualize=False, clip=None):
    """bias_absorption.py:9-80.  `visualize` (matplotlib plots) is ignored.  Returns {relation: c}."""
    print("Start bias absorption")
    todo = [r for r in relations if _is_relu_between(r.layer_second, r.layer_first, graph, bottoms)]
    if not todo:
        return {}
    with torch.no_grad():
        n = len(todo)
        pairs = (L.DfqAbsorbPair * n)()
        dev = graph[todo[0].layer_first].weight.device
        c_all = torch.empty(sum(graph[r.layer_first].weight.size(0) for r in todo), dtype=torch.float32, device=dev)
        wc_all = torch.empty(sum(graph[r.layer_second].weight.size(0) for r in todo), dtype=torch.float32,
                             device=dev)
        co = wo = 0
        max_dim = 0
        out = {}
        for i, rel in enumerate(todo):
            first, second, bn = rel.get_idxs()
            l1, l2, bnm = graph[first], graph[second], graph[bn]
            _ensure_bias(l1)
            _ensure_bias(l2)
            W2 = l2.weight.data
            C1, O2, I2 = l1.weight.size(0), W2.shape[0], W2.shape[1]
            p = pairs[i]
            p.W2 = L.dev_f32(W2, "layer_second.weight")
            p.b1, p.b2 = L.dev_f32(l1.bias.data, "layer_first.bias"), L.dev_f32(l2.bias.data, "layer_second.bias")
            p.bn_w, p.bn_b = L.dev_f32(bnm.fake_weight, "fake_weight"), L.dev_f32(bnm.fake_bias, "fake_bias")
            p.c, p.wc = c_all.data_ptr() + 4 * co, wc_all.data_ptr() + 4 * wo
            p.C1, p.O2, p.I2, p.KK = C1, O2, I2, W2.numel() // (O2 * I2)
            out[rel] = c_all[co:co + C1]
            co, wo = co + C1, wo + O2
            max_dim = max(max_dim, C1, O2)
        rows = ctypes.c_int()
        L.check(L.lib().dfq_absorb_plan_host(pairs, n, ctypes.byref(rows)), "dfq_absorb_plan_host")
        table = L.table_to_device(pairs, dev)
        lo, hi = (float(clip[0]), float(clip[1])) if clip is not None else (0.0, 0.0)
        L.check(L.lib().dfq_absorb_clip(table.data_ptr(), n, rows.value, max_dim, float(N), int(clip is not None),
                                        lo, hi, L.stream()), "dfq_absorb_clip")
    return out
