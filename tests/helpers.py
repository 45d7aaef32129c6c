"""Shared test helpers: seeded MobileNetV2 state, numpy views, digests."""
import hashlib

import numpy as np
import torch
import torch.nn as nn


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def fresh_model():
    from data_free_quantization_b200.fixtures import synthetic_state_dict
    from data_free_quantization_b200.modeling.mobilenetv2 import mobilenet_v2
    torch.manual_seed(0)
    model = mobilenet_v2(None)
    sd = synthetic_state_dict(model, seed=0, bn_seed=1)
    model.load_state_dict(sd)
    return model.eval(), sd


def np_state(sd):
    return {k: v.detach().cpu().numpy().copy() for k, v in sd.items() if v.dtype == torch.float32}


def switch(model, dev=None, quant_n=True):
    """switch_layers with the harness' module_dict (ReLU6->ReLU, Conv/Linear -> QuantN*)."""
    from data_free_quantization_b200 import graph as G
    from data_free_quantization_b200.utils import quantize as Q
    if dev is not None:
        model = model.to(dev)
    pair = (Q.QuantNConv2d, Q.QuantNLinear) if quant_n else (Q.QuantConv2d, Q.QuantLinear)
    md = {0: [(nn.ReLU6, nn.ReLU)], 1: [(nn.Conv2d, pair[0]), (nn.Linear, pair[1])]}
    tr = G.TorchTransformer()
    data = torch.ones(4, 3, 224, 224, device=dev)
    model, tr = G.switch_layers(model, tr, data, md, ignore_layer=[Q.QuantMeasure], quant_op=True)
    return model, tr, list(pair)


def names_of(model):
    return {id(m): n for n, m in model.named_modules()}


def snapshot(model, graph, targ):
    """{name.weight/.bias/.fake_weight/.fake_bias: numpy} like oracle/ref_harness.RefRun.snapshot."""
    names = names_of(model)
    out = {}
    for k, m in graph.items():
        if isinstance(m, str):
            continue
        n = names[k]
        if type(m) in targ:
            out[n + ".weight"] = m.weight.detach().cpu().numpy()
            if m.bias is not None:
                out[n + ".bias"] = m.bias.detach().cpu().numpy()
        elif isinstance(m, nn.BatchNorm2d) and hasattr(m, "fake_weight"):
            out[n + ".fake_weight"] = m.fake_weight.detach().cpu().numpy()
            out[n + ".fake_bias"] = m.fake_bias.detach().cpu().numpy()
    return out


def rel_err(a, b):
    """Relative error against max(|a|, 1% of the tensor's largest magnitude) -- gen_golden.compare()."""
    denom = np.maximum(np.abs(a), 1e-2 * max(float(np.max(np.abs(a))), 1e-30))
    return float(np.max(np.abs(a - b) / denom))


def dfq_quantised(which, dev, golden_npz):
    """Seeded fixture model taken through main_dfq.py:188-231 on the GPU exactly as the golden generators ran the
    reference (oracle/gen_golden.py, gen_eval_golden.py): the reference's absorbed biases (golden `absorb/*`) are
    carried forward, so that weights AND biases of the quantised model equal the reference's bit for bit.
    Returns (model, graph, bottoms, targ, transformer)."""
    from data_free_quantization_b200 import graph as G
    from data_free_quantization_b200.bias_absorption import bias_absorption
    from data_free_quantization_b200.clip_weight import clip_weight
    from data_free_quantization_b200.Cross_layer_equal import cross_layer_equalization
    from data_free_quantization_b200.fixtures import synthetic_state_dict
    from data_free_quantization_b200.utils import quantize as Q
    from data_free_quantization_b200.utils.layer_transform import (merge_batchnorm, quantize_targ_layer,
                                                                   set_quant_minmax)
    from data_free_quantization_b200.utils.relation import create_relation
    torch.manual_seed(0)
    if which == "deeplab":
        from data_free_quantization_b200.modeling.deeplab import DeepLab
        model, trace = DeepLab(), torch.ones(1, 3, 129, 129, device=dev)
    else:
        from data_free_quantization_b200.modeling.mobilenetv2 import mobilenet_v2
        model, trace = mobilenet_v2(None), torch.ones(4, 3, 224, 224, device=dev)
    model.load_state_dict(synthetic_state_dict(model, seed=0, bn_seed=1))
    model = model.eval().to(dev)
    targ = [Q.QuantNConv2d, Q.QuantNLinear]
    md = {0: [(nn.ReLU6, nn.ReLU)], 1: [(nn.Conv2d, targ[0]), (nn.Linear, targ[1])]}
    model, tr = G.switch_layers(model, G.TorchTransformer(), trace, md, ignore_layer=[Q.QuantMeasure], quant_op=True)
    graph, bottoms = tr.log.getGraph(), tr.log.getBottoms()
    merge_batchnorm(model, graph, bottoms, targ)
    rels = create_relation(graph, bottoms, targ)
    cross_layer_equalization(graph, rels, targ, Save_state=False, Treshhold=2e-7)
    bias_absorption(graph, rels, bottoms, N=3)
    mods = dict(model.named_modules())
    for k in golden_npz.files:
        if k.startswith("absorb/"):
            mod, attr = k[len("absorb/"):].rsplit(".", 1)
            getattr(mods[mod], attr).data.copy_(torch.from_numpy(golden_npz[k]))
    Q.set_layer_bits(graph, 8, 8, 8, targ)
    merge_batchnorm(model, graph, bottoms, targ)
    quantize_targ_layer(graph, 8, 8, targ)
    set_quant_minmax(graph, bottoms, verbose=False)
    clip_weight(graph, [-15, 15], targ)
    model.eval()                                   # main_dfq.py:240
    return model, graph, bottoms, targ, tr
