"""Bias correction on B200 -- host mirror of the reference's bias_correction.py.

The fused kernel (csrc/correct.cu, dfq_bias_correct) recomputes eps = Q(W) - W in registers and reduces it
against a per-input-channel vector, so eps never reaches HBM.

Modes:
  mode="ref"   (default) the reference's control flow, bug for bug (bias_correction.py:147-258).  With the
               id()-keyed graph every script builds, `idx_layer not in bottoms` is always true and the call is a
               no-op returning ({}, {}); with a position-keyed graph the reference BODY runs
               ("ref-body": bias += mean_i(eps[o,i] + E[x][i]), the next BN's fake_bias gets the negated term).
  mode="paper" the method of Nagel et al.: bias -= eps . E[x], correctly keyed, E[x] from the preceding BN.
"""
import ctypes
import logging

import torch
from scipy.stats import norm

from . import _lib as L
from .utils.minmax import find_prev_bn

logging.basicConfig(level=logging.INFO)
logger = logging.getLogger(__name__)


def _pdf(x):
    return torch.from_numpy(norm(0, 1).pdf(x)).float()


def _cdf(x):
    return torch.from_numpy(norm.cdf(x)).float()


def _relu_mean(weight, bias):       # bias_correction.py:172
    return weight * _pdf(-bias / weight) + bias * (1 - _cdf(-bias / weight))


def expected_input(branch):
    """E[x] per input channel from the BatchNorms feeding a layer (bias_correction.py:15-58).  CPU fp32."""
    total, kind = None, None
    for bn, relu, kind in branch:
        beta, gamma = bn.fake_bias.detach().cpu().clone(), bn.fake_weight.detach().cpu().clone()
        if relu:
            e = _relu_mean(gamma, beta)
            e[e < 0] = 0
        else:
            e = beta
        if total is None:
            total = e
        elif kind == 'cat':
            total = torch.cat([total, e])
        else:
            total += e
    return kind, total


def eps_dot(layers, vectors, bits_weight=8, signed=False, alphas=None, biases=None, fold_mods=None):
    """out[l][o] = sum_{i,k} (Q(W_l)[o,i,k] - W_l[o,i,k]) * v_l[i]  for a list of layers in ONE launch set.
    v None => 1/I (row mean of the window-summed error).  biases/alphas: bias_l += alpha_l * out_l fused."""
    n = len(layers)
    if n == 0:
        return []
    dev = layers[0].weight.device
    descs = (L.DfqCorrDesc * n)()
    outs, keep, folds = [], [], []
    for i, layer in enumerate(layers):
        W = layer.weight.data
        O, I2 = W.shape[0], W.shape[1]
        out = torch.empty(O, dtype=torch.float32, device=dev)
        outs.append(out)
        d = descs[i]
        d.W = L.dev_f32(W, "weight")
        v = vectors[i]
        if v is not None:
            v = v.to(dev, torch.float32).contiguous()
            groups = getattr(layer, "groups", 1) if v.numel() != I2 else 1
            assert v.numel() == I2 * groups, "E[x] must have one entry per input channel"
            d.groups = groups
            keep.append(v)
            d.v = v.data_ptr()
        d.out = out.data_ptr()
        d.bias = L.dev_f32(biases[i], "bias") if biases is not None and biases[i] is not None else None
        d.O, d.I2, d.KK, d.bits = O, I2, W.numel() // (O * I2), int(bits_weight)
        d.flags = L.Q_SYMMETRIC if signed else 0
        d.alpha = float(alphas[i]) if alphas is not None else 0.0
        if fold_mods is not None and fold_mods[i]:
            f = torch.empty(int(fold_mods[i]), dtype=torch.float32, device=dev)
            folds.append(f)
            d.fold, d.fold_mod = f.data_ptr(), int(fold_mods[i])
        else:
            folds.append(None)
    tu, tb, ts = L.i64(), L.i64(), L.i64()
    L.check(L.lib().dfq_corr_plan_host(descs, n, ctypes.byref(tu), ctypes.byref(tb), ctypes.byref(ts)),
            "dfq_corr_plan_host")
    table = L.table_to_device(descs, dev)
    ranges = torch.empty(n, 4, dtype=torch.float32, device=dev)
    L.check(L.lib().dfq_bias_correct(table.data_ptr(), n, tu.value, tb.value, ts.value, ranges.data_ptr(), 0,
                                     L.stream()),
            "dfq_bias_correct")
    return (outs, folds) if fold_mods is not None else outs


def _next_bn_size(graph, keys, pos, bn_type):
    """Channel count of the first BatchNorm after position `pos` in graph order (where bias_prev lands)."""
    for k in keys[pos + 1:]:
        if isinstance(graph[k], bn_type) and hasattr(graph[k], "fake_bias"):
            return graph[k].fake_bias.size(0)
    return 0


def bias_correction(graph, bottoms, targ_type, bits_weight=8, bn_type=torch.nn.BatchNorm2d, signed=False,
                    mode="ref", visualize=False):
    """bias_correction.py:147-258.  Returns (bias_before_correction, bias_after_correction) like the reference.
    `visualize` is accepted (main_dfq.py:231 passes it; the reference signature lacks it and raises TypeError
    there -- documented divergence) and ignored."""
    assert isinstance(graph, dict), "Expected 'graph' to be a dictionary."
    assert isinstance(bottoms, dict), "Expected 'bottoms' to be a dictionary."
    assert isinstance(targ_type, (type, tuple)), "Expected 'targ_type' to be a type or tuple of types."
    if mode == "paper":
        return _bias_correction_paper(graph, bottoms, targ_type, bits_weight, bn_type, signed)
    assert mode == "ref", "mode must be 'ref' or 'paper'"
    logger.info("Starting bias correction...")
    bn_module, relu_attached = {}, {}
    bias_prev = None
    before, after = {}, {}
    keys = list(graph.keys())
    if not any(i in bottoms for i in range(len(keys))):
        # the reference logs one warning per layer here (:186); one summary line keeps the hot path quiet
        logger.warning(f"Layer indices 0..{len(keys) - 1} not found in bottoms (graph keyed by id(module)): "
                       "bias correction as wired in the reference is a no-op")
        logger.info("Bias correction completed.")
        return before, after
    with torch.no_grad():
        for idx_layer, layer in enumerate(graph.values()):
            layer_name = f"layer_{idx_layer}"
            if idx_layer not in bottoms:              # :185-187 -- always true for id()-keyed graphs
                logger.warning(f"Layer index {idx_layer} not found in bottoms")
                continue
            bot = bottoms[idx_layer]
            if isinstance(layer, targ_type):
                if hasattr(layer, 'bias') and layer.bias is not None:
                    before[layer_name] = layer.bias.data.clone()
            if bot is None or bot[0] == 'Data':
                continue
            node = graph[idx_layer]
            if isinstance(node, bn_type):
                bn_module[idx_layer] = node
                relu_attached[idx_layer] = False
                if bias_prev is not None:             # :206-213
                    node.fake_bias.add_(bias_prev.to(node.fake_bias.device))
                    bias_prev = None
                continue
            if isinstance(node, torch.nn.ReLU):
                if bot[0] in bn_module:
                    relu_attached[bot[0]] = True
            if isinstance(node, targ_type):
                bn_list, relu_list, kind_list, _ = find_prev_bn(bn_module, relu_attached, graph, bottoms, bot[:])
                branches = {}
                for i, (bn, bid) in enumerate(bn_list):
                    branches.setdefault(bid[0], []).append((bn, relu_list[i], kind_list[i]))
                W = node.weight.data
                O, I2 = W.shape[0], W.shape[1]
                C = _next_bn_size(graph, keys, idx_layer, bn_type)
                fold_total = None
                for kind, expect in (expected_input(b) for b in branches.values()):
                    if kind == 'cat':
                        raise RuntimeError("Tensors must have same number of dimensions: got 2 and 1")  # :74
                    E = expect.numel()
                    if E == I2:
                        # bias += mean_i(eps[o,i] + E[x][i])   (:76-77, :98-104); eps never leaves the kernel
                        (rowmean,), (fold,) = eps_dot([node], [None], bits_weight, signed, fold_mods=[C])
                        node.bias.data.add_(rowmean + float(expect.mean()))
                        # bias_prev = -(eps + E[x]) folded onto the next BN's channels (:206-213, :251)
                        if C:
                            tiled = expect.repeat(O)
                            efold = tiled.view(-1, C).mean(dim=0) if O * I2 != C else tiled
                            fold_total = -(fold.cpu() / float(O * I2 // C) + efold)
                    elif I2 == 1:
                        # depthwise: eps [O,1] + E[x] [E] broadcasts to [O,E] in the reference (:77)
                        (eps_o,) = eps_dot([node], [None], bits_weight, signed)
                        M = eps_o.cpu()[:, None] + expect[None, :]
                        node.bias.data.add_(M.mean(dim=1).to(node.bias.device))
                        if C:
                            flat = M.reshape(-1)
                            fold_total = -(flat.view(-1, C).mean(dim=0) if flat.numel() != C else flat)
                    else:
                        raise RuntimeError(f"The size of tensor a ({I2}) must match the size of tensor b ({E}) "
                                           "at non-singleton dimension 1")
                bias_prev = fold_total
                if hasattr(layer, 'bias') and layer.bias is not None:
                    after[layer_name] = layer.bias.data.clone()
    logger.info("Bias correction completed.")
    return before, after


def _bias_correction_paper(graph, bottoms, targ_type, bits_weight, bn_type, signed):
    """Nagel et al. (ICCV'19, sec. 4.2): b <- b - eps . E[x], E[x] from the BatchNorm(s) feeding each layer.
    Every layer of the model is corrected by ONE fused launch set; eps is never materialised."""
    bn_module, relu_attached = {}, {}
    layers, vectors, names = [], [], []
    before, after = {}, {}
    with torch.no_grad():
        for pos, (key, node) in enumerate(graph.items()):
            bot = bottoms[key]
            if bot is None or bot[0] == 'Data':
                continue
            if isinstance(node, bn_type):
                bn_module[key] = node
                relu_attached[key] = False
                continue
            if isinstance(node, torch.nn.ReLU) and bot[0] in bn_module:
                relu_attached[bot[0]] = True
            if isinstance(node, targ_type) and node.bias is not None:
                bn_list, relu_list, kind_list, _ = find_prev_bn(bn_module, relu_attached, graph, bottoms, bot[:])
                if not bn_list:
                    continue
                _, expect = expected_input([(bn, relu_list[i], kind_list[i]) for i, (bn, _b) in enumerate(bn_list)])
                groups = getattr(node, "groups", 1)
                if expect.numel() != node.weight.shape[1] * groups:
                    continue                        # E[x] does not line up with the input channels
                layers.append(node)
                vectors.append(expect)
                names.append(f"layer_{pos}")
                before[f"layer_{pos}"] = node.bias.data.clone()
        eps_dot(layers, vectors, bits_weight, signed, alphas=[-1.0] * len(layers),
                biases=[l.bias.data for l in layers])
        for n, l in zip(names, layers):
            after[n] = l.bias.data.clone()
    return before, after
