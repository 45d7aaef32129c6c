"""Oracle DFQ weight-side pipeline over plain numpy state -- TEST INFRASTRUCTURE ONLY.

State layout (`st`, a dict keyed by the reference model's module names):
  st["<layer>.weight"]  float32 [O, I/g, kh, kw] or [O, I]
  st["<layer>.bias"]    float32 [O]   (absent until a step creates it)
  st["<bn>.weight/.bias/.running_mean/.running_var"]  raw BatchNorm tensors
  st["<bn>.fake_weight/.fake_bias"]                   after merge (layer_transform.py:273-274)
`topo` describes the traced graph:
  topo["targets"]   target layer names in graph order
  topo["conv_bn"]   {layer: bn following it}
  topo["relations"] [(first, second, bn, relu_between)]   (utils/relation.py:36-106)
Step order and constants follow main_dfq.py:188-231.
"""
import numpy as np

from . import oracle as O


def merge_batchnorm(st, topo, eps=1e-5):
    """utils/layer_transform.py:240-285."""
    for layer, bn in topo["conv_bn"].items():
        W = st[layer + ".weight"]
        if layer + ".bias" not in st:
            st[layer + ".bias"] = np.zeros(W.shape[0], dtype=np.float32)
        g, b = st[bn + ".weight"], st[bn + ".bias"]
        O.merge_bn(W, st[layer + ".bias"], g, b, st[bn + ".running_mean"], st[bn + ".running_var"],
                   st.get(bn + ".eps", eps))
        st[bn + ".fake_weight"] = np.abs(g).astype(np.float32)
        st[bn + ".fake_bias"] = b.copy()
        g.fill(1), b.fill(0)
        st[bn + ".running_var"].fill(1), st[bn + ".running_mean"].fill(0)
        st[bn + ".eps"] = 0.0


def cross_layer_equalization(st, topo, s_min_max=(1e-8, 1e8), threshold=2e-7, count=20,
                             signed=False, eps=0.0, max_iters=None):
    """Cross_layer_equal.py:63-115.  Returns (outer iterations run, [diff per iteration], S)."""
    diff, stall, iters, diffs = 1e8, 0, 0, []
    S_acc = {}
    while diff > threshold and stall < count:
        if max_iters is not None and iters >= max_iters:
            break
        old = {n: st[n + ".weight"].copy() for n in topo["targets"]}
        for first, second, bn, _ in topo["relations"]:
            if first + ".bias" not in st:
                st[first + ".bias"] = np.zeros(st[first + ".weight"].shape[0], dtype=np.float32)
            S = O.cle_pair(st[first + ".weight"], st[second + ".weight"], st[first + ".bias"],
                           st[bn + ".fake_weight"], st[bn + ".fake_bias"], s_min_max[0], s_min_max[1],
                           signed, eps)
            key = (first, second)
            S_acc[key] = S if key not in S_acc else (S_acc[key] * S).astype(np.float32)
        d = float(np.sum([O.mean_abs_diff(st[n + ".weight"], old[n]) for n in topo["targets"]]))
        diffs.append(d)
        iters += 1
        if abs(diff - d) > 1e-9:
            stall, diff = 0, d
        else:
            stall += 1
    return iters, diffs, S_acc


def bias_absorption(st, topo, N=3):
    """bias_absorption.py:9-80."""
    for first, second, bn, relu in topo["relations"]:
        if not relu:
            continue
        W2 = st[second + ".weight"]
        for l in (first, second):
            if l + ".bias" not in st:
                st[l + ".bias"] = np.zeros(st[l + ".weight"].shape[0], dtype=np.float32)
        O.absorb_pair(W2, st[first + ".weight"].shape[0], st[first + ".bias"], st[second + ".bias"],
                      st[bn + ".fake_weight"], st[bn + ".fake_bias"], float(N))


def quantize_targ_layer(st, topo, bit_weight=8, bits_bias=16, per_channel=False, want_codes=False):
    """utils/layer_transform.py:288-305.  Returns {layer: int32 codes} when asked."""
    codes = {}
    for n in topo["targets"]:
        W = st[n + ".weight"]
        if per_channel:
            r = O.quantize_per_channel(W, bit_weight, want_codes)
        else:
            lo, hi = O.minmax(W)
            r = O.fake_quant(W, bit_weight, lo, hi, want_codes=want_codes)
        if want_codes:
            r, codes[n] = r
        st[n + ".weight"] = r
        if n + ".bias" in st and bits_bias < 32:
            b = st[n + ".bias"]
            lo, hi = O.minmax(b)
            st[n + ".bias"] = O.fake_quant(b, bits_bias, lo, hi)
    return codes


def clip_weight(st, topo, lo=-15.0, hi=15.0):
    """clip_weight.py:4-33."""
    for n in topo["targets"]:
        O.clip(st[n + ".weight"], lo, hi)
