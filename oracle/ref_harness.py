"""Drives the UNMODIFIED reference (/root/reference) step by step -- build container only.

TEST INFRASTRUCTURE: used by oracle/gen_golden.py to pin oracle/dfq_oracle.c and to
produce tests/golden/*.  /root/reference does not exist on the GPU box, so nothing
under tests/ -m gpu, smoke() or bench.py imports this file.

The four import shims are the ones SURVEY.md section 8c documents; the step order is
main_dfq.py:161-258 (the script itself cannot be imported: main_dfq.py:15-16).
"""
import collections
import collections.abc
import sys
import types

import torch
import torch.nn as nn

REF = "/root/reference"


def import_reference():
    collections.Mapping = collections.abc.Mapping          # PyTransformer/transformers/utils.py:491
    for name in ("matplotlib", "matplotlib.pyplot", "pydot", "graphviz"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["graphviz"].Digraph = object
    if REF not in sys.path:
        sys.path.insert(0, REF)
    ns = types.SimpleNamespace()
    from modeling.classification.MobileNetV2 import mobilenet_v2
    from PyTransformer.transformers.torchTransformer import TorchTransformer
    from utils.relation import create_relation
    from utils import layer_transform as lt
    from utils import quantize as q
    from Cross_layer_equal import cross_layer_equalization, _layer_equalization
    from bias_absorption import bias_absorption
    from clip_weight import clip_weight
    import bias_correction as bc
    ns.mobilenet_v2 = mobilenet_v2
    ns.TorchTransformer = TorchTransformer
    ns.create_relation = create_relation
    ns.lt = lt
    ns.q = q
    ns.cross_layer_equalization = cross_layer_equalization
    ns._layer_equalization = _layer_equalization
    ns.bias_absorption = bias_absorption
    ns.clip_weight = clip_weight
    ns.bc = bc
    return ns


class ieee_sqrt:
    """Shim 5: make torch.sqrt correctly rounded while the reference runs.

    ATen-CPU sqrt goes through MKL VML (vsSqrt, HA mode), which returns the value one ulp
    below the IEEE-754 result for ~0.6 % of inputs (measured here, torch 2.11 + MKL 2024.2;
    scalar and vector calls alike).  That third-party deviation cannot be restated, so the
    bit-exact pin is taken with IEEE sqrt; the unpatched run is compared separately.
    """

    def __enter__(self):
        import numpy as np
        self._orig = torch.sqrt

        def _sqrt(x):
            return torch.as_tensor(np.sqrt(x.detach().numpy()))
        torch.sqrt = _sqrt
        return self

    def __exit__(self, *exc):
        torch.sqrt = self._orig
        return False


class RefRun:
    """One reference model taken through the DFQ steps; exposes tensors by module name."""

    def __init__(self, ns, model, trace_input):
        self.ns = ns
        q, lt = ns.q, ns.lt
        self.model = model.cpu().eval()
        self.targ_layer = [q.QuantNConv2d, q.QuantNLinear]
        self.transformer = ns.TorchTransformer()
        module_dict = {0: [(nn.ReLU6, nn.ReLU)],
                       1: [(nn.Conv2d, q.QuantNConv2d), (nn.Linear, q.QuantNLinear)]}
        self.model, self.transformer = lt.switch_layers(
            self.model, self.transformer, trace_input, module_dict,
            ignore_layer=[q.QuantMeasure], quant_op=True)
        self.graph = self.transformer.log.getGraph()
        self.bottoms = self.transformer.log.getBottoms()
        self.names = {id(m): n for n, m in self.model.named_modules()}
        self.relations = None

    # -- views ------------------------------------------------------------
    def key_name(self, k):
        return k if isinstance(k, str) else self.names[k]

    def graph_names(self):
        return [self.key_name(k) for k in self.graph]

    def bottoms_names(self):
        return {self.key_name(k): (None if v is None else [self.key_name(b) for b in v])
                for k, v in self.bottoms.items()}

    def targets(self):
        return [(self.names[k], m) for k, m in self.graph.items() if type(m) in self.targ_layer]

    def snapshot(self):
        out = {}
        for k, m in self.graph.items():
            if isinstance(m, str):
                continue
            n = self.names[k]
            if type(m) in self.targ_layer:
                out[n + ".weight"] = m.weight.detach().clone()
                if m.bias is not None:
                    out[n + ".bias"] = m.bias.detach().clone()
            elif isinstance(m, nn.BatchNorm2d) and hasattr(m, "fake_weight"):
                out[n + ".fake_weight"] = m.fake_weight.detach().clone()
                out[n + ".fake_bias"] = m.fake_bias.detach().clone()
        return out

    # -- steps (main_dfq.py:188-231) ----------------------------------------
    def merge_batchnorm(self):
        self.model = self.ns.lt.merge_batchnorm(self.model, self.graph, self.bottoms, self.targ_layer)

    def create_relation(self):
        self.relations = self.ns.create_relation(self.graph, self.bottoms, self.targ_layer)
        return [(self.names[a], self.names[b], self.names[c])
                for a, b, c in (r.get_idxs() for r in self.relations)]

    def equalize(self, **kw):
        self.ns.cross_layer_equalization(self.graph, self.relations, self.targ_layer,
                                         Save_state=False, **kw)

    def absorb(self, N=3):
        self.ns.bias_absorption(self.graph, self.relations, self.bottoms, N)

    def set_layer_bits(self, bw, ba, bb):
        self.ns.q.set_layer_bits(self.graph, bw, ba, bb, self.targ_layer)

    def quantize(self, bw, bb):
        self.ns.lt.quantize_targ_layer(self.graph, bw, bb, self.targ_layer)

    def set_quant_minmax(self):
        self.ns.lt.set_quant_minmax(self.graph, self.bottoms)

    def clip(self, lo=-15, hi=15):
        self.ns.clip_weight(self.graph, [lo, hi], self.targ_layer)

    def bias_correction_as_wired(self, bw):
        return self.ns.bc.bias_correction(self.graph, self.bottoms, tuple(self.targ_layer), bits_weight=bw)

    def fix_bn_eps(self):
        for m in self.model.modules():
            if isinstance(m, nn.BatchNorm2d) and m.eps == 0:
                m.eps = 1e-12

    def forward_quantized(self, x):
        self.fix_bn_eps()
        # main_dfq.py:240 calls model.eval() after the DFQ steps: the QuantMeasures that set_layer_bits created
        # (utils/quantize.py:373, born in training mode) must be in eval mode for inference
        self.model.eval()
        self.ns.lt.replace_op()
        try:
            with torch.no_grad():
                return self.model(x)
        finally:
            self.ns.lt.restore_op()
