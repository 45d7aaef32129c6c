"""Trace-free graph construction -- produces the reference's `graph` / `bottoms` dicts.

The reference pushes a logging proxy through model.forward (PyTransformer, out of scope).  Here the
structure comes from one torch.fx symbolic trace; the dict formats are the reference's:
  graph   : OrderedDict  key -> nn.Module | str     ('Data' first; modules keyed by id(module); tensor ops
            keyed '<op>_<position>' with op in add / torch.cat / torch.mean / F.interpolate / F.pad / F.softmax)
  bottoms : same keys -> list of producer keys (None for 'Data')
PyTransformer/transformers/utils.py:75-105,207-211; torchTransformer.py:485-488,584-602.

switch_layers() also records the tensor ops whose operands the reference re-quantises at eval time
(utils/layer_transform.py:171-195) and installs a forward that executes the traced graph, applying those
QuantMeasures while replace_op() is in force -- no global monkey-patching, no inspect.stack().
"""
import operator
import types
from collections import OrderedDict

import torch
import torch.fx
import torch.nn as nn
import torch.nn.functional as F

from .utils.quantize import (QuantConv2d, QuantLinear, QuantMeasure, QuantNConv2d, QuantNLinear)

_QUANT_LEAVES = (QuantConv2d, QuantLinear, QuantNConv2d, QuantNLinear, QuantMeasure)

# fx target -> reference op name; value = True when every tensor argument is an operand (add, cat)
_ADD = {operator.add, operator.iadd, torch.add}
_OPS = {torch.cat: "torch.cat", torch.mean: "torch.mean", F.interpolate: "F.interpolate", F.pad: "F.pad",
        F.softmax: "F.softmax"}
_METHOD_OPS = {"add": "add", "add_": "add", "mean": "torch.mean", "softmax": "F.softmax"}
_QUANT_OP_TAG = {"add": "add", "torch.cat": "torch_cat", "torch.mean": "torch_mean",
                 "F.interpolate": "F_interpolate", "F.softmax": "F_softmax"}


class _Tracer(torch.fx.Tracer):
    def is_leaf_module(self, m, qualname):
        return isinstance(m, _QUANT_LEAVES) or super().is_leaf_module(m, qualname)


class Log:
    """The slice of PyTransformer's Log the DFQ steps read (getGraph / getBottoms / getRecordTensorOP)."""

    def __init__(self):
        self.graph = OrderedDict()
        self.bottoms = OrderedDict()
        self.record_tensor_op = []
        self.names = {}

    def getGraph(self):
        return self.graph

    def getBottoms(self):
        return self.bottoms

    def getRecordTensorOP(self):
        return self.record_tensor_op


def _node_args(node):
    out = []

    def visit(a):
        if isinstance(a, torch.fx.Node):
            out.append(a)
        elif isinstance(a, (list, tuple)):
            for b in a:
                visit(b)
    for a in node.args:
        visit(a)
    for a in node.kwargs.values():
        visit(a)
    return out


def _op_name(node):
    if node.op == "call_function":
        if node.target in _ADD:
            return "add"
        return _OPS.get(node.target)
    if node.op == "call_method":
        return _METHOD_OPS.get(node.target)
    return None


def _operands(node, opname):
    """fx Nodes that are tensor operands of a recorded op, in call order."""
    if opname == "torch.cat":
        first = node.args[0] if node.args else node.kwargs.get("tensors")
        return [a for a in first if isinstance(a, torch.fx.Node)]
    if opname == "add":
        return [a for a in node.args[:2] if isinstance(a, torch.fx.Node)]
    first = node.args[0] if node.args else None
    return [first] if isinstance(first, torch.fx.Node) else []


_NON_TENSOR_METHODS = {"size", "dim", "numel"}


class TorchTransformer:
    """Layer swapping + graph construction with the call surface of PyTransformer's TorchTransformer
    (register :32, trans_layers :42-92, _build_graph :321-368, .log)."""

    def __init__(self):
        self._register_dict = OrderedDict()
        self.log = Log()
        self.fx_graph = None
        self._producer = {}

    def register(self, origin_class, target_class):
        self._register_dict[origin_class] = target_class

    # -- layer swap ----------------------------------------------------------------------------------
    @staticmethod
    def _swap(old, target, update):
        if isinstance(old, nn.Conv2d) and issubclass(target, nn.Conv2d):
            new = target(old.in_channels, old.out_channels, old.kernel_size, old.stride, old.padding, old.dilation,
                         old.groups, old.bias is not None)
        elif isinstance(old, nn.Linear) and issubclass(target, nn.Linear):
            new = target(old.in_features, old.out_features, old.bias is not None)
        elif isinstance(old, (nn.ReLU6, nn.ReLU)) and issubclass(target, (nn.ReLU, nn.ReLU6)):
            return target(inplace=old.inplace)
        else:
            new = target()
        if update:                                   # carry parameters/buffers over (dict_merge, utils.py:477)
            for name, p in old._parameters.items():
                new._parameters[name] = p
            for name, b in old._buffers.items():
                new._buffers[name] = b
            dev = next((p.device for p in old.parameters()), None)
            if dev is not None:
                new.to(dev)
        new.train(old.training)
        return new

    def trans_layers(self, model, update=True):
        for name, child in list(model.named_children()):
            if type(child) in self._register_dict:
                setattr(model, name, self._swap(child, self._register_dict[type(child)], update))
            else:
                self.trans_layers(child, update)
        return model

    # -- graph ---------------------------------------------------------------------------------------
    def _build_graph(self, model, data=None, ignore_layer=()):
        tracer = _Tracer()
        fxg = tracer.trace(model)
        self.fx_graph = fxg
        self.model = model
        log = self.log = Log()
        mods = dict(model.named_modules())
        producer = self._producer = {}
        for node in fxg.nodes:
            if node.op == "placeholder":
                if "Data" not in log.graph:
                    log.graph["Data"], log.bottoms["Data"] = "Data", None
                producer[node] = "Data"
            elif node.op == "call_module":
                m = mods[node.target]
                key = id(m)
                log.graph[key] = m
                log.bottoms[key] = [producer[a] for a in _node_args(node) if producer.get(a) is not None]
                log.names[key] = node.target
                producer[node] = key
            elif node.op in ("call_function", "call_method"):
                opname = _op_name(node)
                ops = [a for a in _operands(node, opname) if producer.get(a) is not None] if opname else []
                recorded = opname is not None and (len(ops) >= 2 if opname == "add" else len(ops) >= 1)
                if recorded:
                    key = "{}_{}".format(opname, len(log.graph))
                    log.graph[key] = key
                    log.bottoms[key] = [producer[a] for a in ops]
                    producer[node] = key
                    node.meta["dfq_key"], node.meta["dfq_operands"] = key, ops
                    if opname in _QUANT_OP_TAG:
                        log.record_tensor_op.append(
                            (key, "{}_{}_{}".format(_QUANT_OP_TAG[opname], len(log.graph) - 1, len(ops))))
                elif node.op == "call_method" and node.target in _NON_TENSOR_METHODS:
                    producer[node] = None
                else:                                # view / reshape / getitem / flatten ...: transparent
                    src = [producer[a] for a in _node_args(node) if producer.get(a) is not None]
                    producer[node] = src[0] if src else None
            elif node.op == "get_attr":
                producer[node] = None
        return log


class CustomTensorOP(nn.Module):
    """utils/layer_transform.py:200-237 -- the QuantMeasures of the recorded tensor ops, in call order."""

    def __init__(self, tensor_op, name_tensor_op):
        super(CustomTensorOP, self).__init__()
        for idx, op in enumerate(tensor_op):
            self.add_module(str(idx), op)
        self.idx_tensor_op = 0
        self.len = len(tensor_op)
        self.name_tensor_op = name_tensor_op
        self.idx_name_tensor_op = 0
        self.num_op = len(name_tensor_op)

    def add_idx_tensor_op(self):
        self.idx_tensor_op = (self.idx_tensor_op + 1) % self.len

    def add_idx_name_tensor_op(self):
        self.idx_name_tensor_op = (self.idx_name_tensor_op + 1) % self.num_op

    def get_module_name(self):
        return self.name_tensor_op[self.idx_name_tensor_op][1]

    def get_graph_name(self):
        return self.name_tensor_op[self.idx_name_tensor_op][0]

    def get_module_next(self):
        mod = self._modules[str(self.idx_tensor_op)]
        self.add_idx_tensor_op()
        return mod

    def forward(self, x):
        x = self._modules[str(self.idx_tensor_op)](x)
        self.add_idx_tensor_op()
        return x


# ---- eval-time tensor-op quantisation (utils/layer_transform.py:18-158) -----------------------------
_state = {"quant_ops": False}


def replace_op():
    """utils/layer_transform.py:128-144: from now on the recorded add / cat / mean / interpolate / softmax
    operands go through their QuantMeasures.  (The reference monkey-patches torch globally; here it is a flag
    read by the graph executor installed by switch_layers.)"""
    _state["quant_ops"] = True


def restore_op():
    """utils/layer_transform.py:147-158."""
    _state["quant_ops"] = False


def _load_arg(a, env):
    if isinstance(a, torch.fx.Node):
        return env[a]
    if isinstance(a, (list, tuple)):
        return type(a)(_load_arg(b, env) for b in a)
    if isinstance(a, dict):
        return {k: _load_arg(v, env) for k, v in a.items()}
    if isinstance(a, slice):
        return slice(_load_arg(a.start, env), _load_arg(a.stop, env), _load_arg(a.step, env))
    return a


def _load_arg_quantised(a, env, operands, values, k):
    """_load_arg with the recorded operands replaced, occurrence by occurrence and in order, by their quantised values:
    `x + x` consumes TWO measures of the CustomTensorOP, one per occurrence (utils/layer_transform.py:18-32)."""
    if isinstance(a, torch.fx.Node):
        if k[0] < len(operands) and a is operands[k[0]]:
            k[0] += 1
            return values[k[0] - 1]
        return env[a]
    if isinstance(a, (list, tuple)):
        return type(a)(_load_arg_quantised(b, env, operands, values, k) for b in a)
    if isinstance(a, dict):
        return {key: _load_arg_quantised(v, env, operands, values, k) for key, v in a.items()}
    return _load_arg(a, env)


def run_graph(fxg, model, *inputs):
    """Execute the traced graph node by node; recorded tensor ops quantise their operands first."""
    env = {}
    it = iter(inputs)
    mods = dict(model.named_modules())
    top = getattr(model, "custom_tensor_op", None)
    quant = _state["quant_ops"] and top is not None
    for node in fxg.nodes:
        if node.op == "placeholder":
            env[node] = next(it)
        elif node.op == "get_attr":
            obj = model
            for part in node.target.split("."):
                obj = getattr(obj, part)
            env[node] = obj
        elif node.op == "output":
            return _load_arg(node.args[0], env)
        else:
            if quant and "dfq_key" in node.meta and top.num_op and top.get_graph_name() == node.meta["dfq_key"]:
                operands = node.meta["dfq_operands"]
                values = [top(env[a]) for a in operands]          # one measure per occurrence, in operand order
                top.add_idx_name_tensor_op()
                k = [0]
                args = _load_arg_quantised(node.args, env, operands, values, k)
                kwargs = _load_arg_quantised(node.kwargs, env, operands, values, k)
            else:
                args, kwargs = _load_arg(node.args, env), _load_arg(node.kwargs, env)
            if node.op == "call_module":
                env[node] = mods[node.target](*args, **kwargs)
            elif node.op == "call_function":
                env[node] = node.target(*args, **kwargs)
            else:
                env[node] = getattr(args[0], node.target)(*args[1:], **kwargs)
    raise RuntimeError("graph has no output node")


def switch_layers(model, transformer, data, module_dict, ignore_layer=[], ignore_op=['pad'], quant_op=True):
    """utils/layer_transform.py:161-197: swap layer classes, build graph/bottoms, and (quant_op) create one
    8-bit QuantMeasure per operand of every recorded tensor op as `model.custom_tensor_op`."""
    for key in module_dict:
        for source, target in module_dict[key]:
            transformer.register(source, target)
        model = transformer.trans_layers(model, update=True if key == 1 else False)
    transformer._build_graph(model, data, ignore_layer)
    fxg = transformer.fx_graph
    model.forward = types.MethodType(lambda self, *inputs: run_graph(fxg, self, *inputs), model)
    if not quant_op:
        return model, transformer
    tmp = [t for t in transformer.log.getRecordTensorOP() if not any(ig in t[1] for ig in ignore_op)]
    dev = next((p.device for p in model.parameters()), torch.device("cpu"))
    tensor_op_quantize = []
    for _layer_name, t in tmp:
        for _ in range(int(t.split('_')[-1])):
            tensor_op_quantize.append(QuantMeasure(num_bits=8, momentum=0.1).to(dev))
    module_tensor_op = CustomTensorOP(tensor_op_quantize, tmp)
    model.add_module('custom_tensor_op', module_tensor_op)
    setattr(model, 'name_tensor_op', tmp)
    setattr(model, 'idx_name_tensor_op', 0)
    setattr(model, 'idx_tensor_op', 0)
    transformer.module_tensor_op = module_tensor_op
    _state["module_tensor_op"] = module_tensor_op           # the reference keeps it in a module global (:171)
    return model, transformer
