"""Compiled form of the weight-side DFQ pipeline (main_dfq.py:188-231): the second time `run_dfq_steps` sees the same
traced model in the same fresh state, the reference-shaped functions are recorded ONCE into two CUDA graphs
(merge_batchnorm .. quantize_targ_layer, and clip_weight), and every later call replays them.

Why: the kernels of a MobileNetV2 pass take ~1 ms; rebuilding 9 descriptor tables, 52 bias Parameters, 53
QuantMeasures and the graph walk of set_quant_minmax in Python took 4 ms.  The graph is static, so the tables are
uploaded once (the captured copies re-read a pinned arena that lives with the plan) and the Python-visible side
effects of the functions -- new bias Parameters, fake_weight / fake_bias buffers, the identity-BN eps, the
QuantMeasures set_layer_bits creates, Relation.S -- are recorded as a list of (dict, key, value) assignments and
re-applied after each replay, so the model object ends up exactly as the eager functions leave it.

What is NOT captured: set_quant_minmax (host arithmetic on BatchNorm statistics: one D2H copy queued behind the first
graph, the numpy work overlapped with the second graph, one H2D copy) and bias_correction (host logic; a no-op as
wired in the reference).  A plan is only used when every tensor the captured kernels read sits at the address it
had at capture time and the model is in the same fresh state (biases that did not exist do not exist, BatchNorm eps
as before); anything else falls back to the eager functions.  DFQ_PIPELINE_PLAN=0 disables the mechanism.
"""
import contextlib
import io
import os

import torch
import torch.nn as nn

from . import _lib as L

_ATTRS = ("eps", "num_bits", "num_bits_bias")
_DATA_PTR = torch.Tensor.data_ptr
_plans = {}            # id(graph) -> PipelinePlan | signature seen once (first call runs eagerly)
stats = {"eager": 0, "captured": 0, "replayed": 0}


def enabled():
    return os.environ.get("DFQ_PIPELINE_PLAN", "1") != "0"


def _args_key(args, targ_pair):
    return (bool(args.quantize), bool(args.equalize), bool(args.absorption), bool(args.clip_weight),
            bool(args.correction), getattr(args, "correction_mode", "ref"), int(args.bits_weight),
            int(args.bits_activation), int(args.bits_bias), tuple(targ_pair))


def _reads(graph):
    """(module dict, key) of every tensor the captured kernels read from the caller's model."""
    slots = []
    for m in graph.values():
        if isinstance(m, nn.BatchNorm2d):
            slots += [(m._parameters, "weight"), (m._parameters, "bias"), (m._buffers, "running_mean"),
                      (m._buffers, "running_var")]
        elif isinstance(m, (nn.Conv2d, nn.Linear)):
            slots += [(m._parameters, "weight"), (m._parameters, "bias")]
    return slots


def _signature(graph, slots, bns):
    ptrs = tuple(None if d[k] is None else d[k].data_ptr() for d, k in slots)
    return ptrs, tuple(m.eps for m in bns)


def _snapshot(graph):
    snap = []
    for m in graph.values():
        if isinstance(m, nn.Module):
            snap.append((m, dict(m._parameters), dict(m._buffers), dict(m._modules),
                         {k: m.__dict__[k] for k in _ATTRS if k in m.__dict__}))
    return snap


def _diff(snap):
    """Assignments that turn the snapshot state into the current state."""
    fix = []
    for m, params, bufs, mods, attrs in snap:
        for before, now in ((params, m._parameters), (bufs, m._buffers), (mods, m._modules)):
            for k, v in now.items():
                if k not in before or before[k] is not v:
                    fix.append((now, k, v))
        for k in _ATTRS:
            if k in m.__dict__ and (k not in attrs or attrs[k] != m.__dict__[k]):
                fix.append((m.__dict__, k, m.__dict__[k]))
    return fix


def _restore(snap):
    """Undo every assignment made since `snap` (a capture that failed half-way ran no kernel)."""
    for m, params, bufs, mods, attrs in snap:
        for before, now in ((params, m._parameters), (bufs, m._buffers), (mods, m._modules)):
            for k in list(now):
                if k not in before:
                    del now[k]
            now.update(before)
        for k in _ATTRS:
            if k in attrs:
                m.__dict__[k] = attrs[k]


class PipelinePlan:
    def __init__(self, model, graph, bottoms, args, targ_pair, steps_fn):
        """Records one pass of `steps_fn(phase)` (phase 1: everything up to quantize_targ_layer, phase 2:
        clip_weight) into CUDA graphs.  The model must be in the fresh state; it is left as after a normal pass."""
        from .utils.minmax import minmax_plan
        self.graph, self.bottoms, self.args_key = graph, bottoms, _args_key(args, targ_pair)
        self.slots = _reads(graph)
        self.bns = [m for m in graph.values() if isinstance(m, nn.BatchNorm2d)]
        self.sig = _signature(graph, self.slots, self.bns)
        self._fast = None
        self.node_ids = tuple(map(id, graph.values()))
        snap = _snapshot(graph)
        self.cap = L.begin_capture()
        l0 = L.launch_count()
        try:
            self.pool = torch.cuda.graph_pool_handle()
            self.g1 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.g1, pool=self.pool, capture_error_mode="relaxed"):
                self.relations = steps_fn(1)
            self.g2 = None
            if args.clip_weight:
                self.g2 = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self.g2, pool=self.pool, capture_error_mode="relaxed"):
                    steps_fn(2)
        finally:
            L.end_capture()
        self.launches = L.launch_count() - l0            # kernels recorded in the two graphs
        self.fix = _diff(snap)
        # the measures set_layer_bits created: reattached on every replay in the state they were born with
        self.measures = [(v, v.update_stat, v.num_bits, v.training) for d, k, v in self.fix
                         if k == "quant" and isinstance(v, nn.Module)]
        self.quantize = bool(args.quantize)
        self.mm = minmax_plan(graph, bottoms) if self.quantize else None

    def matches(self, graph, bottoms, args, targ_pair):
        """The recorded state?  On the device's critical path of every replay (nothing is queued before it returns), so
        the 300-odd address reads go through one map() over the tensors that existed at capture time."""
        if not (graph is self.graph and bottoms is self.bottoms and _args_key(args, targ_pair) == self.args_key
                and tuple(map(id, graph.values())) == self.node_ids):
            return False
        if self._fast is None:
            ptrs = self.sig[0]
            self._fast = ([s for s, p in zip(self.slots, ptrs) if p is not None],
                          [s for s, p in zip(self.slots, ptrs) if p is None], tuple(p for p in ptrs if p is not None))
        tens, nones, ptrs = self._fast
        try:
            if tuple(map(_DATA_PTR, [d[k] for d, k in tens])) != ptrs:
                return False
        except TypeError:                                   # a tensor of the capture is None now
            return False
        if any(d[k] is not None for d, k in nones):
            return False
        return tuple(m.eps for m in self.bns) == self.sig[1]

    def replay(self, finish=True):
        """One pass: first graph, statistics copy, second graph, side effects, ranges (finish=False: the caller has
        host work to do while the graphs run and calls finish_ranges() itself)."""
        self.g1.replay()
        if self.mm is not None:
            self.mm.start()                       # D2H of the BatchNorm statistics behind the first graph
        if self.g2 is not None:
            self.g2.replay()
        for d, k, v in self.fix:
            d[k] = v
        for q, us, nb, tr in self.measures:
            q.update_stat, q.num_bits, q.training = us, nb, tr
        L.bump_weights_epoch()
        L.add_launches(self.launches)
        if self.mm is not None:
            self.mm.speculate()                   # the previous pass's ranges, queued behind the graphs (see MinmaxPlan)
        if finish:
            self.finish_ranges()
        return self.relations

    def finish_ranges(self):
        if self.mm is not None:
            self.mm.finish()                      # set_quant_minmax: host arithmetic + one H2D per range table


def lookup(graph, bottoms, args, targ_pair):
    """('replay', plan) when a matching plan exists, ('capture', None) when this fresh state was seen before (record
    now), ('eager', None) otherwise (first sight, or not the recorded state)."""
    if not enabled():
        return "eager", None
    entry = _plans.get(id(graph))
    if isinstance(entry, tuple) and entry and entry[0] == "failed":
        return "eager", None
    if isinstance(entry, PipelinePlan):
        if entry.matches(graph, bottoms, args, targ_pair):
            return "replay", entry
        return "eager", None
    slots = _reads(graph)
    bns = [m for m in graph.values() if isinstance(m, nn.BatchNorm2d)]
    seen = (graph, _args_key(args, targ_pair), _signature(graph, slots, bns))
    if entry is not None and entry[0] is graph and entry[1:] == seen[1:]:
        return "capture", None
    if len(_plans) >= 4:
        _plans.clear()
    _plans[id(graph)] = seen
    return "eager", None


def store(graph, plan):
    _plans[id(graph)] = plan


def forget(graph=None):
    if graph is None:
        _plans.clear()
    else:
        _plans.pop(id(graph), None)
