"""set_quant_minmax / find_prev_bn -- host mirror of utils/layer_transform.py:308-618.

Scalar graph logic over per-channel BatchNorm statistics (<= ~1k floats per node): it stays on the host
(SURVEY.md 8a row a11).  BN vectors are pulled to the CPU once (one D2H copy of merge_batchnorm's stash), the
per-channel arithmetic runs on numpy fp32 arrays -- the same IEEE operations in the same order as the reference's
torch-CPU ops, a few us per call instead of tens -- the Gaussian / clipped-Gaussian moments use the fp64 pdf/cdf
exactly as the reference does, and the resulting (min, max) pairs go to the QuantMeasure buffers with one H2D copy.

One documented divergence: sqrt.  numpy's fp32 sqrt is correctly rounded; torch-CPU's vectorised sqrt is not (1 ulp
off on ~0.7% of inputs, and which ones depends on the element's position in the SIMD loop), so ranges that pass
through sqrt(var + eps) on an add branch (:420-431) can differ from a given torch-CPU run by 1 ulp.  The goldens
(tests/golden/minmax_random_bn.json, quant_minmax of both models) are met exactly / within 1e-6.
"""
import math

import numpy as np
import torch
import torch.nn as nn
from scipy.special import ndtr

from .quantize import (QConv2d, QLinear, QuantConv2d, QuantLinear, QuantNConv2d, QuantNLinear)

_CONV_TYPES = [nn.Conv2d, QConv2d, QuantConv2d, QuantNConv2d]
_LINEAR_TYPES = [nn.Linear, QLinear, QuantLinear, QuantNLinear]
_TENSOR_OPS = ['__add__', 'add', '__iadd__', 'cat', 'mean', 'interpolate', 'softmax']   # layer_transform.py:11-15


def find_prev_bn(bn_module, relu_attached, graph, bottoms, bot):
    """utils/layer_transform.py:308-353: breadth-first walk up from `bot` to the nearest BatchNorms.
    Returns (bn_list [(bn, branch id)], relu flags, connection types 'one'/'add[_relu]'/'cat', {branch: layer})."""
    queue = [(b, str(i)) for i, b in enumerate(bot)]
    kind = {str(i): 'one' for i in range(len(bot))}
    no_bn = {}
    bn_list, relu_list, kind_list = [], [], []
    joined = False
    while queue:
        key, bid = queue.pop(0)
        node = graph[key]
        if type(node) == str:
            if 'add' in key:
                kind[bid] = 'add_{}'.format(relu_attached[key]) if key in relu_attached else 'add'
                joined = True
            elif 'cat' in key:
                kind[bid] = 'cat'
                joined = True
        elif not joined and type(node) in _CONV_TYPES + _LINEAR_TYPES:
            print("Warning: {} layer before first batch norm layer detected. The calculated value range might "
                  "be off.".format(type(node)))
            if bid[0] in no_bn:
                assert False, "Multiple conv/linear layer without batch_norm is not supported."
            no_bn[bid[0]] = ("conv" if type(node) in _CONV_TYPES else "linear", node)
        if key not in bn_module:
            child = bid + bid[0]
            queue.extend((b, child) for b in bottoms[key])
            kind[child] = kind[bid]
        else:
            bn_list.append((bn_module[key], bid))
            relu_list.append(relu_attached[key])
            kind_list.append(kind[bid])
    return bn_list, relu_list, kind_list, no_bn


# ---- moments of a (clipped) rectified Gaussian, fp64 pdf/cdf rounded to fp32 like the reference (:413-431)
_SQRT_2PI = math.sqrt(2 * math.pi)


def _pdf(x):
    # scipy.stats.norm(0, 1).pdf(x) == exp(-x**2 / 2) / sqrt(2 pi) in fp64 (scipy/stats/_continuous_distns.py
    # _norm_pdf), without the per-call frozen-distribution set-up
    x = np.asarray(x, dtype=np.float64)
    return (np.exp(-x ** 2 / 2.0) / _SQRT_2PI).astype(np.float32)


def _cdf(x):
    # scipy.stats.norm.cdf(x) == scipy.special.ndtr(x) in fp64
    return ndtr(np.asarray(x, dtype=np.float64)).astype(np.float32)


def relu_mean(w, b):
    return w * _pdf(-b / w) + b * (1 - _cdf(-b / w))


def relu_var(w, b, m):
    return (1 - _cdf(-b / w)) * (b * b + w * w + m * m - 2 * m * b) + w * (b - 2 * m) * (_pdf(-b / w)) + \
        m * m * _cdf(-b / w)


def relu6_mean(w, b):
    return w * (_pdf(-b / w) - _pdf((6 - b) / w)) + b * (_cdf((6 - b) / w) - _cdf(-b / w)) + \
        6 * (1 - _cdf((6 - b) / w))


def relu6_var(w, b, m):
    return (_cdf((6 - b) / w) - _cdf(-b / w)) * (b * b + w * w + m * m - 2 * m * b) + \
        w * (-6) * _pdf((6 - b) / w) + w * (b - 2 * m) * (_pdf(-b / w) - _pdf((6 - b) / w)) + \
        m * m * _cdf(-b / w) + ((6 - m) ** 2) * (1 - _cdf((6 - b) / w))


def _hi(bias, weight, n):
    return float(np.max(bias + n * weight))


def _lo(bias, weight, n):
    return float(np.min(bias - n * weight))


_ctx = {"stats": None, "pending": None, "record": None}


class _LayerQuant:
    """`node.quant`, looked up when the range is written: set_layer_bits replaces the measures of every target layer
    (utils/quantize.py:373), and a cached walk must follow it."""
    __slots__ = ("node",)

    def __init__(self, node):
        self.node = node


def _bn_stats(bn):
    """(fake_bias, fake_weight) of a BatchNorm as fresh fp32 numpy vectors (one batched D2H copy per call of
    set_quant_minmax, see _gather_stats)."""
    cache = _ctx["stats"]
    if cache is not None and bn in cache:
        b, w = cache[bn]
        return b.copy(), w.copy()
    return bn.fake_bias.detach().cpu().numpy().copy(), bn.fake_weight.detach().cpu().numpy().copy()


class _Gather:
    """Device->host copy of every BatchNorm's (fake_bias, fake_weight) through one pinned buffer.  start() queues
    the copy on the current stream, fetch() waits for it and hands back {bn: (bias, weight)} as fp32 numpy views
    (plus `flat` / `spans` for the vectorised 1-to-1 ranges).  The span table is rebuilt only when the vectors move
    (merge_batchnorm's pooled stash is a new allocation per eager call, a fixed one inside a captured pipeline)."""

    def __init__(self, graph, bn_type):
        self.bns = [m for m in graph.values() if type(m) == bn_type and hasattr(m, "fake_bias")]
        self.key = None
        self.host = self.done = self.src = None
        self.spans, self.pooled_total = [], None

    def _layout(self):
        bns = self.bns
        first = bns[0]._buffers["fake_bias"]
        base = first._base
        key = (None if base is None else base.data_ptr(), first.data_ptr(), bns[-1]._buffers["fake_weight"].data_ptr())
        if key == self.key and base is not None:
            return
        vecs = [t for m in bns for t in (m._buffers["fake_bias"], m._buffers["fake_weight"])]
        if base is not None and base.is_contiguous() and all(t._base is base and t.is_contiguous() for t in vecs):
            # merge_batchnorm's pooled stash: the vectors are slices of one allocation
            self.src, b0 = base.detach().reshape(-1), base.storage_offset()
            self.spans = [(m, m._buffers["fake_bias"].storage_offset() - b0,
                           m._buffers["fake_weight"].storage_offset() - b0, m._buffers["fake_bias"].numel()) for m in bns]
            self.key = key
            # [all fake_weight | all fake_bias], BatchNorms back to back: per-BN extrema by two reduceat calls
            total = self.src.numel() // 2
            off, ok = 0, self.src.numel() == 2 * total
            for _m, ob, ow, n in self.spans:
                ok = ok and ow == off and ob == total + off
                off += n
            self.pooled_total = total if ok and off == total else None
        else:
            self.src = None                     # gathered per call (torch.cat)
            self.key = None
            self.pooled_total = None
            self.spans, off = [], 0
            for m in bns:
                n = m._buffers["fake_bias"].numel()
                self.spans.append((m, off, off + n, n))
                off += 2 * n

    def start(self):
        if not self.bns:
            return
        self._layout()
        src = self.src
        if src is None:
            src = torch.cat([t.detach().reshape(-1) for m in self.bns
                             for t in (m._buffers["fake_bias"], m._buffers["fake_weight"])])
        if src.is_cuda:
            if self.host is None or self.host.numel() != src.numel():
                self.host = torch.empty(src.numel(), dtype=src.dtype, pin_memory=True)
            self.host.copy_(src, non_blocking=True)
            if self.done is None:
                self.done = torch.cuda.Event()
            self.done.record()
            self._flat = None
        else:
            self._flat = src.numpy()

    def fetch(self):
        if not self.bns:
            self.flat = None
            return {}
        if self._flat is None:
            self.done.synchronize()
            self.flat = self.host.numpy()
        else:
            self.flat = self._flat
        flat = self.flat
        return {m: (flat[ob:ob + n], flat[ow:ow + n]) for m, ob, ow, n in self.spans}


def _start_gather(graph, bn_type):
    """Phase 1 (uncached form): queue the device->host copy of the BatchNorm statistics, return the fetch function."""
    g = _Gather(graph, bn_type)
    g.start()
    return g.fetch


def _clipped_range(bias, weight, relu, N):
    lo = max(0., _lo(bias, weight, N)) if 'relu' in relu else _lo(bias, weight, N)
    hi = min(6., _hi(bias, weight, N)) if 'relu6' in relu else _hi(bias, weight, N)
    return lo, hi


def _quant_modules_of(node, top):
    """The QuantMeasure(s) feeding `node` (:381-406): a layer's own `.quant`, or the next operand measures of the
    recorded tensor op when `node` is that op's graph name."""
    if type(node) == str:
        if top is not None and top.num_op and top.get_graph_name() == node:
            for op_name in (n.replace('_', '') for n in _TENSOR_OPS):
                if op_name in node:
                    count = int(top.get_module_name().split('_')[-1])
                    mods = [top.get_module_next() for _ in range(count)]
                    top.add_idx_name_tensor_op()
                    return mods
    elif hasattr(node, 'quant'):
        return [_LayerQuant(node)]
    return None


def _merge_branch(entries, N, eps):
    """One operand branch of a tensor op: fold its BatchNorms (deepest first) into either Gaussian moments
    ('add') or a (min, max) range ('cat' / 'one')  (:500-566)."""
    todo = sorted(entries, key=lambda e: len(e[0][1]), reverse=True)
    (bn, bid), relu, kind = todo.pop(0)
    depth = len(bid)
    bias, weight = _bn_stats(bn)
    if 'add' in kind:
        if relu == 'relu':
            mean = relu_mean(weight, bias)
            var = relu_var(weight, bias, mean)
        elif relu == 'relu6':
            mean = relu6_mean(weight, bias)
            var = relu6_var(weight, bias, mean)
        else:
            mean, var = bias, weight * weight
    else:
        lo, hi = _clipped_range(bias, weight, relu, N)
    while todo:
        bound = 0
        while bound < len(todo) and len(todo[bound][0][1]) == depth:
            bound += 1
        if bound == 0:
            depth = len(todo[0][0][1])
            continue
        for (bn_t, _bid), relu_t, kind in todo[:bound]:
            bias, weight = _bn_stats(bn_t)
            if 'add' in kind:
                if relu_t == 'relu':
                    m = relu_mean(weight, bias)
                    mean += m
                    var += relu_var(weight, bias, m)
                elif relu_t == 'relu6':
                    m = relu6_mean(weight, bias)
                    mean += m
                    var += relu6_var(weight, bias, m)
                else:
                    mean += bias
                    var += weight * weight
                if 'relu6' in kind:
                    prev = mean
                    mean = relu6_mean(np.sqrt(var + eps), mean)
                    var = relu6_var(np.sqrt(var + eps), prev, mean)
                elif 'relu' in kind:
                    prev = mean
                    mean = relu_mean(np.sqrt(var + eps), mean)
                    var = relu_var(np.sqrt(var + eps), prev, mean)
            elif kind == 'cat':
                l2, h2 = _clipped_range(bias, weight, relu_t, N)
                lo, hi = min(lo, l2), max(hi, h2)
            else:
                # `if use_relu_tmp` in the reference tests a non-empty string: always true (:553)
                lo += max(0., _lo(bias, weight, N)) if relu_t else _lo(bias, weight, N)
                hi += _hi(bias, weight, N)
        todo = todo[bound:]
        if kind == 'one':
            lo /= (bound + 1)
            hi /= (bound + 1)
    return (kind, mean, var) if 'add' in kind else (kind, lo, hi)


def _range_of(res, N, eps):
    if 'add' in res[0]:
        _, mean, var = res
        return _lo(mean, np.sqrt(var + eps), N), _hi(mean, np.sqrt(var + eps), N)
    return res[1], res[2]


def _set(qm, lo, hi):
    if _ctx["record"] is not None:
        _ctx["record"].append((qm, lo, hi))
    if type(qm) is _LayerQuant:
        qm = qm.node.quant
    if _ctx["pending"] is not None:
        _ctx["pending"][qm] = (lo, hi)
        return
    qm.running_max.fill_(hi)
    qm.running_min.fill_(lo)


_flush_cache = {}       # id(running_min) -> (running_min, running_max, table, row): layout of pooled measure buffers


def _flush_pending():
    """running_min.fill_(lo) / running_max.fill_(hi) for every QuantMeasure with ONE host->device copy per
    backing table.  Measures whose buffers already are rows of a shared [n, 2] (min, max) table (set_layer_bits
    pools them; an earlier call of this function does too) are written through that table; the others become
    [1]-views of a new table (contiguous, reused by the eval plan)."""
    pending, _ctx["pending"] = _ctx["pending"], None
    if not pending:
        return None
    pooled, loose = {}, []
    if len(_flush_cache) > 8192:
        _flush_cache.clear()
    for q, (lo, hi) in pending.items():
        rmin, rmax = q._buffers["running_min"], q._buffers["running_max"]
        hit = _flush_cache.get(id(rmin))
        if hit is not None and hit[0] is rmin and hit[1] is rmax:          # same buffer objects as last time
            pooled.setdefault(id(hit[2]), (hit[2], {}))[1][hit[3]] = (lo, hi)
            continue
        base = rmin._base
        if (base is not None and rmax._base is base and base.dim() == 2 and base.size(1) == 2
                and base.is_contiguous() and base.dtype == torch.float32
                and rmax.storage_offset() == rmin.storage_offset() + 1):
            row = (rmin.storage_offset() - base.storage_offset()) // 2
            pooled.setdefault(id(base), (base, {}))[1][row] = (lo, hi)
            _flush_cache[id(rmin)] = (rmin, rmax, base, row)
        else:
            loose.append(q)
    table = None
    uploads, complete = [], not loose
    for base, rows in pooled.values():
        if len(rows) == base.size(0):                        # every row rewritten: plain copy
            host = torch.tensor([rows[i] for i in range(base.size(0))], dtype=torch.float32)
            base.copy_(host)
            uploads.append((base, host))
        else:
            idx = sorted(rows)
            base.index_copy_(0, torch.tensor(idx, device=base.device),
                             torch.tensor([rows[i] for i in idx], dtype=torch.float32).to(base.device))
            complete = False
        table = base
    # (MinmaxPlan.speculate) what this call uploaded, when it was whole tables only
    _ctx["uploads"] = uploads if complete and uploads else None
    if loose:
        dev = loose[0]._buffers["running_min"].device
        table = torch.tensor([pending[q] for q in loose], dtype=torch.float32).to(dev)
        for i, q in enumerate(loose):
            q._buffers["running_min"] = table[i, 0:1]         # registered buffers: skip Module.__setattr__
            q._buffers["running_max"] = table[i, 1:2]
    return table


class MinmaxPlan:
    """The graph walk of set_quant_minmax done once (it depends on the graph structure only), the arithmetic per call.
    start() queues the device->host copy of the BatchNorm statistics; finish() waits for it, does the per-channel
    arithmetic and writes every range.  Ranges of the 1-to-1 case (:476-505: min(beta - N gamma), max(beta + N gamma)
    per BatchNorm) are formed for ALL BatchNorms by four vector operations over the pooled statistics -- the same
    fp32 operations element by element -- and two reduceat calls."""

    def __init__(self, graph, bottoms, is_detection=False, bn_type=torch.nn.BatchNorm2d, N=6):
        from .. import graph as G
        self.graph, self.bottoms = graph, bottoms
        self.top = G._state.get("module_tensor_op")
        self.key = self._key(graph, bottoms, is_detection, bn_type, N, self.top)
        self.N, self.eps = N, 1e-6
        self.jobs = _walk(graph, bottoms, is_detection, bn_type, self.top)
        self.gather = _Gather(graph, bn_type)
        self.bn_pos = {m: i for i, m in enumerate(self.gather.bns)}
        self._memo = None
        self._spec = self._spec_key = None        # uploads of the memoised ranges: [(table, pinned rows)], buffer ids
        self._spec_done = False

    @staticmethod
    def _key(graph, bottoms, is_detection, bn_type, N, top):
        return (id(graph), id(bottoms), tuple(map(id, graph.values())), bool(is_detection), bn_type, N, id(top))

    def start(self):
        self.gather.start()

    @staticmethod
    def _buffers_key(rec):
        try:
            return tuple(id((q.node.quant if type(q) is _LayerQuant else q)._buffers["running_min"]) for q, _lo, _hi in rec)
        except (AttributeError, KeyError):
            return None

    def speculate(self):
        """Queue the upload of the ranges the PREVIOUS call computed, before the statistics of this call have arrived:
        the ranges are a pure function of the BatchNorm statistics, and in the reference's flow those are the same on
        every pass.  finish() still compares the fetched statistics with the memoised ones and recomputes + uploads
        again when they differ, so the speculation only removes the host round trip (wait for the copy, compare,
        upload) from the device's critical path."""
        self._spec_done = False
        if self._memo is None or self._spec is None or not torch.cuda.is_available():
            return
        if self._spec_key is None or self._buffers_key(self._memo[1]) != self._spec_key:   # measures re-bound: no shortcut
            self._spec = None
            return
        for base, host in self._spec:
            base.copy_(host, non_blocking=True)
        self._spec_done = True

    def _remember_uploads(self):
        ups = _ctx.pop("uploads", None)
        if ups and self._memo is not None and all(b.is_cuda for b, _h in ups):
            self._spec = [(b, h.pin_memory()) for b, h in ups]
            self._spec_key = self._buffers_key(self._memo[1])
        else:
            self._spec = self._spec_key = None

    def finish(self):
        stats = self.gather.fetch()
        # The ranges are a pure function of the BatchNorm statistics: when the fetched statistics are bit-identical to
        # the previous call's (always, in the reference's flow: the second merge_batchnorm resets them to |1| / 0,
        # SURVEY.md 0.3), the recorded (measure, min, max) list is written again without redoing the arithmetic.
        flat = self.gather.flat
        if flat is not None and self._memo is not None and flat.shape == self._memo[0].shape and \
                np.array_equal(flat.view(np.uint32), self._memo[0].view(np.uint32)):
            if self._spec_done:                     # speculate() already queued exactly these ranges
                self._spec_done = False
                return
            _ctx["pending"] = {}
            try:
                for qref, lo, hi in self._memo[1]:
                    _set(qref, lo, hi)
            finally:
                _flush_pending()
                self._remember_uploads()
            return
        self._spec_done = False
        _ctx["stats"], _ctx["pending"] = stats, {}
        _ctx["record"] = rec = []
        N, eps = self.N, self.eps
        try:
            with np.errstate(all="ignore"):              # x/0 -> inf, silently, as torch does
                lo_bn = hi_bn = None
                total = self.gather.pooled_total
                if total:
                    flat = self.gather.flat
                    w_all, b_all = flat[:total], flat[total:2 * total]
                    starts = np.fromiter((ow for _m, _ob, ow, _n in self.gather.spans), dtype=np.intp)
                    nw = N * w_all
                    lo_bn = np.minimum.reduceat(b_all - nw, starts).tolist()
                    hi_bn = np.maximum.reduceat(b_all + nw, starts).tolist()
                for job in self.jobs:
                    if job[0] == "one" and lo_bn is not None and not job[4]:
                        _, qms, bn_list, relu_list, _no_bn = job
                        for i, (bn, _bid) in enumerate(bn_list):
                            k = self.bn_pos[bn]
                            lo, hi, relu = lo_bn[k], hi_bn[k], relu_list[i]
                            if 'relu' in relu:
                                lo = max(0., lo)
                            if 'relu6' in relu:
                                hi = min(6., hi)
                            _set(qms[i], lo, hi)
                    else:
                        _run_job(job, N, eps)
            self._memo = (flat.copy(), rec) if flat is not None else None
        finally:
            _ctx["stats"] = None
            _ctx["record"] = None
            _flush_pending()
            self._remember_uploads()


_plans = {}


def minmax_plan(graph, bottoms, is_detection=False, bn_type=torch.nn.BatchNorm2d, N=6):
    """Cached MinmaxPlan of this graph (rebuilt when the graph's nodes, the bottoms or the tensor-op table change)."""
    from .. import graph as G
    key = MinmaxPlan._key(graph, bottoms, is_detection, bn_type, N, G._state.get("module_tensor_op"))
    plan = _plans.get(id(graph))
    if plan is None or plan.key != key or plan.graph is not graph:
        if len(_plans) >= 8:
            _plans.clear()
        plan = _plans[id(graph)] = MinmaxPlan(graph, bottoms, is_detection, bn_type, N)
    return plan


def set_quant_minmax(graph, bottoms, is_detection=False, bn_type=torch.nn.BatchNorm2d, N=6, verbose=True):
    """utils/layer_transform.py:356-618: fill every QuantMeasure's running_min / running_max from the statistics
    (fake_weight = |gamma|, fake_bias = beta) of the BatchNorm(s) that feed it.

    Three phases so that the host work overlaps the kernels still queued on the stream (the BN statistics are
    the last thing the preceding steps write): (1) start the device->host copy of the statistics, (2) the graph
    walk -- which BatchNorms / ReLUs / tensor ops feed which QuantMeasure, pure graph logic, cached per graph --
    (3) wait for the copy and do the per-channel arithmetic (MinmaxPlan)."""
    if verbose:
        print("SET QUANT MIN MAX")
    cached = _plans.get(id(graph))
    if cached is not None and cached.graph is graph and cached.bottoms is bottoms:
        plan = minmax_plan(graph, bottoms, is_detection, bn_type, N)
        plan.start()
    else:                                   # first call on this graph: copy in flight while the walk runs
        g = _Gather(graph, bn_type)
        g.start()
        plan = minmax_plan(graph, bottoms, is_detection, bn_type, N)
        plan.gather = g
        plan.bn_pos = {m: i for i, m in enumerate(g.bns)}
    plan.finish()


def _walk(graph, bottoms, is_detection, bn_type, top):
    """Phase 2: [(kind, ...)] jobs in graph order; no BatchNorm value is read here."""
    bn_module, relu_attached = {}, {}
    jobs = []
    for key in graph:
        node, bot = graph[key], bottoms[key]
        if bot is None:
            continue
        if type(node) == bn_type:
            bn_module[key] = node
            relu_attached[key] = 'none'
            continue
        if type(node) == torch.nn.ReLU:
            relu_attached[bot[0]] = 'relu'
        elif type(node) == torch.nn.ReLU6:
            relu_attached[bot[0]] = 'relu6'
        qms = _quant_modules_of(node, top)
        if len(bot) == 1 and bot[0] == 'Data':
            if is_detection:
                jobs.append(("const", qms[0], -1, 1))
            else:                                   # (0 - mean)/std .. (1 - mean)/std of the data preprocessing
                jobs.append(("const", qms[0], -2.11790393, 2.64))
            continue
        if qms is None:
            continue
        bn_list, relu_list, kind_list, no_bn = find_prev_bn(bn_module, relu_attached, graph, bottoms, bot[:])
        if len(qms) == len(bn_list):                # 1 to 1 (:476-505)
            # relu flags are snapshotted: a later ReLU node must not leak into this job
            jobs.append(("one", qms, bn_list, list(relu_list), no_bn))
            continue
        branches = {}
        for i, item in enumerate(bn_list):
            branches.setdefault(item[1][0], []).append((item, relu_list[i], kind_list[i]))
        if len(qms) == 1 and len(qms) < len(bn_list):          # 1 to many
            assert len(branches) == 1, "Error occurs when setting min/max, should be 1 to many"
        elif len(qms) < len(bn_list):                          # many to many
            assert len(branches) == len(qms), 'LENGTH NOT EQUAL {} vs {}'.format(len(branches), len(qms))
        else:
            assert False, "Unknown error occured while setting min/max"
        jobs.append(("branches", qms, branches))
    return jobs


def _run_job(job, N, eps):
    """Phase 3: the arithmetic of one QuantMeasure (or of the measures of one tensor op)."""
    if job[0] == "const":
        _set(job[1], job[2], job[3])
    elif job[0] == "one":
        _, qms, bn_list, relu_list, no_bn = job
        for i, (bn, bid) in enumerate(bn_list):
            bias, weight = _bn_stats(bn)
            if bid[0] in no_bn:                 # case (d): a conv/linear without BatchNorm in between
                layer_type, layer = no_bn[bid[0]]
                bias, weight = torch.from_numpy(bias), torch.from_numpy(weight)
                lw, lb = layer.weight.detach().cpu(), layer.bias.detach().cpu()
                if layer_type == 'conv':
                    lw = lw.view(lw.size(0), lw.size(1), -1).sum(-1)
                    lw = lw.view(lw.size(0), lw.size(1), 1, 1)
                    bias = torch.nn.functional.conv2d(bias.view(1, -1, 1, 1), lw, lb, 1, 0, 1, layer.groups)
                    weight = torch.nn.functional.conv2d(weight.view(1, -1, 1, 1), lw, lb, 1, 0, 1, layer.groups)
                else:
                    bias = torch.nn.functional.linear(bias.view(1, -1), lw, lb)
                    weight = torch.nn.functional.linear(weight.view(1, -1), lw, lb)
                bias, weight = bias.numpy(), weight.numpy()
                lo, hi = _lo(bias, weight, N), _hi(bias, weight, N)
            else:
                lo, hi = _clipped_range(bias, weight, relu_list[i], N)
            _set(qms[i], lo, hi)
    else:
        _, qms, branches = job
        res = {b: _merge_branch(entries, N, eps) for b, entries in branches.items()}
        if len(qms) == 1:
            _set(qms[0], *_range_of(list(res.values())[0], N, eps))
        else:
            for i in range(len(res)):
                _set(qms[i], *_range_of(res[str(i)], N, eps))
