"""Cross-layer equalisation on B200 -- host mirror of the reference's Cross_layer_equal.py.

`cross_layer_equalization` keeps the reference signature (Cross_layer_equal.py:63) and runs the whole
`while diff > Treshhold and iter_count < Count` loop in ONE cooperative persistent launch
(csrc/cle.cu, dfq_cle_run).  Relations that touch a common tensor are ordered exactly as the reference's
sequential loop orders them (they get increasing `stage` numbers); relations that share nothing run in
the same stage, in parallel.
"""
import ctypes
import os

import torch
import torch.nn as nn

from . import _lib as L

DEBUG_STAMPS = None       # tuning aid: an int64 CUDA tensor [blocks * phases * 3] receives per-phase timer stamps
MAX_ITERS = 8192          # the reference loop is unbounded; MobileNetV2 converges in 44 iterations


def _ensure_bias(layer):
    if layer.bias is None:   # Cross_layer_equal.py:93-94
        layer.bias = nn.Parameter(data=torch.zeros((layer.weight.size(0)), dtype=torch.float32,
                                                   device=layer.weight.device), requires_grad=False)


def _stages(relations):
    """stage[r] = 1 + max stage of every earlier relation that touches one of r's tensors."""
    last_use = {}
    stages = []
    for rel in relations:
        first, second, bn = rel.get_idxs()
        keys = (first, second, ("bn", bn))
        st = max((last_use[k] + 1 for k in keys if k in last_use), default=0)
        for k in keys:
            last_use[k] = st
        stages.append(st)
    return stages


def _w2_dims(w):
    O2, I2 = w.shape[0], w.shape[1]
    return O2, I2, w.numel() // (O2 * I2)


def cross_layer_equalization(graph, relations, Target_list, s_min_max=[1e-8, 1e8], Treshhold=2e-7, Count=20,
                             signed=False, eps=0, Save_state=True, max_iters=MAX_ITERS, return_info=False):
    """Cross_layer_equal.py:63-115.  Weights, biases, BN fake_weight/fake_bias and Relation.S are updated in
    place on the GPU.  Save_state (per-iteration weight plots) is a visualisation feature and is ignored."""
    print("Cross layer equalization")
    if len(relations) == 0:
        return {"iters": 0, "diffs": []} if return_info else None
    with torch.no_grad():
        stages = _stages(relations)
        n = len(relations)
        pairs = (L.DfqClePair * n)()
        keep = []
        metric = OrderedDictNoDup()
        dev = None
        # one fill for every missing bias / scale vector instead of a tiny launch per relation
        dev = graph[relations[0].get_idxs()[0]].weight.device
        need_b = [graph[r.get_idxs()[0]] for r in relations if graph[r.get_idxs()[0]].bias is None]
        need_b = list({id(m): m for m in need_b}.values())
        need_s = [r for r in relations if r.S is None]
        zeros = torch.zeros(sum(m.weight.size(0) for m in need_b), dtype=torch.float32, device=dev)
        ones = torch.ones(sum(graph[r.get_idxs()[0]].weight.size(0) for r in need_s), dtype=torch.float32, device=dev)
        off = 0
        for m in need_b:                            # Cross_layer_equal.py:93-94
            c = m.weight.size(0)
            m.bias = nn.Parameter(data=zeros[off:off + c], requires_grad=False)
            off += c
        off = 0
        for r in need_s:
            c = graph[r.get_idxs()[0]].weight.size(0)
            r.S = ones[off:off + c]
            off += c
        for i, rel in enumerate(relations):
            first, second, bn = rel.get_idxs()
            l1, l2, bnm = graph[first], graph[second], graph[bn]
            _ensure_bias(l1)
            dev = l1.weight.device
            W1, W2 = l1.weight.data, l2.weight.data
            C1 = W1.shape[0]
            O2, I2, KK = _w2_dims(W2)
            if rel.S is None:
                rel.S = torch.ones(C1, dtype=torch.float32, device=dev)
            elif not rel.S.is_cuda:
                rel.S = rel.S.to(dev)
            p = pairs[i]
            p.W1, p.W2 = L.dev_f32(W1, "layer_first.weight"), L.dev_f32(W2, "layer_second.weight")
            p.b1 = L.dev_f32(l1.bias.data, "layer_first.bias")
            p.bn_w = L.dev_f32(bnm.fake_weight, "fake_weight")
            p.bn_b = L.dev_f32(bnm.fake_bias, "fake_bias")
            p.S = L.dev_f32(rel.S, "Relation.S")
            p.C1, p.K1, p.O2, p.I2, p.KK, p.stage = C1, W1.numel() // C1, O2, I2, KK, stages[i]
            metric.add(first, W1)
            metric.add(second, W2)
            keep.append((W1, W2))
        # layers outside every relation never change: their mean|dW| term is exactly 0 (:107)
        ns, tu = ctypes.c_int(), ctypes.c_int()
        L.check(L.lib().dfq_cle_plan_host(pairs, n, ctypes.byref(ns), ctypes.byref(tu)), "dfq_cle_plan_host")
        ws = metric.values()
        layers = (L.DfqCleLayer * len(ws))()
        scratch = torch.empty(sum(w.numel() for w in ws), dtype=torch.float32, device=dev)
        off = 0
        for i, w in enumerate(ws):
            layers[i].W, layers[i].n = w.data_ptr(), w.numel()
            layers[i].W_old = scratch.data_ptr() + 4 * off
            off += w.numel()
        tc = ctypes.c_int()
        L.check(L.lib().dfq_cle_layers_plan_host(layers, len(ws), ctypes.byref(tc)), "dfq_cle_layers_plan_host")
        pairs_dev = L.table_to_device(pairs, dev)
        acc = torch.empty(2 * len(ws), dtype=torch.float64, device=dev)
        nbytes = int(L.lib().dfq_cle_result_bytes(max_iters + 1))
        result = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
        ev0 = ev1 = None
        if return_info:                 # device time of the loop itself (CUDA events on the launch stream)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            _EVENTS[:] = [ev0, ev1]
        want = os.environ.get("DFQ_CLE_IMPL", "resident")
        if want == "resident" and _run_resident(pairs, n, layers, len(ws), s_min_max, signed, eps, Treshhold, Count,
                                                max_iters, result, dev):
            impl = "resident"
        elif want in ("resident", "stream") and _run_streaming(
                pairs, n, pairs_dev, layers, len(ws), s_min_max, signed, eps, Treshhold, Count, max_iters, acc,
                result, dev):
            impl = "stream"
        else:
            impl = "units"
        if impl == "units":
            layers_dev = L.table_to_device(layers, dev)
            _tick(0)
            L.check(L.lib().dfq_cle_run(pairs_dev.data_ptr(), n, ns.value, layers_dev.data_ptr(), len(ws), tc.value,
                                        float(s_min_max[0]), float(s_min_max[1]), int(bool(signed)), float(eps),
                                        float(Treshhold), int(Count), int(max_iters), acc.data_ptr(),
                                        result.data_ptr(), L.stream()), "dfq_cle_run")
            _tick(1)
        _EVENTS[:] = []
        if not return_info:
            return None
        host = result.cpu().numpy()
        iters = int(host[:4].view("int32")[0])
        diffs = host[8:].view("float64")
        return {"iters": iters, "stall": int(host[4:8].view("int32")[0]), "diff": float(diffs[0]),
                "diffs": diffs[1:1 + iters].tolist(), "stages": ns.value, "impl": impl,
                "kernel_ms": ev0.elapsed_time(ev1)}


def _run_resident(pairs, n, layers, n_layers, s_min_max, signed, eps, Treshhold, Count, max_iters, result, dev):
    """The shared-memory-resident form (csrc/cle3.cu).  Returns False when the plan says cudaErrorNotSupported (not
    simple chains of ungrouped / depthwise layers, or the chains do not fit the chip's shared memory)."""
    lib = L.lib()
    tensors = (L.DfqCle3Tensor * n_layers)()
    chains = (L.DfqCle3Chain * n_layers)()
    plan = L.DfqCle3Plan()
    err = lib.dfq_cle3_plan_host(pairs, n, layers, n_layers, 0, 0, tensors, chains, ctypes.byref(plan))
    if err == 801:                                  # cudaErrorNotSupported
        return False
    L.check(err, "dfq_cle3_plan_host")
    tensors_dev = L.table_to_device(tensors, dev)
    chains_dev = L.table_to_device((L.DfqCle3Chain * plan.n_chains).from_buffer(chains), dev)
    scratch = torch.empty(int(lib.dfq_cle3_scratch_bytes(ctypes.byref(plan))), dtype=torch.uint8, device=dev)
    _tick(0)
    L.check(lib.dfq_cle3_run(tensors_dev.data_ptr(), chains_dev.data_ptr(), ctypes.byref(plan), scratch.data_ptr(),
                             float(s_min_max[0]), float(s_min_max[1]), int(bool(signed)), float(eps),
                             float(Treshhold), int(Count), int(max_iters), result.data_ptr(),
                             None if DEBUG_STAMPS is None else DEBUG_STAMPS.data_ptr(), L.stream()),
            "dfq_cle3_run")
    _tick(1)
    global LAST_PLAN
    LAST_PLAN = {"grid": plan.grid, "smem_bytes": plan.smem_bytes, "chains": [(chains[i].len, chains[i].ncta)
                                                                               for i in range(plan.n_chains)]}
    return True


LAST_PLAN = None
_EVENTS = []           # [start, stop] CUDA events of a return_info call: recorded right around the kernel launch(es)


def _tick(i):
    if _EVENTS:
        _EVENTS[i].record()


def _run_streaming(pairs, n, pairs_dev, layers, n_layers, s_min_max, signed, eps, Treshhold, Count, max_iters, acc,
                   result, dev):
    """The streaming form (csrc/cle2.cu).  Returns False when the relations are not simple chains (the plan says
    cudaErrorNotSupported); the caller then runs the per-channel-unit kernel."""
    lib = L.lib()
    tensors = (L.DfqCleTensor * n_layers)()
    max_tasks = 3 * sum((int(layers[i].n) // 512) + 8 for i in range(n_layers)) + 3 * sum(
        int(pairs[i].C1) // 8 + int(pairs[i].O2) // 8 + 8 for i in range(n))
    tasks = (L.DfqCleTask * max_tasks)()
    nt, ni, nph = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    phase_begin = (L.i32 * 65)()
    chan_off = (L.i32 * (n + 1))()
    err = lib.dfq_cle2_plan_host(pairs, n, layers, n_layers, tensors, tasks, max_tasks, ctypes.byref(nt),
                                 ctypes.byref(ni), phase_begin, ctypes.byref(nph), chan_off)
    if err == 801:                                  # cudaErrorNotSupported
        return False
    L.check(err, "dfq_cle2_plan_host")
    used = nt.value + ni.value
    tasks_dev = L.table_to_device((L.DfqCleTask * used).from_buffer(tasks), dev)
    tensors_dev = L.table_to_device(tensors, dev)
    chan_dev = L.table_to_device(chan_off, dev)
    total_c = int(chan_off[n])
    r1 = torch.empty(2 * total_c, dtype=torch.float32, device=dev)
    r2 = torch.empty(4 * total_c, dtype=torch.int32, device=dev)
    fat = torch.empty(int(lib.dfq_cle2_fat_bytes(used)) // 16 + 1, 4, dtype=torch.int32, device=dev)   # 16 B aligned
    _tick(0)
    L.check(lib.dfq_cle2_run(pairs_dev.data_ptr(), n, tensors_dev.data_ptr(), n_layers, tasks_dev.data_ptr(), nt.value,
                             ni.value, phase_begin, nph.value, chan_dev.data_ptr(), total_c, r1.data_ptr(),
                             r2.data_ptr(), float(s_min_max[0]), float(s_min_max[1]), int(bool(signed)), float(eps),
                             float(Treshhold), int(Count), int(max_iters), acc.data_ptr(), result.data_ptr(),
                             int(os.environ.get("DFQ_CLE_BLOCKS_PER_SM", "2")), fat.data_ptr(),
                             None if DEBUG_STAMPS is None else DEBUG_STAMPS.data_ptr(), L.stream()), "dfq_cle2_run")
    _tick(1)
    if DEBUG_STAMPS is not None:
        DEBUG_STAMPS.n_phases = nph.value
        DEBUG_STAMPS.tasks = [(tasks[i].tensor, tasks[i].row0, tasks[i].nrows, tasks[i].kind & 255, tasks[i].kind >> 8,
                               tensors[tasks[i].tensor].O, tensors[tasks[i].tensor].row_len,
                               tensors[tasks[i].tensor].up, tensors[tasks[i].tensor].down) for i in range(used)]
    return True


class OrderedDictNoDup:
    """Insertion-ordered set of weight tensors keyed by graph key."""

    def __init__(self):
        self._d = {}

    def add(self, key, w):
        self._d.setdefault(key, w)

    def values(self):
        return list(self._d.values())
