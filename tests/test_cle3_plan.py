"""CPU: the host planner of the shared-memory-resident CLE kernel (dfq_cle3_plan_host, csrc/cle3.cu) on the relation
sets of the two fixture models (tests/golden/*.json topologies, fake device addresses -- planning only)."""
import ctypes
import json
import os

import pytest

from conftest import ROOT


def _tables(model, gold_path):
    from data_free_quantization_b200 import _lib as L
    topo = json.load(open(gold_path))["topo"]
    mods = dict(model.named_modules())
    rels = topo["relations"]
    pairs = (L.DfqClePair * len(rels))()
    addr, base = {}, [0x10000000]

    def A(name, nbytes):
        if name not in addr:
            addr[name] = base[0]
            base[0] += (nbytes + 255) // 256 * 256
        return addr[name]
    ws, last = {}, {}
    for i, (a, b, bn, _relu) in enumerate(rels):
        W1, W2 = mods[a].weight, mods[b].weight
        p = pairs[i]
        p.W1, p.W2 = A(a, W1.numel() * 4), A(b, W2.numel() * 4)
        C1 = W1.shape[0]
        p.b1, p.bn_w, p.bn_b, p.S = A(a + ".b", 4 * C1), A(bn + ".w", 4 * C1), A(bn + ".b", 4 * C1), A(f"S{i}", 4 * C1)
        p.C1, p.K1, p.O2, p.I2 = C1, W1.numel() // C1, W2.shape[0], W2.shape[1]
        p.KK = W2.numel() // (W2.shape[0] * W2.shape[1])
        keys = (a, b, ("bn", bn))
        st = max((last[k] + 1 for k in keys if k in last), default=0)
        for k in keys:
            last[k] = st
        p.stage = st
        ws.setdefault(a, W1)
        ws.setdefault(b, W2)
    layers = (L.DfqCleLayer * len(ws))()
    for i, (k, w) in enumerate(ws.items()):
        layers[i].W, layers[i].n = addr[k], w.numel()
    return pairs, layers, ws


def _plan(pairs, layers, n_sm=148, smem=232448):
    from data_free_quantization_b200 import _lib as L
    n = len(layers)
    tensors, chains, plan = (L.DfqCle3Tensor * n)(), (L.DfqCle3Chain * n)(), L.DfqCle3Plan()
    err = L.lib().dfq_cle3_plan_host(pairs, len(pairs), layers, n, n_sm, smem, tensors, chains, ctypes.byref(plan))
    return err, tensors, chains, plan


@pytest.mark.parametrize("which", ["mobilenetv2", "deeplab"])
def test_resident_plan_covers_every_row_and_fits(which):
    if which == "mobilenetv2":
        from data_free_quantization_b200.modeling.mobilenetv2 import mobilenet_v2
        model = mobilenet_v2(None)
    else:
        from data_free_quantization_b200.modeling.deeplab import DeepLab
        model = DeepLab()
    pairs, layers, ws = _tables(model, os.path.join(ROOT, "tests", "golden", f"{which}_dfq.json"))
    err, tensors, chains, plan = _plan(pairs, layers)
    assert err == 0
    assert plan.n_tensors == len(ws) and 1 <= plan.grid <= 148 and plan.smem_bytes <= 232448 - 4096
    assert sum(chains[c].ncta for c in range(plan.n_chains)) == plan.grid
    seen = set()
    for c in range(plan.n_chains):
        ch = chains[c]
        assert ch.cta0 == sum(chains[k].ncta for k in range(c))
        for i in range(ch.len):
            T = tensors[ch.t0 + i]
            seen.add(T.layer)
            assert T.O * T.L == layers[T.layer].n and T.G == ch.ncta
            # consecutive tensors of a chain are the W1 / W2 of one pair
            if i + 1 < ch.len:
                assert T.down >= 0 and tensors[ch.t0 + i + 1].up == T.down
            else:
                assert T.down < 0
            # the row split used by the kernel covers [0, O) without gaps
            edges = [g * T.O // ch.ncta for g in range(ch.ncta + 1)]
            assert edges[0] == 0 and edges[-1] == T.O and all(b >= a for a, b in zip(edges, edges[1:]))
            # per-CTA footprint (data + per-row vectors) within the planned shared memory
        for g in range(ch.ncta):
            floats = 0
            for i in range(ch.len):
                T = tensors[ch.t0 + i]
                nr = (g + 1) * T.O // ch.ncta - g * T.O // ch.ncta
                floats += (nr * T.L + 3) // 4 * 4 + (4 * nr if T.down >= 0 else 0)
            assert floats * 4 + 4 * (plan.fvec_floats + plan.cp_sm_words + 5 * plan.max_rows +
                                   2 * (plan.dp_doubles + 3 * plan.n_tensors) + 8) + 16 <= plan.smem_bytes
    assert seen == set(range(len(ws)))
    if which == "mobilenetv2":      # the classifier chain (2.0 of 3.5 M weights) gets most of the chip
        assert max(chains[c].ncta for c in range(plan.n_chains)) >= 64


def test_resident_plan_refuses_what_it_cannot_hold():
    from data_free_quantization_b200.modeling.mobilenetv2 import mobilenet_v2
    pairs, layers, _ = _tables(mobilenet_v2(None), os.path.join(ROOT, "tests", "golden", "mobilenetv2_dfq.json"))
    assert _plan(pairs, layers, n_sm=148, smem=48 * 1024)[0] == 801       # 13.9 MB do not fit 148 x 48 KB
    assert _plan(pairs, layers, n_sm=8)[0] == 801                          # fewer SMs than chains
    pairs[3].W2 = pairs[1].W2                                              # a tensor that is W2 of two relations
    assert _plan(pairs, layers)[0] == 801
