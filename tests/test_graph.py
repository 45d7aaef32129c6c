"""CPU: the fx-based graph builder yields the reference tracer's graph / bottoms / relations (golden json),
and set_quant_minmax reproduces the reference's activation ranges."""
import json
import os

import numpy as np
import torch

from conftest import ROOT
from helpers import fresh_model, names_of, switch


def _named(model, tr):
    names = names_of(model)
    kn = lambda k: k if isinstance(k, str) else names[k]
    g, b = tr.log.getGraph(), tr.log.getBottoms()
    return [kn(k) for k in g], {kn(k): (None if v is None else [kn(x) for x in v]) for k, v in b.items()}, kn


def test_graph_and_bottoms_match_reference_tracer(gold):
    model, _ = fresh_model()
    model, tr, _ = switch(model)
    g, b, _ = _named(model, tr)
    assert g == gold["graph"]
    assert b == gold["bottoms"]
    assert len(tr.log.getRecordTensorOP()) == 11          # 10 residual adds + 1 mean
    assert len(list(model.custom_tensor_op.children())) == 21


def test_relations_match_reference(gold):
    from data_free_quantization_b200.utils.relation import create_relation
    model, _ = fresh_model()
    model, tr, targ = switch(model)
    _, _, kn = _named(model, tr)
    rels = create_relation(tr.log.getGraph(), tr.log.getBottoms(), targ)
    mine = [[kn(a), kn(b), kn(c)] for a, b, c in (r.get_idxs() for r in rels)]
    assert mine == [r[:3] for r in gold["topo"]["relations"]]
    from data_free_quantization_b200.bias_absorption import _is_relu_between
    relu = [_is_relu_between(r.layer_second, r.layer_first, tr.log.getGraph(), tr.log.getBottoms()) for r in rels]
    assert relu == [r[3] for r in gold["topo"]["relations"]]


def test_cle_stage_assignment_keeps_reference_order():
    from data_free_quantization_b200.Cross_layer_equal import _stages
    from data_free_quantization_b200.utils.relation import Relation
    rels = [Relation("a", "b", "bn_a"), Relation("b", "c", "bn_b"), Relation("x", "y", "bn_x"),
            Relation("c", "d", "bn_c"), Relation("y", "z", "bn_y")]
    assert _stages(rels) == [0, 1, 0, 2, 1]


def test_set_quant_minmax_as_wired(gold):
    """main_dfq.py:211 resets fake_weight/fake_bias to 1/0 before set_quant_minmax: ranges from the golden run."""
    from data_free_quantization_b200.utils.minmax import set_quant_minmax
    model, _ = fresh_model()
    model, tr, targ = switch(model)
    graph, bottoms = tr.log.getGraph(), tr.log.getBottoms()
    for m in model.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.register_buffer("fake_weight", torch.ones(m.num_features))
            m.register_buffer("fake_bias", torch.zeros(m.num_features))
    set_quant_minmax(graph, bottoms, verbose=False)
    names = names_of(model)
    for k, m in graph.items():
        if type(m) in targ:
            lo, hi = gold["quant_minmax"][names[k]]
            assert float(m.quant.running_min) == lo and float(m.quant.running_max) == hi, names[k]


def test_set_quant_minmax_random_bn_stats():
    """Non-trivial BN statistics: fixture generated from the reference by oracle/gen_minmax_golden.py."""
    path = os.path.join(ROOT, "tests", "golden", "minmax_random_bn.json")
    fx = json.load(open(path))
    from data_free_quantization_b200.utils.minmax import set_quant_minmax
    model, _ = fresh_model()
    model, tr, targ = switch(model)
    g = torch.Generator().manual_seed(fx["seed"])
    for name, m in model.named_modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.register_buffer("fake_weight", torch.rand(m.num_features, generator=g) * 1.5 + 0.25)
            m.register_buffer("fake_bias", torch.randn(m.num_features, generator=g) * 0.5)
    set_quant_minmax(tr.log.getGraph(), tr.log.getBottoms(), verbose=False)
    got = {n: [float(m.running_min), float(m.running_max)] for n, m in model.named_modules()
           if type(m).__name__ == "QuantMeasure"}
    assert set(got) == set(fx["ranges"])
    for n, (lo, hi) in fx["ranges"].items():
        np.testing.assert_allclose(got[n], [lo, hi], rtol=1e-6, atol=1e-7, err_msg=n)
