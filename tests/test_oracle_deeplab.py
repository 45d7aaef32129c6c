"""CPU: BASELINE.json configs[3] without a GPU -- the C oracle reproduces the DeepLabV3+ golden digests taken from
the UNMODIFIED reference (tests/golden/deeplab_dfq.*), and the trace-free graph builder produces the reference's
graph (201 nodes, 61 target layers, 35 relations) for our alias-free twin of the model."""
import json
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

from helpers import np_state, rel_err, sha

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def dgold():
    with open(os.path.join(GOLD, "deeplab_dfq.json")) as f:
        return json.load(f), np.load(os.path.join(GOLD, "deeplab_dfq.npz"))


def _model():
    from data_free_quantization_b200.fixtures import synthetic_state_dict
    from data_free_quantization_b200.modeling.deeplab import DeepLab
    torch.manual_seed(0)
    model = DeepLab()
    sd = synthetic_state_dict(model, seed=0, bn_seed=1)
    model.load_state_dict(sd)
    return model.eval(), sd


def test_graph_matches_reference(dgold):
    from data_free_quantization_b200 import graph as G
    from data_free_quantization_b200.utils import quantize as Q
    from data_free_quantization_b200.utils.relation import create_relation
    gold, _ = dgold
    model, _ = _model()
    targ = [Q.QuantNConv2d, Q.QuantNLinear]
    md = {0: [(nn.ReLU6, nn.ReLU)], 1: [(nn.Conv2d, targ[0]), (nn.Linear, targ[1])]}
    model, tr = G.switch_layers(model, G.TorchTransformer(), torch.ones(1, 3, 65, 65), md,
                                ignore_layer=[Q.QuantMeasure], quant_op=True)
    graph, bottoms = tr.log.getGraph(), tr.log.getBottoms()
    names = {id(m): n for n, m in model.named_modules()}
    assert len(graph) == len(gold["graph"]) == 201
    # module nodes carry the reference's names in the reference's order (tensor-op keys differ in their position
    # suffix only: PyTransformer numbers them by call count, the fx builder by node index)
    mine = [names[k] for k in graph if not isinstance(k, str)]
    ref = [n for n in gold["graph"] if n in set(mine)]
    assert mine == ref
    assert [names[k] for k, m in graph.items() if type(m) in targ] == gold["topo"]["targets"]
    rels = create_relation(graph, bottoms, targ)
    assert [[names[a], names[b], names[c]] for a, b, c in (r.get_idxs() for r in rels)] == \
        [r[:3] for r in gold["topo"]["relations"]]
    ops = [k.split("_")[0] for k in graph if isinstance(k, str) and k != "Data"]
    assert ops.count("add") == 10 and ops.count("F.interpolate") == 3 and ops.count("torch.cat") == 2


def test_oracle_pipeline_reproduces_reference_digests(dgold, oracle):
    from oracle import pipeline as P
    gold, npz = dgold
    _, sd = _model()
    st = np_state(sd)
    assert sha(np.concatenate([v.ravel() for v in st.values()])) == gold["init_digest"]
    topo = {"targets": gold["topo"]["targets"], "conv_bn": gold["topo"]["conv_bn"],
            "relations": [tuple(r) for r in gold["topo"]["relations"]]}
    P.merge_batchnorm(st, topo)
    for k, d in gold["steps"]["merge_batchnorm"].items():
        assert sha(st[k]) == d, k
    iters, diffs, S = P.cross_layer_equalization(st, topo)
    assert iters == gold["cle_iters"] == 43
    np.testing.assert_allclose(diffs, gold["cle_diffs"], rtol=0, atol=0)
    for k, d in gold["steps"]["cross_layer_equalization"].items():
        assert sha(st[k]) == d, k
    for (a, b), s in S.items():
        assert s.tobytes() == npz[f"S/{a}->{b}"].tobytes()
    P.bias_absorption(st, topo, 3)
    for k in npz.files:
        if k.startswith("absorb/"):
            name = k[len("absorb/"):]
            assert rel_err(npz[k], st[name]) <= 1e-5, name
            st[name] = npz[k].copy()
    for bn in set(topo["conv_bn"].values()):
        st[bn + ".fake_weight"].fill(1.0)
        st[bn + ".fake_bias"].fill(0.0)
    codes = P.quantize_targ_layer(st, topo, 8, 8, want_codes=True)
    for k, d in gold["steps"]["quantize_targ_layer"].items():
        assert sha(st[k]) == d, k
    for k, d in gold["codes"].items():
        assert sha(codes[k].astype(np.uint8)) == d, k
    P.clip_weight(st, topo)
    for k, d in gold["steps"]["final"].items():
        assert sha(st[k]) == d, k
