"""GPU: bias_correction in its two active modes against fixtures generated from the UNMODIFIED reference
(oracle/gen_biascorr_golden.py): 'ref-body' (reference function body on a position-keyed graph) and 'paper'
(b -= eps . E[x] assembled from the reference's helpers).  Tolerance: 1e-5 relative (north_star)."""
import os

import numpy as np
import pytest
import torch

from conftest import ROOT
from helpers import fresh_model, rel_err, snapshot, switch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def fx():
    return np.load(os.path.join(ROOT, "tests", "golden", "biascorr_mobilenetv2.npz"))


def _merged():
    from data_free_quantization_b200.utils.layer_transform import merge_batchnorm
    model, _ = fresh_model()
    model, tr, targ = switch(model, DEV)
    graph, bottoms = tr.log.getGraph(), tr.log.getBottoms()
    merge_batchnorm(model, graph, bottoms, targ)
    return model, graph, bottoms, targ


def test_paper_mode(fx):
    from data_free_quantization_b200.bias_correction import bias_correction
    model, graph, bottoms, targ = _merged()
    before, after = bias_correction(graph, bottoms, tuple(targ), bits_weight=8, mode="paper")
    assert len(after) == 52                      # every target layer except the first conv (bottom = Data)
    snap = snapshot(model, graph, targ)
    n = 0
    for k in fx.files:
        if k.startswith("paper/"):
            assert rel_err(fx[k], snap[k[len("paper/"):]]) <= 1e-5, k
            n += 1
    assert n == 53


def test_ref_body_mode(fx):
    from data_free_quantization_b200.bias_correction import bias_correction
    model, graph, bottoms, targ = _merged()
    keys = list(graph.keys())
    pos = {k: i for i, k in enumerate(keys)}
    re = lambda k: k if isinstance(graph[k], str) else pos[k]
    graph_pos = type(graph)((re(k), v) for k, v in graph.items())
    bottoms_pos = {re(k): (None if v is None else [re(b) for b in v]) for k, v in bottoms.items()}
    before, after = bias_correction(graph_pos, bottoms_pos, tuple(targ), bits_weight=8)
    assert len(after) == 52
    snap = snapshot(model, graph, targ)
    for k in fx.files:
        if k.startswith("refbody/"):
            assert rel_err(fx[k], snap[k[len("refbody/"):]]) <= 1e-5, k


def test_as_wired_is_noop():
    from data_free_quantization_b200.bias_correction import bias_correction
    model, graph, bottoms, targ = _merged()
    s0 = {k: v.copy() for k, v in snapshot(model, graph, targ).items()}
    assert bias_correction(graph, bottoms, tuple(targ), bits_weight=8, visualize=False) == ({}, {})
    s1 = snapshot(model, graph, targ)
    assert all(np.array_equal(s0[k], s1[k]) for k in s0)
