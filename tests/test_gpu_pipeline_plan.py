"""GPU: the captured form of the weight-side pipeline (pipeline_plan.py) -- the third and later passes of
run_dfq_steps over the same traced model replay two CUDA graphs; weights, biases, BN stashes, QuantMeasure ranges,
Relation.S and the module attributes must be what the eager functions produce (and what the reference produced:
tests/golden/mobilenetv2_dfq.json)."""
import contextlib
import io

import numpy as np
import pytest
import torch
import torch.nn as nn

from helpers import names_of, sha, snapshot

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


class Args:
    quantize = relu = equalize = absorption = clip_weight = correction = True
    visualize = False
    bits_weight = bits_activation = bits_bias = 8
    correction_mode = "ref"


def _state(case):
    snap = snapshot(case.model, case.graph, list(case.targ))
    out = {k: sha(v) for k, v in snap.items()}
    names = names_of(case.model)
    for key, m in case.graph.items():
        if type(m) in case.targ:
            out[names[key] + ".range"] = (float(m.quant.running_min), float(m.quant.running_max), m.quant.update_stat,
                                          m.quant.num_bits)
    for i, q in enumerate(case.model.custom_tensor_op.children()):
        out[f"top{i}.range"] = (float(q.running_min), float(q.running_max))
    return out


def test_replayed_pipeline_equals_eager_and_reference(gold, monkeypatch):
    import bench
    from data_free_quantization_b200 import pipeline_plan as PP
    from data_free_quantization_b200 import _lib as L
    PP.forget()
    case = bench.PipelineCase(torch.device(DEV))
    states, rels = [], []
    before = dict(PP.stats)
    with contextlib.redirect_stdout(io.StringIO()):
        for it in range(4):
            case.rearm(False)
            l0 = L.launch_count()
            r = case.step()
            torch.cuda.synchronize()
            launches = L.launch_count() - l0
            states.append(_state(case))
            rels.append([x.S.cpu().numpy().tobytes() for x in r])
            assert launches >= 8, launches
    assert PP.stats["eager"] - before["eager"] == 1 and PP.stats["captured"] - before["captured"] == 1
    assert PP.stats["replayed"] - before["replayed"] == 3
    for it in range(1, 4):
        assert states[it] == states[0], [k for k in states[0] if states[0][k] != states[it].get(k)][:5]
        assert rels[it] == rels[0]
    # and against the reference: the quantised weights (biases differ from the golden's only through the absorption
    # GEMV's summation order, which the golden test replaces by the reference's values; weights do not depend on it)
    for k, d in gold["steps"]["final"].items():
        if k.endswith(".weight"):
            assert states[0][k] == d, k
    for name, (lo, hi) in gold["quant_minmax"].items():
        got = states[3][name + ".range"]
        assert abs(got[0] - lo) <= 1e-6 * max(1.0, abs(lo)) and abs(got[1] - hi) <= 1e-6 * max(1.0, abs(hi)), name
        assert got[2] == 8 and got[3] == 8          # set_layer_bits bug-compat: update_stat = bits_activation
    # module-visible side effects of the eager functions are in place after a replay
    for m in case.model.modules():
        if isinstance(m, nn.BatchNorm2d):
            assert m.eps == 1e-12 and hasattr(m, "fake_weight") and hasattr(m, "fake_bias")
    assert all(m.bias is not None for m in case.targets)


def test_plan_not_used_when_the_model_is_not_fresh():
    """A second pass WITHOUT re-arming (biases exist, BN already identity) must not replay the recorded fresh pass."""
    import bench
    from data_free_quantization_b200 import pipeline_plan as PP
    PP.forget()
    case = bench.PipelineCase(torch.device(DEV))
    with contextlib.redirect_stdout(io.StringIO()):
        for _ in range(3):
            case.rearm(False)
            case.step()
        n = PP.stats["replayed"]
        assert n >= 1
        case.step()                                # stale state: eager
        assert PP.stats["replayed"] == n
        case.rearm(False)
        case.step()
        assert PP.stats["replayed"] == n + 1
    torch.cuda.synchronize()


def test_speculative_range_upload_is_checked_against_the_statistics():
    """MinmaxPlan.speculate() queues the previous pass's ranges before this pass's BatchNorm statistics have arrived;
    finish() must leave the ranges of THESE statistics: the memoised ones when nothing moved (even over a clobbered
    table), recomputed ones when a statistic changed."""
    import bench
    from data_free_quantization_b200 import pipeline_plan as PP
    PP.forget()
    case = bench.PipelineCase(torch.device(DEV))
    with contextlib.redirect_stdout(io.StringIO()):
        for _ in range(4):
            case.rearm(False)
            case.step()
    torch.cuda.synchronize()
    plan = next(v for v in PP._plans.values() if isinstance(v, PP.PipelinePlan))
    mm = plan.mm
    assert mm._memo is not None and mm._spec, "the replays should have armed the speculation"
    ranges = lambda: _state(case)
    want = {k: v for k, v in ranges().items() if k.endswith(".range")}
    for base, _host in mm._spec:                       # clobber the device tables: only the upload can restore them
        base.fill_(123.0)
    mm.start(); mm.speculate()
    assert mm._spec_done
    mm.finish()
    torch.cuda.synchronize()
    assert {k: v for k, v in ranges().items() if k.endswith(".range")} == want
    # a statistic moves: the speculation is wrong and must be overwritten by the recomputed ranges
    bn = mm.gather.bns[3]
    bn._buffers["fake_bias"].add_(0.75)
    mm.start(); mm.speculate(); mm.finish()
    torch.cuda.synchronize()
    moved = {k: v for k, v in ranges().items() if k.endswith(".range")}
    assert moved != want
    mm._memo = mm._spec = None                         # the same statistics without any shortcut
    mm.start(); mm.speculate(); mm.finish()
    torch.cuda.synchronize()
    assert {k: v for k, v in ranges().items() if k.endswith(".range")} == moved
    # and the shortcut is armed again for the new statistics
    mm.start(); mm.speculate()
    assert mm._spec_done
    mm.finish()
    torch.cuda.synchronize()
    assert {k: v for k, v in ranges().items() if k.endswith(".range")} == moved
