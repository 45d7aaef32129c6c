"""GPU: eval parity at the tolerance north_star states, layer by layer ("teacher-forced").

tests/golden/eval_tf_<model>.npz holds, for ONE forward of the UNMODIFIED reference (oracle/gen_eval_golden.py:
utils/quantize.py:106-123, 250-256, 358-363 with replace_op() in force, main_dfq.py:96-103), for EVERY target layer
(53 / 61) and every patched tensor op (11 / 15): the uint8 codes its QuantMeasure produced for the input, the
(min, max) it quantised with, and a strided sample of its output.  Each kernel of the static eval plan is fed the
reference's input and range and must reproduce the reference's output within 1e-3 of the layer's largest output
(north_star's bar; measured ~1e-6: the codes are identical, only the summation order of the convolution differs).

Why layer by layer: the end-to-end logits of this random-init network are chaotic under 8-bit activations -- the
reference's own logits move by 1.6e-2 of the largest logit (MobileNetV2; 3.5e-2 DeepLab) when its input is scaled by
(1 + 1e-7) (tests/golden/eval_top1_*.json, measured with the reference itself) -- so no implementation with another
summation order can meet 1e-3 at the logits, the reference included.  End-to-end agreement is asserted on the
classes instead (test_gpu_eval_top1.py) next to that measured noise floor."""
import json
import os

import numpy as np
import pytest
import torch

from helpers import dfq_quantised

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 1e-3            # north_star: eval within 1e-3 relative
EXPECT = 5e-5         # what identical codes + fp32 accumulation actually give (asserted on the median layer)


def _dequant(codes, lo, hi):
    """code * scale + min with the two fp32 roundings of utils/quantize.py:76 (scale formed in double, :66-68)."""
    scale = np.float32(max((float(hi) - float(lo)) / 255.0, 1e-8))
    return (codes.astype(np.float32) * scale).astype(np.float32) + np.float32(lo)


@pytest.mark.parametrize("which", ["mobilenetv2", "deeplab"])
def test_every_layer_and_tensor_op_matches_the_reference_given_its_input(which):
    from data_free_quantization_b200.eval_plan import EvalPlan
    tf = np.load(os.path.join(GOLD, f"eval_tf_{which}.npz"))
    meta = json.load(open(os.path.join(GOLD, f"eval_tf_{which}.json")))
    stride = meta["stride"]
    gold_npz = np.load(os.path.join(GOLD, f"{which}_dfq.npz"))
    model, graph, bottoms, targ, tr = dfq_quantised(which, DEV, gold_npz)
    for m in model.modules():                       # ranges are forced: no measure kernel in this test
        if hasattr(m, "set_update_stat"):
            m.set_update_stat(False)
    in_shape = tuple(meta["layers"][0]["in_shape"][1:])
    plan = EvalPlan(model, tr, batch=1, in_shape=in_shape)
    errs = {}
    # ---- the 53 / 61 target layers
    for rec in meta["layers"]:
        name = rec["name"]
        lo, hi = (float(v) for v in tf[f"L/{name}/range"])
        x = torch.from_numpy(_dequant(tf[f"L/{name}/xq"], lo, hi)).to(DEV)
        io = plan.step_io[name]
        assert tuple(io["srcs"][0].shape) == tuple(rec["in_shape"]), name
        got = plan.teacher_force(name, [x], [(lo, hi)]).cpu().numpy().ravel()[::stride]
        want = tf[f"L/{name}/y"]
        if io["relu"]:
            want = np.maximum(want, 0.0)            # the plan fuses the ReLU that follows the (identity) BatchNorm
        assert got.shape == want.shape, name
        errs[name] = float(np.abs(got - want).max() / max(np.abs(want).max(), 1e-30))
    # ---- the patched tensor ops, in execution order
    tops = [n for n, io in plan.step_io.items() if io["kind"] in ("add", "mean", "interp", "cat")]
    assert len(tops) == len(meta["tensor_ops"])
    kind_of = {"add": "add", "mean": "mean", "interp": "interp", "cat": "cat"}
    for i, (name, rec) in enumerate(zip(tops, meta["tensor_ops"])):
        io = plan.step_io[name]
        assert kind_of[io["kind"]] == rec["kind"] and len(io["srcs"]) == rec["operands"], (name, rec)
        xs, rs = [], []
        for j in range(rec["operands"]):
            lo, hi = (float(v) for v in tf[f"T/{i}/range{j}"])
            xs.append(torch.from_numpy(_dequant(tf[f"T/{i}/xq{j}"], lo, hi)).to(DEV))
            rs.append((lo, hi))
        got = plan.teacher_force(name, xs, rs).cpu().numpy().ravel()[::stride]
        want = tf[f"T/{i}/y"]
        assert got.shape == want.shape, name
        errs[f"{rec['kind']}[{i}]"] = float(np.abs(got - want).max() / max(np.abs(want).max(), 1e-30))
    worst = max(errs, key=errs.get)
    print(f"{which}: {len(errs)} ops teacher-forced, worst {worst} {errs[worst]:.2e}, "
          f"median {np.median(list(errs.values())):.2e}")
    bad = {k: v for k, v in errs.items() if v > TOL}
    assert not bad, bad
    assert np.median(list(errs.values())) <= EXPECT
    # every layer ran on a libdfq_b200 kernel of the plan (no cuDNN / ATen arithmetic on this path)
    assert len(errs) == len(meta["layers"]) + len(meta["tensor_ops"])
