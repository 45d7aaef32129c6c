"""GPU: the packed integer checkpoint (checkpoint.py) -- pack -> save -> load -> unpack restores the
quantised weights and biases of quantize_targ_layer bit for bit (against the reference's digests at W8 / W6 / W4),
and the file is 4x (W8) to 8x (W4) smaller than the fp32 tensors."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import ROOT
from helpers import fresh_model, sha, snapshot, switch
from test_gpu_sweep import absorbed, _restore  # noqa: F401  (module-scoped fixture + helper)

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("bits", [8, 6, 4])
def test_roundtrip_bit_exact(absorbed, tmp_path, bits):  # noqa: F811
    from data_free_quantization_b200 import checkpoint as C
    from data_free_quantization_b200.utils.layer_transform import merge_batchnorm
    from data_free_quantization_b200.utils.quantize import set_layer_bits
    with open(os.path.join(ROOT, "tests", "golden", "mobilenetv2_dfq" + ("" if bits == 8 else f"_w{bits}") + ".json")) as f:
        gold = json.load(f)
    model, graph, bottoms, targ = _restore(absorbed)
    set_layer_bits(graph, bits, 8, 8, targ)
    merge_batchnorm(model, graph, bottoms, targ)
    packed = C.pack_quantized(model, graph, targ, bits, 8)
    snap = snapshot(model, graph, targ)
    for k, d in gold["steps"]["quantize_targ_layer"].items():      # pack_quantized == quantize_targ_layer in place
        assert sha(snap[k]) == d, k
    path = str(tmp_path / "model.dfq")
    C.save_packed(packed, path)
    small, full = C.packed_nbytes(packed)
    assert full / small >= (7.5 if bits <= 4 else 3.9), (small, full)
    assert os.path.getsize(path) < 1.1 * small + 200_000
    # a fresh, un-quantised model of the same architecture receives the state
    model2, graph2, bottoms2, targ2 = _restore(absorbed)
    merge_batchnorm(model2, graph2, bottoms2, targ2)
    C.unpack_into(model2, C.load_packed(path))
    snap2 = snapshot(model2, graph2, targ2)
    for k, d in gold["steps"]["quantize_targ_layer"].items():
        if k.endswith(("weight", ".bias")) and "fake" not in k:
            assert sha(snap2[k]) == d, k


def test_rejects_foreign_file(tmp_path):
    from data_free_quantization_b200 import checkpoint as C
    p = str(tmp_path / "x.pt")
    torch.save({"format": "something else"}, p)
    with pytest.raises(ValueError):
        C.load_packed(p)


def test_true_integer_biases(gold_npz):
    """int32 biases in units of s_x * s_w.  Each is within half a unit of the reference-exact bias,
    and the static u8 plan run with them keeps every class of the reference-exact run."""
    import numpy as np
    from helpers import dfq_quantised
    from data_free_quantization_b200 import checkpoint as C
    from data_free_quantization_b200.eval_q8 import EvalPlanQ8
    model, graph, bottoms, targ, tr = dfq_quantised("mobilenetv2", "cuda:0", gold_npz)
    for m in model.modules():
        if hasattr(m, "set_update_stat"):
            m.set_update_stat(False)
    x = torch.randn(32, 3, 224, 224, generator=torch.Generator().manual_seed(7)).to("cuda:0")
    base = EvalPlanQ8(model, tr, batch=32)(x).clone()
    table = C.integer_biases(model, graph, targ)
    assert len(table) == 53
    mods = dict(model.named_modules())
    for name, e in table.items():
        b = mods[name].bias.data.double().cpu()
        assert e["bias_i32"].dtype == torch.int32
        assert float((e["bias_i32"].double() * e["scale"] - b).abs().max()) <= 0.5 * e["scale"] * (1 + 1e-6)
    C.apply_integer_biases(model, table)
    got = EvalPlanQ8(model, tr, batch=32)(x)
    assert torch.equal(got.argmax(1), base.argmax(1))
    assert float((got - base).abs().max()) <= 5e-2 * float(base.abs().max())
