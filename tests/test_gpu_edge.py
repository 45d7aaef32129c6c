"""GPU: the reference's edge cases on the kernels.
  * NaN / +-inf in the tensor handed to quantize(): torch.min / torch.max propagate NaN and Python's max(scale, 1e-8)
    keeps a NaN scale (utils/quantize.py:29-37, 66-76), so a single NaN poisons the whole tensor; an infinite range
    gives scale = inf and 0 * inf = NaN everywhere.
  * all-equal tensors at every bit-width: scale clamps to 1e-8 (:68).
  * a dead output channel in cross-layer equalisation: r1 = 0 -> s = (1/0) * sqrt(0) = NaN -> Python min/max leave
    s_max (Cross_layer_equal.py:38-43); replicate, do not fix."""
import numpy as np
import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def ref_quantize(x, bits, lo=None, hi=None):
    """utils/quantize.py:24-78 restated with torch CPU ops (asymmetric, Python-float range)."""
    t = torch.from_numpy(x.copy())
    lo = float(t.min()) if lo is None else lo
    hi = float(t.max()) if hi is None else hi
    qmax = 2. ** bits - 1.
    scale = max((hi - lo) / qmax, 1e-8)
    t.add_(-lo).div_(scale).clamp_(0., qmax).round_().mul_(scale).add_(lo)
    return t.numpy()


@pytest.mark.parametrize("poison", [float("nan"), float("inf"), -float("inf")])
@pytest.mark.parametrize("n", [257, 1 << 16])
def test_nan_and_inf_poison_the_range_like_torch(poison, n):
    from data_free_quantization_b200.utils.quantize import fake_quant_tensors, minmax, quantize
    rng = np.random.default_rng(1)
    x = rng.standard_normal(n).astype(np.float32)
    x[n // 3] = poison
    xd = torch.from_numpy(x).to(DEV)
    mm = minmax(xd).cpu().numpy()
    t = torch.from_numpy(x)
    want_mm = np.array([float(t.min()), float(t.max())], np.float32)
    assert np.array_equal(mm, want_mm, equal_nan=True), (mm, want_mm)
    # quantize_targ_layer's path: range reduced on the device, scale formed on the device
    out = torch.empty_like(xd)
    fake_quant_tensors([xd], 8, 0, outs=[out])
    want = ref_quantize(x, 8)
    got = out.cpu().numpy()
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.array_equal(got[~np.isnan(got)], want[~np.isnan(want)])
    # a given finite range with a poisoned element: only what torch's elementwise chain does to that element
    got = quantize(xd, 8, -2.0, 2.0).cpu().numpy()
    want = ref_quantize(x, 8, -2.0, 2.0)
    assert np.array_equal(got, want, equal_nan=True)


@pytest.mark.parametrize("bits", [2, 4, 6, 8, 16])
@pytest.mark.parametrize("value", [0.0, -3.25, 1e-12])
def test_all_equal_tensor(bits, value):
    from data_free_quantization_b200.utils.quantize import fake_quant_tensors
    x = np.full(5000, value, np.float32)
    xd = torch.from_numpy(x).to(DEV)
    out = torch.empty_like(xd)
    fake_quant_tensors([xd], bits, 0, outs=[out])
    assert np.array_equal(out.cpu().numpy(), ref_quantize(x, bits))


@pytest.mark.parametrize("impl", ["resident", "stream", "units"])
def test_cle_dead_channel_replicates_the_reference(oracle, impl, monkeypatch):
    from data_free_quantization_b200.Cross_layer_equal import cross_layer_equalization
    from data_free_quantization_b200.utils.relation import Relation
    monkeypatch.setenv("DFQ_CLE_IMPL", impl)
    rng = np.random.default_rng(3)
    W1 = (rng.standard_normal((48, 20)) * 0.4).astype(np.float32)
    W2 = (rng.standard_normal((24, 48, 1)) * 0.3).astype(np.float32)
    W1[5] = 0.0                     # dead output channel: r1 = 0
    W1[11] = 0.7                    # constant row: r1 = 0 as well
    W2[:, 17, :] = 0.0              # dead input column of the second layer: r2 = 0 -> s = 0 -> clamps to s_min
    b1 = rng.standard_normal(48).astype(np.float32)
    g, be = (rng.random(48) + 0.5).astype(np.float32), rng.standard_normal(48).astype(np.float32)
    l1, l2, bn = nn.Conv2d(1, 1, 1), nn.Conv2d(1, 1, 1), nn.BatchNorm2d(48)
    l1.weight = nn.Parameter(torch.from_numpy(W1.copy()).to(DEV), requires_grad=False)
    l1.bias = nn.Parameter(torch.from_numpy(b1.copy()).to(DEV), requires_grad=False)
    l2.weight = nn.Parameter(torch.from_numpy(W2.copy()).to(DEV), requires_grad=False)
    bn.register_buffer("fake_weight", torch.from_numpy(g.copy()).to(DEV))
    bn.register_buffer("fake_bias", torch.from_numpy(be.copy()).to(DEV))
    graph = {"a": l1, "b": l2, "n": bn}
    rels = [Relation("a", "b", "n")]
    info = cross_layer_equalization(graph, rels, [nn.Conv2d], Treshhold=-1.0, Count=10 ** 9, Save_state=False,
                                    max_iters=2, return_info=True)
    assert info["impl"] == impl and info["iters"] == 2
    S = np.ones(48, np.float32)
    with np.errstate(all="ignore"):
        for _ in range(2):
            S = (S * oracle.cle_pair(W1, W2, b1, g, be)).astype(np.float32)
    for got, want in ((l1.weight, W1), (l2.weight, W2), (l1.bias, b1), (bn.fake_weight, g), (bn.fake_bias, be),
                      (rels[0].S, S)):
        assert got.detach().cpu().numpy().tobytes() == want.tobytes()
    assert S[5] == np.float32(1e8) ** 2 or not np.isfinite(S[5]) or S[5] >= 1e8       # s_max kept for the dead channel


def test_straight_through_gradient_reaches_the_weights():
    """utils/quantize.py:81-85: UniformQuantize.backward passes the gradient straight through, so a QuantConv2d /
    QuantLinear whose weight requires grad must receive one (the weight goes through UniformQuantize, not a detach)."""
    from data_free_quantization_b200.utils.quantize import QuantLinear
    lin = QuantLinear(16, 8).to(DEV)
    lin.quant.eval()
    x = torch.randn(4, 16, device=DEV)
    y = lin(x)
    y.sum().backward()
    assert lin.weight.grad is not None and lin.weight.grad.shape == lin.weight.shape
    assert float(lin.weight.grad.abs().sum()) > 0


def test_same_tensor_twice_in_an_add_consumes_two_measures():
    """`x + x` records two operands; each occurrence goes through its own CustomTensorOP measure."""
    from data_free_quantization_b200 import graph as G
    from data_free_quantization_b200.utils import quantize as Q

    class M(nn.Module):
        def __init__(self):
            super().__init__()
            self.c = nn.Conv2d(3, 4, 1)

        def forward(self, x):
            y = self.c(x)
            return y + y

    m = M().to(DEV).eval()
    md = {1: [(nn.Conv2d, Q.QuantNConv2d), (nn.Linear, Q.QuantNLinear)]}
    m, tr = G.switch_layers(m, G.TorchTransformer(), torch.ones(1, 3, 8, 8, device=DEV), md,
                            ignore_layer=[Q.QuantMeasure], quant_op=True)
    top = m.custom_tensor_op
    qms = list(top.children())
    assert len(qms) == 2
    for q in qms:
        q.update_stat = True
        q.eval()
    G.replace_op()
    try:
        with torch.no_grad():
            m(torch.randn(2, 3, 8, 8, device=DEV))
    finally:
        G.restore_op()
    assert all(float(q.running_max) != 0.0 or float(q.running_min) != 0.0 for q in qms), "second measure never ran"
