"""GPU: the MobileNetV2 DFQ pipeline through the reference-shaped API against the golden digests taken from
the UNMODIFIED reference (tests/golden/), plus CLE / absorption on synthetic pairs against the C oracle."""
import numpy as np
import pytest
import torch
import torch.nn as nn

from helpers import fresh_model, names_of, rel_err, sha, snapshot, switch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def run():
    """merge -> relations -> CLE on the GPU; later tests continue from this state."""
    from data_free_quantization_b200.Cross_layer_equal import cross_layer_equalization
    from data_free_quantization_b200.utils.layer_transform import merge_batchnorm
    from data_free_quantization_b200.utils.relation import create_relation
    model, _ = fresh_model()
    model, tr, targ = switch(model, DEV)
    graph, bottoms = tr.log.getGraph(), tr.log.getBottoms()
    merge_batchnorm(model, graph, bottoms, targ)
    out = {"model": model, "graph": graph, "bottoms": bottoms, "targ": targ}
    out["merge"] = {k: sha(v) for k, v in snapshot(model, graph, targ).items()}
    rels = create_relation(graph, bottoms, targ)
    info = cross_layer_equalization(graph, rels, targ, Save_state=False, Treshhold=2e-7, return_info=True)
    out["rels"], out["info"] = rels, info
    out["cle"] = {k: sha(v) for k, v in snapshot(model, graph, targ).items()}
    return out


def test_merge_batchnorm_bit_exact(gold, run):
    for k, d in gold["steps"]["merge_batchnorm"].items():
        assert run["merge"][k] == d, k


def test_cle_bit_exact_same_iterations(gold, gold_npz, run):
    assert run["info"]["iters"] == gold["cle_iters"]
    # per-iteration metric: the reference forms it with fp32 torch.mean reductions (Cross_layer_equal.py:107), the
    # golden values are the oracle's double-accumulated restatement; the resident kernel sums a few dozen terms per
    # thread in fp32 and the rest in double.  What must be identical is the stopping iteration (asserted above).
    np.testing.assert_allclose(run["info"]["diffs"], gold["cle_diffs"], rtol=2e-6)
    for k, d in gold["steps"]["cross_layer_equalization"].items():
        assert run["cle"][k] == d, k
    names = names_of(run["model"])
    for r in run["rels"]:
        a, b, _ = r.get_idxs()
        assert r.S.cpu().numpy().tobytes() == gold_npz[f"S/{names[a]}->{names[b]}"].tobytes()


def test_absorb_quantize_clip_correct(gold, gold_npz, run):
    from data_free_quantization_b200.bias_absorption import bias_absorption
    from data_free_quantization_b200.bias_correction import bias_correction
    from data_free_quantization_b200.clip_weight import clip_weight
    from data_free_quantization_b200.utils.layer_transform import (merge_batchnorm, quantize_targ_layer,
                                                                   set_quant_minmax)
    from data_free_quantization_b200.utils.quantize import set_layer_bits
    model, graph, bottoms, targ = run["model"], run["graph"], run["bottoms"], run["targ"]
    bias_absorption(graph, run["rels"], bottoms, N=3)
    snap = snapshot(model, graph, targ)
    mods = dict(model.named_modules())
    for k in gold_npz.files:
        if k.startswith("absorb/"):
            name = k[len("absorb/"):]
            assert rel_err(gold_npz[k], snap[name]) <= 1e-5, name
            # carry the reference's values forward so the bit-exact checks below are not polluted by the
            # allowed summation-order difference of the absorption GEMV (same as oracle/gen_golden.py)
            mod, attr = name.rsplit(".", 1)
            getattr(mods[mod], attr).data.copy_(torch.from_numpy(gold_npz[k]))
    set_layer_bits(graph, 8, 8, 8, targ)
    merge_batchnorm(model, graph, bottoms, targ)              # main_dfq.py:211
    codes = quantize_targ_layer(graph, 8, 8, targ, want_codes=True)
    # want_codes quantised in place too
    snap = snapshot(model, graph, targ)
    for k, d in gold["steps"]["quantize_targ_layer"].items():
        assert sha(snap[k]) == d, k
    names = names_of(model)
    for key, (c, _rng) in codes.items():
        assert sha(c.cpu().numpy()) == gold["codes"][names[key]], names[key]
    set_quant_minmax(graph, bottoms, verbose=False)
    for key, m in graph.items():
        if type(m) in targ:
            lo, hi = gold["quant_minmax"][names[key]]
            assert float(m.quant.running_min) == lo and float(m.quant.running_max) == hi
    clip_weight(graph, [-15, 15], targ)
    before, after = bias_correction(graph, bottoms, tuple(targ), bits_weight=8)    # as wired: no-op
    assert [len(before), len(after)] == gold["bias_correction_as_wired"]
    snap = snapshot(model, graph, targ)
    for k, d in gold["steps"]["final"].items():
        assert sha(snap[k]) == d, k


def _pair(rng, C1, K1, O2, I2, KK):
    W1 = (rng.standard_normal((C1, K1)) * rng.uniform(0.05, 2.0, (C1, 1))).astype(np.float32)
    W2 = (rng.standard_normal((O2, I2, KK)) * 0.3).astype(np.float32)
    return W1, W2


@pytest.mark.parametrize("impl", ["resident", "stream", "units"])
@pytest.mark.parametrize("signed", [False, True])
def test_cle_pairs_against_oracle(oracle, signed, impl, monkeypatch):
    """Dense->dense, dense->depthwise, depthwise->dense, ragged channel counts; 3 iterations, bit-exact -- on the
    shared-memory-resident kernel (cle3.cu), the streaming kernel (cle2.cu) and the per-channel-unit kernel (cle.cu)."""
    from data_free_quantization_b200.Cross_layer_equal import cross_layer_equalization
    monkeypatch.setenv("DFQ_CLE_IMPL", impl)
    from data_free_quantization_b200.utils.relation import Relation
    rng = np.random.default_rng(4)
    shapes = [(32, 27, 32, 1, 9), (32, 9, 16, 32, 1), (16, 32, 96, 16, 1), (96, 16, 96, 1, 9), (96, 9, 24, 96, 1),
              (13, 7, 5, 13, 9), (8, 40, 1000, 8, 1)]
    graph, rels, ref = {}, [], []
    for i, (C1, K1, O2, I2, KK) in enumerate(shapes):
        W1, W2 = _pair(rng, C1, K1, O2, I2, KK)
        b1 = rng.standard_normal(C1).astype(np.float32)
        g, be = (rng.random(C1) + 0.5).astype(np.float32), rng.standard_normal(C1).astype(np.float32)
        l1, l2, bn = nn.Conv2d(1, 1, 1), nn.Conv2d(1, 1, 1), nn.BatchNorm2d(C1)
        l1.weight = nn.Parameter(torch.from_numpy(W1).to(DEV), requires_grad=False)
        l1.bias = nn.Parameter(torch.from_numpy(b1).to(DEV), requires_grad=False) if i % 2 == 0 else None
        l2.weight = nn.Parameter(torch.from_numpy(W2).to(DEV), requires_grad=False)
        bn.register_buffer("fake_weight", torch.from_numpy(g).to(DEV))
        bn.register_buffer("fake_bias", torch.from_numpy(be).to(DEV))
        graph[f"a{i}"], graph[f"b{i}"], graph[f"n{i}"] = l1, l2, bn
        rels.append(Relation(f"a{i}", f"b{i}", f"n{i}"))
        ref.append([W1.copy(), W2.copy(), b1.copy() if i % 2 == 0 else np.zeros(C1, np.float32), g.copy(), be.copy()])
    info = cross_layer_equalization(graph, rels, [nn.Conv2d], Treshhold=-1.0, Count=10 ** 9, signed=signed,
                                    Save_state=False, max_iters=3, return_info=True)
    assert info["iters"] == 3
    assert info["impl"] == impl
    for i, (W1, W2, b1, g, be) in enumerate(ref):
        S = np.ones(W1.shape[0], np.float32)
        for _ in range(3):
            S = (S * oracle.cle_pair(W1, W2, b1, g, be, signed=signed)).astype(np.float32)
        assert graph[f"a{i}"].weight.cpu().numpy().tobytes() == W1.tobytes(), i
        assert graph[f"b{i}"].weight.cpu().numpy().tobytes() == W2.tobytes(), i
        assert graph[f"a{i}"].bias.cpu().numpy().tobytes() == b1.tobytes(), i
        assert graph[f"n{i}"].fake_weight.cpu().numpy().tobytes() == g.tobytes(), i
        assert graph[f"n{i}"].fake_bias.cpu().numpy().tobytes() == be.tobytes(), i
        assert rels[i].S.cpu().numpy().tobytes() == S.tobytes(), i


@pytest.mark.parametrize("clip", [None, (-0.4, 0.4)])
def test_absorption_against_oracle(oracle, clip):
    from data_free_quantization_b200.bias_absorption import bias_absorption
    from data_free_quantization_b200.utils.relation import Relation
    rng = np.random.default_rng(8)
    shapes = [(32, 32, 1, 9), (32, 16, 32, 1), (96, 24, 96, 1), (960, 960, 1, 9), (64, 1000, 64, 1)]
    graph, bottoms, rels, ref = {}, {}, [], []
    for i, (C1, O2, I2, KK) in enumerate(shapes):
        W2 = (rng.standard_normal((O2, I2, KK)) * 0.3).astype(np.float32)
        b1, b2 = rng.standard_normal(C1).astype(np.float32), rng.standard_normal(O2).astype(np.float32)
        g, be = (rng.random(C1) * 0.3 + 0.05).astype(np.float32), rng.standard_normal(C1).astype(np.float32)
        l1, l2, bn, relu = nn.Conv2d(1, 1, 1), nn.Conv2d(1, 1, 1), nn.BatchNorm2d(C1), nn.ReLU()
        l1.weight = nn.Parameter(torch.zeros(C1, 1, device=DEV), requires_grad=False)
        l1.bias = nn.Parameter(torch.from_numpy(b1).to(DEV), requires_grad=False)
        l2.weight = nn.Parameter(torch.from_numpy(W2).to(DEV), requires_grad=False)
        l2.bias = nn.Parameter(torch.from_numpy(b2).to(DEV), requires_grad=False)
        bn.register_buffer("fake_weight", torch.from_numpy(g).to(DEV))
        bn.register_buffer("fake_bias", torch.from_numpy(be).to(DEV))
        for k, v, bot in ((f"a{i}", l1, ["Data"]), (f"n{i}", bn, [f"a{i}"]), (f"r{i}", relu, [f"n{i}"]),
                          (f"b{i}", l2, [f"r{i}"])):
            graph[k], bottoms[k] = v, bot
        rels.append(Relation(f"a{i}", f"b{i}", f"n{i}"))
        ref.append([W2, b1.copy(), b2.copy(), g, be.copy(), C1])
    got_c = bias_absorption(graph, rels, bottoms, N=3, clip=clip)
    for i, (W2, b1, b2, g, be, C1) in enumerate(ref):
        c = oracle.absorb_pair(W2, C1, b1, b2, g, be, 3.0)
        assert (c > 0).any()
        assert got_c[rels[i]].cpu().numpy().tobytes() == c.tobytes()
        assert graph[f"a{i}"].bias.cpu().numpy().tobytes() == b1.tobytes()
        assert graph[f"n{i}"].fake_bias.cpu().numpy().tobytes() == be.tobytes()
        assert rel_err(b2, graph[f"b{i}"].bias.cpu().numpy()) <= 1e-5
        want_w = np.clip(W2, *clip) if clip else W2
        assert graph[f"b{i}"].weight.cpu().numpy().tobytes() == want_w.tobytes()


@pytest.mark.parametrize("impl", ["resident", "stream"])
def test_cle_long_chain_many_ctas_against_oracle(oracle, impl, monkeypatch):
    """One chain of five tensors large enough that the resident kernel spreads it over a multi-CTA group (group
    barriers, column-range partials through L2) next to a tiny single-CTA chain; 4 iterations, bit-exact against
    the oracle's sequential loop (Cross_layer_equal.py:85-99 order)."""
    from data_free_quantization_b200 import Cross_layer_equal as CLE
    from data_free_quantization_b200.utils.relation import Relation
    monkeypatch.setenv("DFQ_CLE_IMPL", impl)
    rng = np.random.default_rng(11)
    # (O, I, KK, groups): pw 96->576, dw 576, pw 576->160, pw 160->960, linear-like 960->700; then a tiny pair
    specs = [[(576, 96, 1, 1), (576, 1, 9, 576), (160, 576, 1, 1), (960, 160, 1, 1), (700, 960, 1, 1)],
             [(8, 3, 9, 1), (8, 1, 9, 8)]]
    graph, rels, ref = {}, [], []
    for ci, chain in enumerate(specs):
        Ws = []
        for (O, I, KK, _g) in chain:
            Ws.append((rng.standard_normal((O, I, KK)) * rng.uniform(0.05, 1.5, (O, 1, 1))).astype(np.float32))
        for li, W in enumerate(Ws):
            conv = nn.Conv2d(1, 1, 1, bias=False)
            conv.weight = nn.Parameter(torch.from_numpy(W.copy()).to(DEV), requires_grad=False)
            graph[f"c{ci}l{li}"] = conv
            if li + 1 < len(Ws):
                bn = nn.BatchNorm2d(W.shape[0])
                bn.register_buffer("fake_weight", torch.from_numpy((rng.random(W.shape[0]) + 0.5).astype(np.float32)).to(DEV))
                bn.register_buffer("fake_bias", torch.from_numpy(rng.standard_normal(W.shape[0]).astype(np.float32)).to(DEV))
                graph[f"c{ci}n{li}"] = bn
                rels.append(Relation(f"c{ci}l{li}", f"c{ci}l{li + 1}", f"c{ci}n{li}"))
        ref.append(Ws)
    state = {k: (m.fake_weight.cpu().numpy().copy(), m.fake_bias.cpu().numpy().copy())
             for k, m in graph.items() if isinstance(m, nn.BatchNorm2d)}
    info = CLE.cross_layer_equalization(graph, rels, [nn.Conv2d], Treshhold=-1.0, Count=10 ** 9, Save_state=False,
                                        max_iters=4, return_info=True)
    assert info["iters"] == 4 and info["impl"] == impl
    if impl == "resident":
        assert max(n for _len, n in CLE.LAST_PLAN["chains"]) > 8, CLE.LAST_PLAN      # a real multi-CTA group
    diffs = []
    for _ in range(4):
        old = [[w.copy() for w in Ws] for Ws in ref]
        for ci, Ws in enumerate(ref):
            for li in range(len(Ws) - 1):
                W1 = Ws[li].reshape(Ws[li].shape[0], -1)
                b1 = np.zeros(W1.shape[0], np.float32)
                g, be = state[f"c{ci}n{li}"]
                oracle.cle_pair(W1, Ws[li + 1], b1, g, be)
        diffs.append(sum(float(np.float32(np.mean(np.abs(w.astype(np.float64) - o.astype(np.float64)))))
                         for Ws, Os in zip(ref, old) for w, o in zip(Ws, Os)))
    for ci, Ws in enumerate(ref):
        for li, W in enumerate(Ws):
            got = graph[f"c{ci}l{li}"].weight.cpu().numpy()
            assert got.tobytes() == W.tobytes(), (ci, li, int((got != W).sum()))
    for k, (g, be) in state.items():
        assert graph[k].fake_weight.cpu().numpy().tobytes() == g.tobytes(), k
        assert graph[k].fake_bias.cpu().numpy().tobytes() == be.tobytes(), k
    np.testing.assert_allclose(info["diffs"], diffs, rtol=1e-5)
