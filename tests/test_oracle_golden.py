"""CPU: the C oracle (oracle/dfq_oracle.c) reproduces the golden digests that oracle/gen_golden.py took from
the UNMODIFIED reference (tests/golden/mobilenetv2_dfq.{json,npz})."""
import numpy as np
import pytest

from helpers import fresh_model, np_state, rel_err, sha


@pytest.fixture(scope="module")
def oracle_run(gold, oracle):
    """Oracle pipeline merge -> CLE -> absorb on the seeded MobileNetV2 state (about 1.5 s)."""
    from oracle import pipeline as P
    _, sd = fresh_model()
    st = np_state(sd)
    topo = {"targets": gold["topo"]["targets"], "conv_bn": gold["topo"]["conv_bn"],
            "relations": [tuple(r) for r in gold["topo"]["relations"]]}
    out = {"topo": topo, "st": st}
    P.merge_batchnorm(st, topo)
    out["merge"] = {k: sha(v) for k, v in st.items() if hasattr(v, "tobytes")}
    iters, diffs, S = P.cross_layer_equalization(st, topo)
    out["cle"] = {k: sha(v) for k, v in st.items() if hasattr(v, "tobytes")}
    out["iters"], out["diffs"], out["S"] = iters, diffs, S
    return out


def test_init_state_is_the_golden_one(gold):
    _, sd = fresh_model()
    st = np_state(sd)
    assert sha(np.concatenate([v.ravel() for v in st.values()])) == gold["init_digest"]


def test_merge_batchnorm_bit_exact(gold, oracle_run):
    for k, d in gold["steps"]["merge_batchnorm"].items():
        assert oracle_run["merge"][k] == d, k


def test_cle_bit_exact_and_iteration_count(gold, gold_npz, oracle_run):
    assert oracle_run["iters"] == gold["cle_iters"] == 44
    np.testing.assert_allclose(oracle_run["diffs"], gold["cle_diffs"], rtol=0, atol=0)
    for k, d in gold["steps"]["cross_layer_equalization"].items():
        assert oracle_run["cle"][k] == d, k
    for (a, b), S in oracle_run["S"].items():
        assert S.tobytes() == gold_npz[f"S/{a}->{b}"].tobytes()


def test_absorb_quantize_clip(gold, gold_npz, oracle_run, oracle):
    from oracle import pipeline as P
    st = {k: (v.copy() if hasattr(v, "copy") else v) for k, v in oracle_run["st"].items()}
    topo = oracle_run["topo"]
    P.bias_absorption(st, topo, 3)
    for k in gold_npz.files:
        if k.startswith("absorb/"):
            name = k[len("absorb/"):]
            assert rel_err(gold_npz[k], st[name]) <= 1e-5, name
            st[name] = gold_npz[k].copy()          # carry the reference's biases forward (gen_golden.py)
    for bn in set(topo["conv_bn"].values()):       # main_dfq.py:211: second merge resets the BN stash
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


@pytest.mark.parametrize("bits", [2, 4, 6, 8, 16])
@pytest.mark.parametrize("symmetric", [False, True])
def test_fake_quant_matches_numpy_restatement(oracle, bits, symmetric):
    """utils/quantize.py:51-76 restated in numpy (fp32 ops, true division, clamp then round-half-even)."""
    rng = np.random.default_rng(bits * 2 + symmetric)
    x = (rng.standard_normal(20011) * 0.3).astype(np.float32)
    lo, hi = float(x.min()), float(x.max())
    if symmetric:
        qmin, qmax = -2.0 ** (bits - 1), 2 ** (bits - 1) - 1
        scale, mn = max(abs(hi), abs(lo)) / qmax, 0.0
    else:
        qmin, qmax = 0.0, 2.0 ** bits - 1.0
        scale, mn = (hi - lo) / (qmax - qmin), lo
    scale = max(scale, 1e-8)
    t = (x + np.float32(-mn)) / np.float32(scale)
    q = np.rint(np.clip(t, np.float32(qmin), np.float32(qmax)))
    want = q * np.float32(scale) + np.float32(mn)
    got, codes = oracle.fake_quant(x, bits, lo, hi, symmetric, want_codes=True)
    assert got.tobytes() == want.astype(np.float32).tobytes()
    assert np.array_equal(codes, (q - np.float32(qmin)).astype(np.int32))


def test_edge_cases(oracle):
    # constant tensor -> scale clamps to 1e-8; single element; empty
    x = np.full(17, 0.25, dtype=np.float32)
    y = oracle.fake_quant(x, 8, 0.25, 0.25)
    assert np.all(np.abs(y - 0.25) <= 1e-7)
    assert oracle.fake_quant(np.zeros(0, np.float32), 8, 0.0, 1.0).size == 0
    one = oracle.fake_quant(np.array([3.0], np.float32), 4)
    assert one.shape == (1,)


@pytest.mark.parametrize("bits", [4, 6])
def test_bit_width_sweep_per_tensor(gold_npz, oracle_run, oracle, bits):
    """BASELINE.json configs[2]: the W4 / W6 quantised state of the unmodified reference
    (oracle/gen_golden.py --bits b --tag _wb) from the same post-absorption state as the W8 golden."""
    import json
    import os
    from conftest import ROOT
    from oracle import pipeline as P
    with open(os.path.join(ROOT, "tests", "golden", f"mobilenetv2_dfq_w{bits}.json")) as f:
        sweep = json.load(f)
    assert sweep["bits"] == bits
    st = {k: (v.copy() if hasattr(v, "copy") else v) for k, v in oracle_run["st"].items()}
    topo = oracle_run["topo"]
    P.bias_absorption(st, topo, 3)
    for k in gold_npz.files:
        if k.startswith("absorb/"):
            st[k[len("absorb/"):]] = gold_npz[k].copy()
    for bn in set(topo["conv_bn"].values()):
        st[bn + ".fake_weight"].fill(1.0)
        st[bn + ".fake_bias"].fill(0.0)
    codes = P.quantize_targ_layer(st, topo, bits, 8, want_codes=True)
    for k, d in sweep["steps"]["quantize_targ_layer"].items():
        assert sha(st[k]) == d, k
    for k, d in sweep["codes"].items():
        assert int(codes[k].max()) <= 2 ** bits - 1
        assert sha(codes[k].astype(np.uint8)) == d, k
    P.clip_weight(st, topo)
    for k, d in sweep["steps"]["final"].items():
        assert sha(st[k]) == d, k
