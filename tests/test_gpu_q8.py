"""GPU: the static-range u8 kernels (csrc/q8.cu) one by one against restatements built from the oracle's quantiser, then
the whole static plan (eval_q8.EvalPlanQ8) against the UNMODIFIED reference run with set_update_stat(False) on every
QuantMeasure (utils/quantize.py:125-126; goldens from oracle/gen_eval_golden.py --what tf_static,static).

Codes are integers: the bar is IDENTICAL codes.  A producer's fp32 value differs from the reference's by the summation
order of its convolution (~1e-7 relative), which moves a value across a rounding boundary of the consumer's 8-bit
grid for about one element in 10^4; such elements may differ by ONE code, never more."""
import ctypes
import json
import os

import numpy as np
import pytest
import torch

from helpers import dfq_quantised

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def slot(lo, hi):
    from data_free_quantization_b200 import _lib as L
    p = L.DfqQuantParams()
    lo, hi = float(np.float32(lo)), float(np.float32(hi))        # ranges live in fp32 buffers (QuantMeasure.running_*)
    L.lib().dfq_quant_params_host(lo, hi, 8, 0, ctypes.byref(p))
    return torch.tensor([float(np.float32(lo)), float(np.float32(hi)), p.scale, 0.0], dtype=torch.float32, device=DEV), p.scale


def deq(codes, lo, scale):
    return (codes.astype(np.float32) * np.float32(scale)).astype(np.float32) + np.float32(lo)


def check_codes(got, want, what, max_frac=2e-3):
    got, want = np.asarray(got, np.int32), np.asarray(want, np.int32)
    if got.size == want.size and got.shape[:2] == want.shape[:2]:
        want = want.reshape(got.shape)              # (the model flattens H x W before torch.mean)
    assert got.shape == want.shape, (what, got.shape, want.shape)
    d = np.abs(got - want)
    assert d.max() <= 1, (what, int(d.max()))
    assert d.mean() <= max_frac, (what, float(d.mean()))
    return float(d.mean())


@pytest.mark.parametrize("G,K,O,o_cap,relu,dual", [
    (1000, 16, 96, 0, 1, False), (777, 24, 144, 0, 0, True), (4096, 96, 24, 0, 0, True), (300, 144, 32, 0, 1, False),
    (12544, 320, 1280, 0, 1, False), (256, 1280, 1000, 32, 0, False), (129, 960, 160, 0, 0, True),
    (5000, 32, 16, 0, 0, False), (640, 576, 250, 64, 1, True)])
def test_pointwise_q8_against_integer_restatement(oracle, G, K, O, o_cap, relu, dual):
    from data_free_quantization_b200 import _lib as L
    from data_free_quantization_b200.eval_q8 import PackedQ8, to_codes_nhwc, from_codes_nhwc
    rng = np.random.default_rng(G + K + O)
    xc = rng.integers(0, 256, (G, K)).astype(np.uint8)
    lo_x, hi_x = -1.7, 2.9
    in_slot, sx = slot(lo_x, hi_x)
    w0 = (rng.standard_normal((O, K)) * 0.2).astype(np.float32)
    w, wq = oracle.fake_quant(w0, 8, float(w0.min()), float(w0.max()), want_codes=True)
    sw, mw = float(oracle.quant_params(float(w0.min()), float(w0.max()), 8)[1]), float(w0.min())
    bias = (rng.standard_normal(O) * 0.5).astype(np.float32)
    A, B = xc.astype(np.int64), wq.astype(np.int64)
    exact = sx * sw * (A @ B.T) + sx * mw * A.sum(1, keepdims=True) + lo_x * sw * B.sum(1)[None, :] + K * lo_x * mw
    exact = exact + bias[None, :].astype(np.float64)
    if relu:
        exact = np.maximum(exact, 0)
    pk = PackedQ8(torch.from_numpy(w).to(DEV), o_cap=o_cap)
    x = to_codes_nhwc(torch.from_numpy(xc)).reshape(-1).to(DEV)
    Cp = (O + 15) // 16 * 16
    y0 = torch.full((G * Cp,), 77, dtype=torch.int8, device=DEV)
    y1 = torch.full((G * Cp,), 77, dtype=torch.int8, device=DEV)
    yf = torch.full((G, O), float("nan"), device=DEV)
    ranges = [(float(exact.min()) - 0.1, float(exact.max()) * 0.8), (-0.5, float(exact.max()) + 1.0)]
    s0, sc0 = slot(*ranges[0])
    s1, sc1 = slot(*ranges[1])
    bd = torch.from_numpy(bias).to(DEV)
    d = L.DfqPwQ8Desc()
    d.x, d.wpk, d.wr, d.bias, d.in_slot = x.data_ptr(), pk.wpk.data_ptr(), pk.wr.data_ptr(), bd.data_ptr(), in_slot.data_ptr()
    d.y0, d.out_slot0 = y0.data_ptr(), s0.data_ptr()
    if dual:
        d.y1, d.out_slot1 = y1.data_ptr(), s1.data_ptr()
    d.yf, d.yf_g, d.yf_o = yf.data_ptr(), O, 1
    d.G, d.K, d.O, d.relu, d.o_cap, d.s_w, d.c_w = G, K, O, relu, o_cap, pk.s_w, pk.c_w
    L.check(L.lib().dfq_pw_q8(ctypes.byref(d), L.stream()), "dfq_pw_q8")
    torch.cuda.synchronize()
    got = yf.cpu().numpy()
    err = np.abs(got - exact).max() / np.abs(exact).max()
    assert err <= 3e-6, err
    for t, (yb, (lo, hi)) in enumerate(zip((y0, y1), ranges)):
        if t == 1 and not dual:
            break
        codes = from_codes_nhwc(yb.cpu(), (G, O)).numpy()
        _q, want = oracle.fake_quant(got.astype(np.float32), 8, float(np.float32(lo)), float(np.float32(hi)), want_codes=True)
        assert np.array_equal(codes, want.astype(np.uint8)), int((codes != want).sum())     # exact quantiser of OUR fp32
        raw = yb.cpu().view(G, Cp)[:, O:]
        assert not raw.any(), "pad channels must be s8 zero"
    if not dual:
        assert bool((y1 == 77).all())


@pytest.mark.parametrize("N,C,H,W,stride,relu,dual", [(2, 32, 112, 112, 1, 1, False), (3, 96, 56, 56, 2, 1, True),
                                                     (2, 144, 29, 31, 1, 0, False), (4, 960, 7, 7, 1, 1, True),
                                                     (1, 24, 14, 14, 2, 0, False), (2, 576, 14, 14, 2, 1, False)])
def test_depthwise_q8(oracle, N, C, H, W, stride, relu, dual):
    from data_free_quantization_b200 import _lib as L
    from data_free_quantization_b200.eval_q8 import to_codes_nhwc, from_codes_nhwc, pack_depthwise
    rng = np.random.default_rng(C + H)
    xc = rng.integers(0, 256, (N, C, H, W)).astype(np.uint8)
    lo_x, hi_x = -0.3, 5.2
    in_slot, sx = slot(lo_x, hi_x)
    xq = deq(xc, np.float32(lo_x), sx)
    w0 = (rng.standard_normal((C, 1, 3, 3)) * 0.3).astype(np.float32)
    w = oracle.fake_quant(w0, 8, float(w0.min()), float(w0.max()))          # what quantize_targ_layer leaves
    bias = (rng.standard_normal(C) * 0.2).astype(np.float32)
    ref = torch.nn.functional.conv2d(torch.from_numpy(xq).double(), torch.from_numpy(w).double(),
                                     torch.from_numpy(bias).double(), stride, 1, 1, C).numpy()
    if relu:
        ref = np.maximum(ref, 0)
    Cp = (C + 15) // 16 * 16
    wrow, dsw, dcw = pack_depthwise(torch.from_numpy(w).to(DEV))
    bp = torch.zeros(Cp, device=DEV); bp[:C] = torch.from_numpy(bias).to(DEV)
    OH, OW = ref.shape[2], ref.shape[3]
    x = to_codes_nhwc(torch.from_numpy(xc)).reshape(-1).to(DEV)
    y0 = torch.full((N * OH * OW * Cp,), 55, dtype=torch.int8, device=DEV)
    y1 = torch.full((N * OH * OW * Cp,), 55, dtype=torch.int8, device=DEV)
    ranges = [(float(ref.min()) * 0.9, float(ref.max()) * 0.9), (0.0, float(ref.max()) * 1.1 + 0.1)]
    s0, _ = slot(*ranges[0]); s1, _ = slot(*ranges[1])
    d = L.DfqDwQ8Desc()
    d.x, d.wrow, d.bias, d.in_slot, d.y0, d.out_slot0 = x.data_ptr(), wrow.data_ptr(), bp.data_ptr(), in_slot.data_ptr(), y0.data_ptr(), s0.data_ptr()
    d.s_w, d.c_w = dsw, dcw
    if dual:
        d.y1, d.out_slot1 = y1.data_ptr(), s1.data_ptr()
    d.N, d.H, d.W, d.C, d.stride, d.relu = N, H, W, C, stride, relu
    d.pad, d.dil = 1, 1
    L.check(L.lib().dfq_dw_q8(ctypes.byref(d), L.stream()), "dfq_dw_q8")
    torch.cuda.synchronize()
    for t, (yb, (lo, hi)) in enumerate(zip((y0, y1), ranges)):
        if t == 1 and not dual:
            break
        codes = from_codes_nhwc(yb.cpu(), ref.shape).numpy()
        _q, want = oracle.fake_quant(ref.astype(np.float32), 8, float(np.float32(lo)), float(np.float32(hi)), want_codes=True)
        check_codes(codes, want, f"dw out {t}")
        assert not yb.cpu().view(-1, Cp)[:, C:].any()


def test_stem_add_mean_q8(oracle):
    from data_free_quantization_b200 import _lib as L
    from data_free_quantization_b200.eval_q8 import to_codes_nhwc, from_codes_nhwc
    lib = L.lib()
    rng = np.random.default_rng(9)
    # ---- stem: fp32 image, quantised on load
    N, H, W, O = 3, 64, 48, 32
    img = (rng.standard_normal((N, 3, H, W)) * 1.2).astype(np.float32)
    lo_i, hi_i = -2.11790393, 2.64
    in_slot, _ = slot(lo_i, hi_i)
    xq = oracle.fake_quant(img, 8, float(np.float32(lo_i)), float(np.float32(hi_i)))
    w = (rng.standard_normal((O, 3, 3, 3)) * 0.3).astype(np.float32)
    bias = (rng.standard_normal(O) * 0.2).astype(np.float32)
    ref = np.maximum(torch.nn.functional.conv2d(torch.from_numpy(xq).double(), torch.from_numpy(w).double(),
                                                torch.from_numpy(bias).double(), 2, 1).numpy(), 0)
    OH, OW = ref.shape[2:]
    y0 = torch.zeros(N * OH * OW * 32, dtype=torch.int8, device=DEV)
    lo, hi = 0.0, float(ref.max()) * 0.95
    s0, _ = slot(lo, hi)
    xd, wd, bd = torch.from_numpy(img).to(DEV), torch.from_numpy(w).to(DEV), torch.from_numpy(bias).to(DEV)
    d = L.DfqStemQ8Desc()
    d.x, d.w, d.bias, d.in_slot, d.y0, d.out_slot0 = xd.data_ptr(), wd.data_ptr(), bd.data_ptr(), in_slot.data_ptr(), y0.data_ptr(), s0.data_ptr()
    d.N, d.H, d.W, d.O, d.stride, d.pad, d.relu = N, H, W, O, 2, 1, 1
    L.check(lib.dfq_stem_q8(ctypes.byref(d), L.stream()), "dfq_stem_q8")
    _q, want = oracle.fake_quant(ref.astype(np.float32), 8, float(np.float32(lo)), float(np.float32(hi)), want_codes=True)
    check_codes(from_codes_nhwc(y0.cpu(), ref.shape).numpy(), want, "stem")
    # ---- residual add of two grids into two grids
    N, C, H, W = 2, 24, 20, 12
    ca, cb = rng.integers(0, 256, (N, C, H, W)).astype(np.uint8), rng.integers(0, 256, (N, C, H, W)).astype(np.uint8)
    (sa, sca), (sb, scb) = slot(-3.0, 2.0), slot(-0.7, 4.4)
    s = deq(ca, np.float32(-3.0), sca) + deq(cb, np.float32(-0.7), scb)
    outs = [(-4.0, 5.0), (-2.5, 6.6)]
    (t0, _), (t1, _) = slot(*outs[0]), slot(*outs[1])
    xa, xb = to_codes_nhwc(torch.from_numpy(ca)).reshape(-1).to(DEV), to_codes_nhwc(torch.from_numpy(cb)).reshape(-1).to(DEV)
    y0, y1 = torch.zeros_like(xa), torch.zeros_like(xa)
    L.check(lib.dfq_add_q8(xa.data_ptr(), xb.data_ptr(), sa.data_ptr(), sb.data_ptr(), y0.data_ptr(), t0.data_ptr(),
                           y1.data_ptr(), t1.data_ptr(), N * H * W, C, None, None, L.stream()), "dfq_add_q8")
    for yb, (lo, hi) in zip((y0, y1), outs):
        _q, want = oracle.fake_quant(s.astype(np.float32), 8, float(np.float32(lo)), float(np.float32(hi)), want_codes=True)
        assert np.array_equal(from_codes_nhwc(yb.cpu(), s.shape).numpy(), want.astype(np.uint8))   # elementwise: exact
        assert not yb.cpu().view(-1, 32)[:, C:].any()
    # ---- spatial mean feeding another grid
    N, C, H, W = 5, 1280, 7, 7
    cx = rng.integers(0, 256, (N, C, H, W)).astype(np.uint8)
    sin, scin = slot(0.0, 6.0)
    m = deq(cx, np.float32(0.0), scin).reshape(N, C, -1).astype(np.float64).mean(-1)
    so, _ = slot(0.1, 5.0)
    x = to_codes_nhwc(torch.from_numpy(cx)).reshape(-1).to(DEV)
    y = torch.full((N * C,), 9, dtype=torch.int8, device=DEV)
    yf = torch.zeros(N, C, device=DEV)
    L.check(lib.dfq_mean_q8(x.data_ptr(), sin.data_ptr(), y.data_ptr(), so.data_ptr(), yf.data_ptr(), N, C, H * W,
                            L.stream()), "dfq_mean_q8")
    assert np.abs(yf.cpu().numpy() - m).max() <= 2e-6 * np.abs(m).max()
    _q, want = oracle.fake_quant(yf.cpu().numpy(), 8, float(np.float32(0.1)), float(np.float32(5.0)), want_codes=True)
    assert np.array_equal(from_codes_nhwc(y.cpu(), (N, C)).numpy(), want.astype(np.uint8))


def test_q8_kernels_stay_inside_their_buffers(oracle):
    """compute-sanitizer is closed on this GPU pool (profiles/r2_sanitizer.md): every output buffer of the u8 kernels
    sits between two sentinel regions that must come back untouched; shapes are not multiples of the tile sizes."""
    from data_free_quantization_b200 import _lib as L
    from data_free_quantization_b200.eval_q8 import PackedQ8, pack_depthwise, to_codes_nhwc
    lib = L.lib()
    rng = np.random.default_rng(21)
    GUARD = 4096

    def guarded(nbytes, dtype=torch.int8):
        item = torch.empty(0, dtype=dtype).element_size()
        buf = torch.full((2 * GUARD + nbytes * item,), 0x5a, dtype=torch.uint8, device=DEV)
        return buf, buf[GUARD:GUARD + nbytes * item].view(dtype)

    def intact(buf, nbytes):
        b = buf.cpu().numpy()
        return bool((b[:GUARD] == 0x5a).all() and (b[GUARD + nbytes:] == 0x5a).all())
    in_slot, _ = slot(-1.0, 3.0)
    s0, _ = slot(-2.0, 9.0)
    s1, _ = slot(0.0, 7.0)
    # pointwise: ragged last tile, two passes, two consumers, fp32 output as well
    for G, K, O, cap in ((1, 16, 16, 0), (130, 40, 250, 64), (513, 1280, 33, 0)):
        w0 = (rng.standard_normal((O, K)) * 0.2).astype(np.float32)
        w = oracle.fake_quant(w0, 8, float(w0.min()), float(w0.max()))
        pk = PackedQ8(torch.from_numpy(w).to(DEV), o_cap=cap)
        x = to_codes_nhwc(torch.from_numpy(rng.integers(0, 256, (G, K)).astype(np.uint8))).reshape(-1).to(DEV)
        Cp = (O + 15) // 16 * 16
        b0, y0 = guarded(G * Cp)
        b1, y1 = guarded(G * Cp)
        bf, yf = guarded(G * O, torch.float32)
        d = L.DfqPwQ8Desc()
        d.x, d.wpk, d.wr, d.in_slot = x.data_ptr(), pk.wpk.data_ptr(), pk.wr.data_ptr(), in_slot.data_ptr()
        d.y0, d.out_slot0, d.y1, d.out_slot1 = y0.data_ptr(), s0.data_ptr(), y1.data_ptr(), s1.data_ptr()
        d.yf, d.yf_g, d.yf_o = yf.data_ptr(), O, 1
        d.G, d.K, d.O, d.relu, d.o_cap, d.s_w, d.c_w = G, K, O, 1, cap, pk.s_w, pk.c_w
        L.check(lib.dfq_pw_q8(ctypes.byref(d), L.stream()), "dfq_pw_q8")
        torch.cuda.synchronize()
        assert intact(b0, G * Cp) and intact(b1, G * Cp) and intact(bf, 4 * G * O), (G, K, O)
        assert not bool(torch.isnan(yf).any())
    # depthwise, both strides, odd sizes, two consumers
    for N, C, H, W, stride in ((1, 20, 5, 3, 1), (2, 33, 9, 11, 2), (1, 16, 1, 1, 1)):
        w0 = (rng.standard_normal((C, 1, 3, 3)) * 0.3).astype(np.float32)
        wrow, dsw, dcw = pack_depthwise(torch.from_numpy(oracle.fake_quant(w0, 8, float(w0.min()), float(w0.max()))).to(DEV))
        Cp = (C + 15) // 16 * 16
        OH, OW = (H - 1) // stride + 1, (W - 1) // stride + 1
        x = to_codes_nhwc(torch.from_numpy(rng.integers(0, 256, (N, C, H, W)).astype(np.uint8))).reshape(-1).to(DEV)
        b0, y0 = guarded(N * OH * OW * Cp)
        b1, y1 = guarded(N * OH * OW * Cp)
        bp = torch.zeros(Cp, device=DEV)
        d = L.DfqDwQ8Desc()
        d.x, d.wrow, d.bias, d.in_slot = x.data_ptr(), wrow.data_ptr(), bp.data_ptr(), in_slot.data_ptr()
        d.y0, d.out_slot0, d.y1, d.out_slot1 = y0.data_ptr(), s0.data_ptr(), y1.data_ptr(), s1.data_ptr()
        d.N, d.H, d.W, d.C, d.stride, d.relu, d.s_w, d.c_w = N, H, W, C, stride, 1, dsw, dcw
        d.pad, d.dil = 1, 1
        L.check(lib.dfq_dw_q8(ctypes.byref(d), L.stream()), "dfq_dw_q8")
        torch.cuda.synchronize()
        assert intact(b0, N * OH * OW * Cp) and intact(b1, N * OH * OW * Cp), (N, C, H, W, stride)
    # stem, add, mean
    N, H, W, O = 2, 9, 7, 24
    img = torch.randn(N, 3, H, W, device=DEV)
    wst, bst = torch.randn(O, 3, 3, 3, device=DEV) * 0.2, torch.zeros(O, device=DEV)
    OH, OW = (H + 2 - 3) // 2 + 1, (W + 2 - 3) // 2 + 1
    b0, y0 = guarded(N * OH * OW * 32)
    d = L.DfqStemQ8Desc()
    d.x, d.w, d.bias, d.in_slot, d.y0, d.out_slot0 = img.data_ptr(), wst.data_ptr(), bst.data_ptr(), in_slot.data_ptr(), y0.data_ptr(), s0.data_ptr()
    d.N, d.H, d.W, d.O, d.stride, d.pad, d.relu = N, H, W, O, 2, 1, 1
    L.check(lib.dfq_stem_q8(ctypes.byref(d), L.stream()), "dfq_stem_q8")
    px, C = 77, 40
    xa = to_codes_nhwc(torch.from_numpy(rng.integers(0, 256, (px, C)).astype(np.uint8))).reshape(-1).to(DEV)
    ba, ya = guarded(px * 48)
    bb, yb = guarded(px * 48)
    L.check(lib.dfq_add_q8(xa.data_ptr(), xa.data_ptr(), in_slot.data_ptr(), s1.data_ptr(), ya.data_ptr(), s0.data_ptr(),
                           yb.data_ptr(), s1.data_ptr(), px, C, None, None, L.stream()), "dfq_add_q8")
    Nn, Cm, HW = 3, 100, 13
    xm = to_codes_nhwc(torch.from_numpy(rng.integers(0, 256, (Nn, Cm, HW, 1)).astype(np.uint8))).reshape(-1).to(DEV)
    bm, ym = guarded(Nn * 112)
    bmf, ymf = guarded(Nn * Cm, torch.float32)
    L.check(lib.dfq_mean_q8(xm.data_ptr(), in_slot.data_ptr(), ym.data_ptr(), s0.data_ptr(), ymf.data_ptr(), Nn, Cm, HW,
                            L.stream()), "dfq_mean_q8")
    torch.cuda.synchronize()
    assert intact(b0, N * OH * OW * 32) and intact(ba, px * 48) and intact(bb, px * 48)
    assert intact(bm, Nn * 112) and intact(bmf, 4 * Nn * Cm)


@pytest.fixture(scope="module")
def static_model():
    model, graph, bottoms, targ, tr = dfq_quantised("mobilenetv2", DEV, np.load(os.path.join(GOLD, "mobilenetv2_dfq.npz")))
    for m in model.modules():
        if hasattr(m, "set_update_stat"):
            m.set_update_stat(False)
    return model, tr


def test_static_plan_reproduces_the_reference_codes_edge_by_edge(static_model):
    """Every producer -> consumer edge of MobileNetV2: fed the reference's input codes, a q8 kernel must emit the
    codes the reference's consumer QuantMeasure produced (tests/golden/eval_tf_static_mobilenetv2.npz)."""
    from data_free_quantization_b200.eval_q8 import EvalPlanQ8
    model, tr = static_model
    tf = np.load(os.path.join(GOLD, "eval_tf_static_mobilenetv2.npz"))
    plan = EvalPlanQ8(model, tr, batch=1)

    def gold_codes(key):
        return tf[f"L/{key[1]}/xq"] if key[0] == "L" else tf[f"T/{key[1]}/xq{key[2]}"]

    def gold_range(key):
        return tf[f"L/{key[1]}/range"] if key[0] == "L" else tf[f"T/{key[1]}/range{key[2]}"]
    # the plan's static ranges are the reference's
    slots = plan.slots.cpu().numpy()
    for idx, key in plan.slot_key.items():
        want = gold_range(key).astype(np.float32)
        assert np.allclose(slots[idx, :2], want, rtol=1e-6, atol=1e-7), key      # (sqrt on add branches: utils/minmax.py)
        plan.set_range(idx, float(want[0]), float(want[1]))
    x = torch.randn(4, 3, 224, 224, generator=torch.Generator().manual_seed(2))[:1].to(DEV)
    worst, edges = 0.0, 0
    for name, io in plan.step_io.items():
        if io["kind"] == "stem":
            res = plan.run_step(name, image=x)
        else:
            res = plan.run_step(name, in_codes=[gold_codes(plan.slot_key[i]) for i in io["in_slots"]])
        for idx in io["out_slots"]:
            worst = max(worst, check_codes(res[idx].numpy(), gold_codes(plan.slot_key[idx]), f"{name} -> {plan.slot_key[idx]}"))
            edges += 1
        if "fp32" in res:                                       # the classifier's logits
            want = tf["logits"]
            assert np.abs(res["fp32"].numpy().reshape(want.shape) - want).max() <= 1e-3 * np.abs(want).max()
    print(f"{edges} producer->consumer edges; worst fraction of codes off by one: {worst:.2e}")
    assert edges >= 73          # 74 measures; the stem's own input is the fp32 image


def test_static_plan_end_to_end_against_reference(static_model):
    from data_free_quantization_b200.eval_q8 import EvalPlanQ8
    model, tr = static_model
    g = np.load(os.path.join(GOLD, "eval_static_mobilenetv2.npz"))
    info = json.load(open(os.path.join(GOLD, "eval_static_mobilenetv2.json")))
    n = info["images"]
    plan = EvalPlanQ8(model, tr, batch=n)
    x = torch.randn(n, 3, 224, 224, generator=torch.Generator().manual_seed(7)).to(DEV)
    got = plan(x).cpu().numpy().reshape(g["logits"].shape)
    shift = float(np.abs(got - g["logits"]).max() / np.abs(g["logits"]).max())
    agree = float((got.argmax(1) == g["argmax"]).mean())
    print(f"static plan vs reference: logits {shift:.2e} of the largest (reference vs itself under 1e-7 input noise: "
          f"{info['reference_logit_shift_under_perturbation']:.2e}), top-1 agreement {agree:.4f}")
    assert agree >= 0.999
    assert shift <= 2.0 * info["reference_logit_shift_under_perturbation"]
    # replayed as one CUDA graph: identical
    replay = plan.capture()
    plan.input.buf.copy_(x)
    replay()
    torch.cuda.synchronize()
    assert np.array_equal(plan.output.buf.cpu().numpy().reshape(got.shape), got)


@pytest.mark.parametrize("N,H,W,O,stride,dual", [(3, 64, 48, 32, 2, False), (2, 9, 7, 24, 2, True), (1, 11, 13, 32, 1, True),
                                                 (2, 1, 1, 5, 2, False)])
def test_stem_q8_integer_kernel(oracle, N, H, W, O, stride, dual):
    """The stem on exact integer sums (dp4a; DfqStemQ8Desc.wq) against a float64 convolution of the fake-quantised
    operands: borders (taps in the zero padding are absent, not code 128), ragged O, two output grids."""
    from data_free_quantization_b200 import _lib as L
    from data_free_quantization_b200.eval_q8 import from_codes_nhwc, pack_stem
    lib = L.lib()
    rng = np.random.default_rng(31 + H)
    img = (rng.standard_normal((N, 3, H, W)) * 1.2).astype(np.float32)
    lo_i, hi_i = -2.11790393, 2.64
    in_slot, _ = slot(lo_i, hi_i)
    xq = oracle.fake_quant(img, 8, float(np.float32(lo_i)), float(np.float32(hi_i)))
    w = (rng.standard_normal((O, 3, 3, 3)) * 0.3).astype(np.float32)
    w = oracle.fake_quant(w, 8, float(w.min()), float(w.max()))            # quantize_targ_layer: on a grid
    bias = (rng.standard_normal(O) * 0.2).astype(np.float32)
    ref = np.maximum(torch.nn.functional.conv2d(torch.from_numpy(xq).double(), torch.from_numpy(w).double(),
                                                torch.from_numpy(bias).double(), stride, 1).numpy(), 0)
    OH, OW = ref.shape[2:]
    Cp = (O + 15) // 16 * 16
    outs = [(0.0, float(ref.max()) * 0.95 + 0.1), (0.05, float(ref.max()) * 0.6 + 0.1)]
    (s0, _), (s1, _) = slot(*outs[0]), slot(*outs[1])
    xd, wd, bd = torch.from_numpy(img).to(DEV), torch.from_numpy(w).to(DEV), torch.from_numpy(bias).to(DEV)
    pk = pack_stem(wd)
    assert pk is not None
    y0 = torch.full((N * OH * OW * Cp,), 3, dtype=torch.int8, device=DEV)
    y1 = torch.full((N * OH * OW * Cp,), 3, dtype=torch.int8, device=DEV)
    d = L.DfqStemQ8Desc()
    d.x, d.w, d.bias, d.in_slot, d.y0, d.out_slot0 = xd.data_ptr(), wd.data_ptr(), bd.data_ptr(), in_slot.data_ptr(), y0.data_ptr(), s0.data_ptr()
    if dual:
        d.y1, d.out_slot1 = y1.data_ptr(), s1.data_ptr()
    d.N, d.H, d.W, d.O, d.stride, d.pad, d.relu = N, H, W, O, stride, 1, 1
    d.wq, d.s_w, d.c_w = pk[0].data_ptr(), pk[1], pk[2]
    L.check(lib.dfq_stem_q8(ctypes.byref(d), L.stream()), "dfq_stem_q8")
    torch.cuda.synchronize()
    for yb, (lo, hi) in list(zip((y0, y1), outs))[:2 if dual else 1]:
        _q, want = oracle.fake_quant(ref.astype(np.float32), 8, float(np.float32(lo)), float(np.float32(hi)), want_codes=True)
        check_codes(from_codes_nhwc(yb.cpu(), ref.shape).numpy(), want, f"stem i8 {H}x{W}")
        assert not yb.cpu().view(-1, Cp)[:, O:].any()          # padding channels: s8 zero
    # the dynamic-range plan's form: the same sums as fp32 NCHW + encoded per-sample extrema
    yf = torch.full((N, O, OH, OW), float("nan"), device=DEV)
    stats = torch.zeros(N, 2, dtype=torch.int32, device=DEV)
    d2 = L.DfqStemQ8Desc()
    d2.x, d2.w, d2.bias, d2.in_slot = xd.data_ptr(), wd.data_ptr(), bd.data_ptr(), in_slot.data_ptr()
    d2.N, d2.H, d2.W, d2.O, d2.stride, d2.pad, d2.relu = N, H, W, O, stride, 1, 1
    d2.wq, d2.s_w, d2.c_w, d2.yf, d2.out_stats = pk[0].data_ptr(), pk[1], pk[2], yf.data_ptr(), stats.data_ptr()
    L.check(lib.dfq_stem_q8(ctypes.byref(d2), L.stream()), "dfq_stem_q8 (fp32 out)")
    torch.cuda.synchronize()
    got = yf.cpu().numpy()
    assert np.abs(got - ref).max() <= 2e-6 * max(1.0, np.abs(ref).max())
    enc = stats.cpu().numpy().view(np.uint32)
    dec = lambda u: np.where(u & 0x80000000, u ^ 0x80000000, ~u).astype(np.uint32).view(np.float32)
    assert np.array_equal(-dec(enc[:, 0]), got.reshape(N, -1).min(1)) and np.array_equal(dec(enc[:, 1]), got.reshape(N, -1).max(1))
    # off-grid weights are refused by the packer (the fp32 kernel takes them)
    assert pack_stem(wd * 1.0001 + torch.randn_like(wd) * 1e-3) is None
