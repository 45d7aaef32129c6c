"""GPU: DeepLabV3+ (MobileNetV2 backbone) on the static-range u8 plan (eval_q8.EvalPlanQ8) against the UNMODIFIED
reference run with set_update_stat(False) on every QuantMeasure (utils/quantize.py:125-126; goldens from
oracle/gen_eval_golden.py --model deeplab --what tf_static,static).

What DeepLab adds to MobileNetV2's graph: F.pad ahead of the depthwise convolutions' measures (folded into the
convolution: border = code of 0.0), a dilated depthwise block, dense atrous 3 x 3 convolutions and the decoder's 3 x 3 on
tcgen05 with code output, the global average pool of a RAW tensor, bilinear interpolation and channel concatenation of
code tensors.  The bar is the one of tests/test_gpu_q8.py: identical codes on every producer -> consumer edge, off by one
code on at most 2e-3 of the elements, never by more."""
import json
import os

import numpy as np
import pytest
import torch

from helpers import dfq_quantised
from test_gpu_q8 import check_codes

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def static_deeplab():
    model, graph, bottoms, targ, tr = dfq_quantised("deeplab", DEV, np.load(os.path.join(GOLD, "deeplab_dfq.npz")))
    for m in model.modules():
        if hasattr(m, "set_update_stat"):
            m.set_update_stat(False)
    return model, tr


def test_deeplab_static_plan_edge_by_edge(static_deeplab):
    from data_free_quantization_b200.eval_q8 import EvalPlanQ8
    model, tr = static_deeplab
    tf = np.load(os.path.join(GOLD, "eval_tf_static_deeplab.npz"))
    plan = EvalPlanQ8(model, tr, batch=1, in_shape=(3, 129, 129))

    def gold_codes(key):
        return tf[f"L/{key[1]}/xq"] if key[0] == "L" else tf[f"T/{key[1]}/xq{key[2]}"]

    def gold_range(key):
        return tf[f"L/{key[1]}/range"] if key[0] == "L" else tf[f"T/{key[1]}/range{key[2]}"]
    slots = plan.slots.cpu().numpy()
    assert len(plan.slot_key) == len(plan.measures)
    for idx, key in plan.slot_key.items():
        want = gold_range(key).astype(np.float32)
        assert np.allclose(slots[idx, :2], want, rtol=1e-6, atol=1e-7), key
        plan.set_range(idx, float(want[0]), float(want[1]))
    x = torch.randn(2, 3, 129, 129, generator=torch.Generator().manual_seed(2))[:1].to(DEV)
    worst, edges, kinds = 0.0, 0, {}
    for name, io in plan.step_io.items():
        if io["kind"] == "pool":
            continue                    # reads the raw fp32 side output of its producer: covered by the end-to-end test
        if io["kind"] == "stem":
            res = plan.run_step(name, image=x)
        elif io["kind"] == "pad":
            # F.pad ahead of the consumer's measure: fed the interior of the reference's padded codes, the op must
            # rebuild the whole padded tensor -- the border is ONE code, code(0.0) on that grid -- exactly
            l, r, t, b = io["pad"]
            inner = []
            for idx in io["pre_slots"]:
                full = gold_codes(plan.slot_key[idx])
                inner.append(full[:, :, t:full.shape[2] - b, l:full.shape[3] - r])
            res = plan.run_step(name, in_codes=inner)
            for idx in io["out_slots"]:
                assert np.array_equal(res[idx].numpy(), gold_codes(plan.slot_key[idx])), (name, plan.slot_key[idx])
                edges += 1
            kinds["pad"] = kinds.get("pad", 0) + 1
            continue
        else:
            res = plan.run_step(name, in_codes=[gold_codes(plan.slot_key[i]) for i in io["in_slots"]])
        # Bilinear interpolation between ADJACENT codes with weights 1/4, 1/2, 3/4 (align_corners 9 -> 33: the scale is
        # exactly 1/4) lands exactly half-way between two codes of the consumer's grid when that grid is the operand's own
        # ([0, 6] on both sides here): the rounding of such a tie follows the last bit of the blend, where ATen's CPU
        # kernel (combined weights, vectorised) and its CUDA kernel (separable) differ from each other too.  Still never
        # more than one code; the share of elements is what is relaxed for this op.
        frac = 1e-2 if io["kind"] == "interp" else 2e-3
        for idx in io["out_slots"]:
            worst = max(worst, check_codes(res[idx].numpy(), gold_codes(plan.slot_key[idx]),
                                           f"{name} -> {plan.slot_key[idx]}", max_frac=frac))
            edges += 1
        kinds[io["kind"]] = kinds.get(io["kind"], 0) + 1
        if "fp32" in res and io["out"] is plan.output:
            want = tf["logits"]
            got = res["fp32"].numpy()[:, :, ::4, ::4]
            assert np.abs(got - want).max() <= 1e-3 * np.abs(want).max()
    print(f"{edges} edges over {kinds}; worst fraction of codes off by one: {worst:.2e}")
    assert kinds.get("dense", 0) == 5 and kinds.get("interp", 0) == 3 and kinds.get("cat", 0) == 2
    assert edges >= 85


def test_deeplab_dense_layers_tma_form(static_deeplab, monkeypatch):
    """The TMA-fed form of the dense k x k kernel (DFQ_CONV_Q8_TMA=1: 4-D tensor-map boxes of the code tensor per tap,
    zero padding by the TMA's out-of-bounds fill, warp-specialised tcgen05 pipeline) against the same goldens."""
    from data_free_quantization_b200.eval_q8 import EvalPlanQ8
    model, tr = static_deeplab
    monkeypatch.setenv("DFQ_CONV_Q8_TMA", "1")
    tf = np.load(os.path.join(GOLD, "eval_tf_static_deeplab.npz"))
    plan = EvalPlanQ8(model, tr, batch=1, in_shape=(3, 129, 129))
    gold = lambda key: tf[f"L/{key[1]}/xq"] if key[0] == "L" else tf[f"T/{key[1]}/xq{key[2]}"]
    rng = lambda key: tf[f"L/{key[1]}/range"] if key[0] == "L" else tf[f"T/{key[1]}/range{key[2]}"]
    for idx, key in plan.slot_key.items():
        plan.set_range(idx, float(np.float32(rng(key)[0])), float(np.float32(rng(key)[1])))
    done = 0
    for name, io in plan.step_io.items():
        if io["kind"] != "dense":
            continue
        res = plan.run_step(name, in_codes=[gold(plan.slot_key[i]) for i in io["in_slots"]])
        for idx in io["out_slots"]:
            check_codes(res[idx].numpy(), gold(plan.slot_key[idx]), f"{name} (tma) -> {plan.slot_key[idx]}")
            done += 1
    assert done == 5


def test_deeplab_static_plan_end_to_end(static_deeplab):
    from data_free_quantization_b200.eval_q8 import EvalPlanQ8
    model, tr = static_deeplab
    g = np.load(os.path.join(GOLD, "eval_static_deeplab.npz"))
    info = json.load(open(os.path.join(GOLD, "eval_static_deeplab.json")))
    n = info["images"]
    plan = EvalPlanQ8(model, tr, batch=n, in_shape=(3, 129, 129))
    x = torch.randn(n, 3, 129, 129, generator=torch.Generator().manual_seed(7)).to(DEV)
    got = plan(x).cpu().numpy()
    assert got.shape == (n, 21, 129, 129)
    want = g["logits"]
    shift = float(np.abs(got[:, :, ::4, ::4] - want).max() / np.abs(want).max())
    agree = float((got.argmax(1) == g["argmax"]).mean())
    print(f"DeepLab static plan vs reference: logits {shift:.2e} of the largest, pixel classes agree {agree:.4f}")
    assert agree >= 0.999
    assert shift <= 3e-2
    replay = plan.capture()
    plan.input.buf.copy_(x)
    replay()
    torch.cuda.synchronize()
    assert np.array_equal(plan.output.buf.cpu().numpy(), got)


@pytest.mark.parametrize("N,C,H,W,O,pad,dil,relu,dual,forms", [
    (2, 40, 19, 21, 24, 1, 1, 1, True, (0, 1, 2)),        # ragged channels / ragged tiles, all three kernels
    (1, 304, 33, 17, 256, 1, 1, 1, False, (0, 2)),         # the decoder's channel counts on a small image
    (3, 16, 9, 40, 272, 1, 1, 0, True, (0, 2)),            # two output passes, last one ragged
    (1, 64, 17, 17, 48, 2, 2, 0, False, (0, 1)),           # dilated (ASPP-like): gather and TMA forms only
])
def test_dense_conv_q8_forms_against_float64_restatement(oracle, N, C, H, W, O, pad, dil, relu, dual, forms):
    """dfq_conv_q8 in its three forms (register gather, TMA boxes per tap, halo tile) on random codes: the output codes
    must be those of the float64 convolution of the dequantised operands (zero padding = exact zeros), never more than
    one code off."""
    import ctypes
    import torch.nn.functional as F
    from data_free_quantization_b200 import _lib as L, pw_i8
    from data_free_quantization_b200.eval_q8 import (PackedConvQ8, choose_o_cap, from_codes_nhwc, pack_conv_halo,
                                                     to_codes_nhwc)
    from test_gpu_q8 import slot
    rng = np.random.default_rng(N * 1000 + C + O)
    xc = rng.integers(0, 256, (N, C, H, W)).astype(np.uint8)
    lo_x, hi_x = -0.9, 2.3
    in_slot, sx = slot(lo_x, hi_x)
    w0 = (rng.standard_normal((O, C, 3, 3)) * 0.1).astype(np.float32)
    w, _wq = oracle.fake_quant(w0, 8, float(w0.min()), float(w0.max()), want_codes=True)
    bias = (rng.standard_normal(O) * 0.3).astype(np.float32)
    x_hat = (xc.astype(np.float32) * np.float32(sx)).astype(np.float32) + np.float32(lo_x)          # fl(fl(q s) + min)
    exact = F.conv2d(torch.from_numpy(x_hat).double(), torch.from_numpy(w).double(), torch.from_numpy(bias).double(),
                     1, pad, dil).numpy()
    if relu:
        exact = np.maximum(exact, 0)
    OH, OW = exact.shape[2:]
    ranges = [(float(exact.min()) * 0.8, float(exact.max()) * 0.7), (0.0, float(exact.max()) + 0.5)]
    want = [oracle.fake_quant(exact.astype(np.float32), 8, float(np.float32(lo)), float(np.float32(hi)), want_codes=True)[1]
            for lo, hi in ranges]
    s0, _ = slot(*ranges[0]); s1, _ = slot(*ranges[1])
    wd, bd = torch.from_numpy(w).to(DEV), torch.from_numpy(bias).to(DEV)
    x = to_codes_nhwc(torch.from_numpy(xc)).reshape(-1).to(DEV)
    Op = (O + 15) // 16 * 16
    for form in forms:
        if form == 2:
            pk = pack_conv_halo(wd)
        elif form == 1:
            pk = PackedConvQ8(wd, o_cap=choose_o_cap(N * OH * OW, O))
        else:
            pk = pw_i8.pack_weight_conv(wd, None, o_cap=pw_i8.choose_o_cap(N * OH * OW, O))
        y0 = torch.full((N * OH * OW * Op,), 77, dtype=torch.int8, device=DEV)
        y1 = torch.full((N * OH * OW * Op,), 77, dtype=torch.int8, device=DEV)
        d = L.DfqConvQ8Desc()
        d.x, d.wpk, d.wr, d.wrt, d.bias = x.data_ptr(), pk.wpk.data_ptr(), pk.wr.data_ptr(), pk.wrt.data_ptr(), bd.data_ptr()
        d.in_slot, d.y0, d.out_slot0 = in_slot.data_ptr(), y0.data_ptr(), s0.data_ptr()
        if dual:
            d.y1, d.out_slot1 = y1.data_ptr(), s1.data_ptr()
        d.N, d.C, d.H, d.W, d.O, d.OH, d.OW = N, C, H, W, O, OH, OW
        d.kh, d.kw, d.stride, d.pad, d.dil, d.relu = 3, 3, 1, pad, dil, relu
        d.s_w, d.c_w, d.o_cap, d.tma = pk.s_w, pk.c_w, pk.o_cap, form
        L.check(L.lib().dfq_conv_q8(ctypes.byref(d), L.stream()), f"dfq_conv_q8 form {form}")
        torch.cuda.synchronize()
        for t, yb in enumerate((y0, y1)):
            if t == 1 and not dual:
                assert bool((y1 == 77).all())
                break
            got = from_codes_nhwc(yb.cpu(), (N, O, OH, OW)).numpy()
            check_codes(got, want[t], f"form {form} grid {t}")
            assert not yb.cpu().view(N * OH * OW, Op)[:, O:].any(), "pad channels must be s8 zero"


def test_movement_kernels_on_codes_against_torch(oracle):
    """dfq_interp_q8 / dfq_regrid_q8 / dfq_pool_q8 / dfq_pad_q8 and the generic (dilated) depthwise kernel on random
    code tensors with ragged channel counts, against torch on the dequantised values."""
    import ctypes
    import torch.nn.functional as F
    from data_free_quantization_b200 import _lib as L
    from data_free_quantization_b200.eval_q8 import from_codes_nhwc, pack_depthwise, to_codes_nhwc
    from test_gpu_q8 import slot
    lib = L.lib()
    rng = np.random.default_rng(77)
    deq = lambda codes, lo, s: (codes.astype(np.float32) * np.float32(s)).astype(np.float32) + np.float32(lo)
    quant = lambda v, lo, hi: oracle.fake_quant(np.ascontiguousarray(v, np.float32), 8, float(np.float32(lo)),
                                                float(np.float32(hi)), want_codes=True)[1]
    # ---- bilinear interpolation, both align_corners modes, codes and fp32 out
    N, C, H, W, OH, OW = 2, 21, 7, 9, 20, 13
    xc = rng.integers(0, 256, (N, C, H, W)).astype(np.uint8)
    lo, hi = -1.5, 3.0
    sl, sx = slot(lo, hi)
    so, _ = slot(-1.0, 2.5)
    x = to_codes_nhwc(torch.from_numpy(xc)).reshape(-1).to(DEV)
    for align in (0, 1):
        ref = F.interpolate(torch.from_numpy(deq(xc, lo, sx)), size=(OH, OW), mode="bilinear", align_corners=bool(align)).numpy()
        y = torch.full((N * OH * OW * 32,), 77, dtype=torch.int8, device=DEV)
        yf = torch.full((N, C, OH, OW), float("nan"), device=DEV)
        L.check(lib.dfq_interp_q8(x.data_ptr(), sl.data_ptr(), N, C, H, W, OH, OW, align, y.data_ptr(), so.data_ptr(), None,
                                  None, yf.data_ptr(), L.stream()), "dfq_interp_q8")
        torch.cuda.synchronize()
        assert np.abs(yf.cpu().numpy() - ref).max() <= 2e-6 * np.abs(ref).max()
        check_codes(from_codes_nhwc(y.cpu(), (N, C, OH, OW)).numpy(), quant(ref, -1.0, 2.5), f"interp align={align}")
        assert not y.cpu().view(-1, 32)[:, C:].any()
    # ---- one operand of a concatenation: 40 channels at offset 16 of 64-channel rows, the tail cleared
    G, C, c_off, cp_out = 300, 40, 16, 64
    xc = rng.integers(0, 256, (G, C)).astype(np.uint8)
    x = to_codes_nhwc(torch.from_numpy(xc)).reshape(-1).to(DEV)
    y = torch.full((G, cp_out), 77, dtype=torch.int8, device=DEV)
    L.check(lib.dfq_regrid_q8(x.data_ptr(), sl.data_ptr(), G, C, y.data_ptr(), so.data_ptr(), None, None, cp_out, c_off, 1,
                              L.stream()), "dfq_regrid_q8")
    torch.cuda.synchronize()
    got = y.cpu()
    assert bool((got[:, :c_off] == 77).all()) and not got[:, c_off + C:].any()
    want = quant(deq(xc, lo, sx), -1.0, 2.5)
    assert np.array_equal((got[:, c_off:c_off + C].to(torch.int16) + 128).numpy(), want)      # elementwise: exact
    # ---- global average pool of fp32 rows
    N, HW, C = 3, 1089, 40
    xf = (rng.standard_normal((N, HW, C)) * 2).astype(np.float32)
    y = torch.full((N, 48), 77, dtype=torch.int8, device=DEV)
    L.check(lib.dfq_pool_q8(torch.from_numpy(xf).to(DEV).data_ptr(), C, N, HW, C, y.data_ptr(), so.data_ptr(), None, None,
                            L.stream()), "dfq_pool_q8")
    torch.cuda.synchronize()
    check_codes((y.cpu()[:, :C].to(torch.int16) + 128).numpy(), quant(xf.astype(np.float64).mean(1), -1.0, 2.5), "pool")
    assert not y.cpu()[:, C:].any()
    # ---- depthwise 3 x 3 with dilation 2 (generic kernel), stride 1 and 2
    N, C, H, W = 2, 40, 15, 18
    xc = rng.integers(0, 256, (N, C, H, W)).astype(np.uint8)
    x = to_codes_nhwc(torch.from_numpy(xc)).reshape(-1).to(DEV)
    w0 = (rng.standard_normal((C, 1, 3, 3)) * 0.3).astype(np.float32)
    w, _ = oracle.fake_quant(w0, 8, float(w0.min()), float(w0.max()), want_codes=True)
    wrow, dsw, dcw = pack_depthwise(torch.from_numpy(w).to(DEV))
    bias = (rng.standard_normal(C) * 0.2).astype(np.float32)
    bp = torch.zeros(48, device=DEV); bp[:C] = torch.from_numpy(bias).to(DEV)
    for stride in (1, 2):
        ref = F.conv2d(torch.from_numpy(deq(xc, lo, sx)).double(), torch.from_numpy(w).double(),
                       torch.from_numpy(bias).double(), stride, 2, 2, C).numpy()
        OHd, OWd = ref.shape[2:]
        so2, _ = slot(float(ref.min()) * 0.9, float(ref.max()) * 0.9)
        y = torch.full((N * OHd * OWd * 48,), 77, dtype=torch.int8, device=DEV)
        d = L.DfqDwQ8Desc()
        d.x, d.wrow, d.bias, d.in_slot, d.y0, d.out_slot0 = x.data_ptr(), wrow.data_ptr(), bp.data_ptr(), sl.data_ptr(), y.data_ptr(), so2.data_ptr()
        d.N, d.H, d.W, d.C, d.stride, d.relu, d.s_w, d.c_w, d.pad, d.dil = N, H, W, C, stride, 0, dsw, dcw, 2, 2
        L.check(lib.dfq_dw_q8(ctypes.byref(d), L.stream()), "dfq_dw_q8")
        torch.cuda.synchronize()
        check_codes(from_codes_nhwc(y.cpu(), (N, C, OHd, OWd)).numpy(),
                    quant(ref, float(ref.min()) * 0.9, float(ref.max()) * 0.9), f"dilated depthwise stride {stride}")


def test_static_plan_on_other_image_sizes(static_deeplab):
    """Sizes no golden holds (ragged tiles, non-square images, odd batch): the u8 plan against the reference-shaped module
    forward of the same static-mode model (QuantMeasure + F.conv2d per layer, utils/quantize.py:250-256)."""
    from data_free_quantization_b200.eval_q8 import EvalPlanQ8
    from data_free_quantization_b200.utils import quantize as Q
    model, tr = static_deeplab
    for n, h, w in ((3, 200, 264), (1, 97, 131)):
        x = torch.randn(n, 3, h, w, generator=torch.Generator().manual_seed(n + h)).to(DEV)
        plan = EvalPlanQ8(model, tr, batch=n, in_shape=(3, h, w))
        got = plan(x).clone()
        Q.FUSED_EVAL = False
        try:
            with torch.no_grad():
                ref = model(x)
        finally:
            Q.FUSED_EVAL = True
        agree = float((got.argmax(1) == ref.argmax(1)).float().mean())
        shift = float((got - ref).abs().max() / ref.abs().max())
        print(f"{n}x3x{h}x{w}: classes agree {agree:.4f}, logits {shift:.2e} of the largest")
        assert got.shape == ref.shape and agree >= 0.998 and shift <= 0.1


@pytest.mark.parametrize("N,C,H,W,OH,OW,align", [(2, 256, 33, 33, 129, 129, 1), (1, 21, 129, 129, 513, 513, 1),
                                                   (2, 21, 7, 9, 20, 13, 0), (3, 40, 5, 6, 17, 31, 1),
                                                   (1, 19, 4, 4, 9, 9, 0)])
def test_interp_row_walking_kernels_equal_the_one_row_kernels(N, C, H, W, OH, OW, align, monkeypatch):
    """The upscaling forms of dfq_interp_q8 (a thread walks four output rows and forms the horizontal blends once per
    input row pair) against the one-output-per-thread kernels (DFQ_Q8_INTERP_ROWS=0): same operations per value, so
    the same bytes -- codes on two grids and the fp32 output."""
    from data_free_quantization_b200 import _lib as L
    from data_free_quantization_b200.eval_q8 import to_codes_nhwc
    from test_gpu_q8 import slot
    lib = L.lib()
    rng = np.random.default_rng(5 + C)
    Cp = (C + 15) // 16 * 16
    x = to_codes_nhwc(torch.from_numpy(rng.integers(0, 256, (N, C, H, W)).astype(np.uint8))).reshape(-1).to(DEV)
    (sl, _), (s0, _), (s1, _) = slot(-1.5, 3.0), slot(-1.0, 2.5), slot(-2.0, 3.5)
    res = []
    for rows in ("0", "1"):
        monkeypatch.setenv("DFQ_Q8_INTERP_ROWS", rows)
        y0 = torch.full((N * OH * OW * Cp,), 77, dtype=torch.int8, device=DEV)
        y1 = torch.full((N * OH * OW * Cp,), 77, dtype=torch.int8, device=DEV)
        yf = torch.full((N, C, OH, OW), float("nan"), device=DEV)
        L.check(lib.dfq_interp_q8(x.data_ptr(), sl.data_ptr(), N, C, H, W, OH, OW, align, y0.data_ptr(), s0.data_ptr(),
                                  y1.data_ptr(), s1.data_ptr(), yf.data_ptr(), L.stream()), "dfq_interp_q8")
        torch.cuda.synchronize()
        res.append((y0, y1, yf))
    for a, b in zip(*res):
        assert torch.equal(a, b)
    assert not bool((res[1][0] == 77).all()) and not torch.isnan(res[1][2]).any()


@pytest.mark.parametrize("G,C,c_off,cp_out,fill,dual", [(1089, 256, 512, 1280, 0, False), (4000, 48, 256, 304, 1, True),
                                                          (300, 40, 16, 64, 1, False), (77, 24, 0, 64, 0, True)])
def test_regrid_chunk_kernel_equals_the_word_kernel(G, C, c_off, cp_out, fill, dual, monkeypatch):
    """dfq_regrid_q8 on whole 16-channel chunks against the word-per-thread kernel (DFQ_Q8_REGRID_V16=0): the same
    bytes in the operand's slice, and nothing outside it (the last case is not eligible and stays on the word kernel)."""
    from data_free_quantization_b200 import _lib as L
    from data_free_quantization_b200.eval_q8 import to_codes_nhwc
    from test_gpu_q8 import slot
    lib = L.lib()
    rng = np.random.default_rng(C)
    x = to_codes_nhwc(torch.from_numpy(rng.integers(0, 256, (G, C)).astype(np.uint8))).reshape(-1).to(DEV)
    (sl, _), (s0, _), (s1, _) = slot(-1.5, 3.0), slot(-1.0, 2.5), slot(-2.0, 3.5)
    res = []
    for v16 in ("0", "1"):
        monkeypatch.setenv("DFQ_Q8_REGRID_V16", v16)
        y0 = torch.full((G, cp_out), 77, dtype=torch.int8, device=DEV)
        y1 = torch.full((G, cp_out), 77, dtype=torch.int8, device=DEV)
        L.check(lib.dfq_regrid_q8(x.data_ptr(), sl.data_ptr(), G, C, y0.data_ptr(), s0.data_ptr(),
                                  y1.data_ptr() if dual else None, s1.data_ptr() if dual else None, cp_out, c_off, fill,
                                  L.stream()), "dfq_regrid_q8")
        torch.cuda.synchronize()
        res.append((y0, y1))
    for a, b in zip(*res):
        assert torch.equal(a, b)
    y0 = res[1][0]
    assert bool((y0[:, :c_off] == 77).all()) and not bool((y0[:, c_off:c_off + C] == 77).all())
    if not fill:
        assert bool((y0[:, c_off + (C + 3) // 4 * 4:] == 77).all())


@pytest.mark.parametrize("N,C,H,W,dual", [(2, 960, 37, 37, False), (1, 40, 15, 18, True), (3, 16, 5, 6, False),
                                          (1, 100, 9, 29, True)])
def test_dilated_depthwise_strip_kernel_equals_the_generic_kernel(oracle, N, C, H, W, dual, monkeypatch):
    """Dilation 2, no padding, stride 1 (DeepLab's dilated block after F.pad) on the strip kernel (taps = bytes k, k + 2,
    k + 4 of the transposed row string) against the one-pixel-per-thread kernel (DFQ_Q8_DW_DIL2=0): the same integer
    sums in the same fp32 expression, so the same bytes; and against a float64 convolution."""
    import ctypes
    import torch.nn.functional as F
    from data_free_quantization_b200 import _lib as L
    from data_free_quantization_b200.eval_q8 import from_codes_nhwc, pack_depthwise, to_codes_nhwc
    from test_gpu_q8 import slot
    lib = L.lib()
    rng = np.random.default_rng(C + W)
    Cp = (C + 15) // 16 * 16
    lo, hi = -1.5, 3.0
    sl, sx = slot(lo, hi)
    xc = rng.integers(0, 256, (N, C, H, W)).astype(np.uint8)
    x = to_codes_nhwc(torch.from_numpy(xc)).reshape(-1).to(DEV)
    w0 = (rng.standard_normal((C, 1, 3, 3)) * 0.3).astype(np.float32)
    w, _ = oracle.fake_quant(w0, 8, float(w0.min()), float(w0.max()), want_codes=True)
    wrow, dsw, dcw = pack_depthwise(torch.from_numpy(w).to(DEV))
    bias = (rng.standard_normal(C) * 0.2).astype(np.float32)
    bp = torch.zeros(Cp, device=DEV); bp[:C] = torch.from_numpy(bias).to(DEV)
    xf = (xc.astype(np.float32) * np.float32(sx)).astype(np.float32) + np.float32(lo)
    ref = np.maximum(F.conv2d(torch.from_numpy(xf).double(), torch.from_numpy(w).double(), torch.from_numpy(bias).double(),
                              1, 0, 2, C).numpy(), 0)
    OH, OW = ref.shape[2:]
    outs = [(0.0, float(ref.max()) * 0.9 + 0.1), (0.02, float(ref.max()) * 0.5 + 0.1)]
    (s0, _), (s1, _) = slot(*outs[0]), slot(*outs[1])
    res = []
    for fast in ("0", "1"):
        monkeypatch.setenv("DFQ_Q8_DW_DIL2", fast)
        y0 = torch.full((N * OH * OW * Cp,), 77, dtype=torch.int8, device=DEV)
        y1 = torch.full((N * OH * OW * Cp,), 77, dtype=torch.int8, device=DEV)
        d = L.DfqDwQ8Desc()
        d.x, d.wrow, d.bias, d.in_slot, d.y0, d.out_slot0 = x.data_ptr(), wrow.data_ptr(), bp.data_ptr(), sl.data_ptr(), y0.data_ptr(), s0.data_ptr()
        if dual:
            d.y1, d.out_slot1 = y1.data_ptr(), s1.data_ptr()
        d.N, d.H, d.W, d.C, d.stride, d.relu, d.s_w, d.c_w, d.pad, d.dil = N, H, W, C, 1, 1, dsw, dcw, 0, 2
        L.check(lib.dfq_dw_q8(ctypes.byref(d), L.stream()), "dfq_dw_q8")
        torch.cuda.synchronize()
        res.append((y0, y1))
    for a, b in zip(*res):
        assert torch.equal(a, b)
    for yb, (l, h) in list(zip(res[1], outs))[:2 if dual else 1]:
        want = oracle.fake_quant(ref.astype(np.float32), 8, float(np.float32(l)), float(np.float32(h)), want_codes=True)[1]
        check_codes(from_codes_nhwc(yb.cpu(), (N, C, OH, OW)).numpy(), want, "dilated depthwise, strip kernel")


@pytest.mark.parametrize("stride,pad,dil", [(1, 1, 1), (2, 1, 1), (1, 0, 1), (2, 0, 1), (1, 0, 2)])
def test_depthwise_row_segments_do_not_change_the_bytes(oracle, stride, pad, dil, monkeypatch):
    """dfq_dw_q8 cuts an output row into segments of whole strips at small batches (more threads); one thread per row
    (DFQ_Q8_DW_SEGS=1) and five segments must write the same bytes."""
    import ctypes
    from data_free_quantization_b200 import _lib as L
    from data_free_quantization_b200.eval_q8 import pack_depthwise, to_codes_nhwc
    from test_gpu_q8 import slot
    lib = L.lib()
    rng = np.random.default_rng(stride * 10 + pad + dil)
    N, C, H, W = 2, 40, 23, 61
    Cp = 48
    sl, _ = slot(-1.5, 3.0)
    (s0, _), (s1, _) = slot(-2.0, 2.5), slot(0.0, 1.5)
    x = to_codes_nhwc(torch.from_numpy(rng.integers(0, 256, (N, C, H, W)).astype(np.uint8))).reshape(-1).to(DEV)
    w0 = (rng.standard_normal((C, 1, 3, 3)) * 0.3).astype(np.float32)
    w, _ = oracle.fake_quant(w0, 8, float(w0.min()), float(w0.max()), want_codes=True)
    wrow, dsw, dcw = pack_depthwise(torch.from_numpy(w).to(DEV))
    bp = torch.zeros(Cp, device=DEV); bp[:C] = torch.from_numpy((rng.standard_normal(C) * 0.2).astype(np.float32)).to(DEV)
    OH, OW = (H + 2 * pad - 2 * dil - 1) // stride + 1, (W + 2 * pad - 2 * dil - 1) // stride + 1
    res = []
    for segs in ("1", "5"):
        monkeypatch.setenv("DFQ_Q8_DW_SEGS", segs)
        y0 = torch.full((N * OH * OW * Cp,), 77, dtype=torch.int8, device=DEV)
        y1 = torch.full((N * OH * OW * Cp,), 77, dtype=torch.int8, device=DEV)
        d = L.DfqDwQ8Desc()
        d.x, d.wrow, d.bias, d.in_slot, d.y0, d.out_slot0 = x.data_ptr(), wrow.data_ptr(), bp.data_ptr(), sl.data_ptr(), y0.data_ptr(), s0.data_ptr()
        d.y1, d.out_slot1 = y1.data_ptr(), s1.data_ptr()
        d.N, d.H, d.W, d.C, d.stride, d.relu, d.s_w, d.c_w, d.pad, d.dil = N, H, W, C, stride, 0, dsw, dcw, pad, dil
        L.check(lib.dfq_dw_q8(ctypes.byref(d), L.stream()), "dfq_dw_q8")
        torch.cuda.synchronize()
        res.append((y0, y1))
    for a, b in zip(*res):
        assert torch.equal(a, b) and not bool((a == 77).all())


def test_forked_sibling_convolutions_give_the_same_output(static_deeplab, monkeypatch):
    """The ASPP branches launched side by side on forked streams (EvalPlanQ8._schedule) against the one-stream order:
    same kernels, so the same logits -- eagerly and replayed from a CUDA graph, several times over."""
    from data_free_quantization_b200.eval_q8 import EvalPlanQ8
    model, tr = static_deeplab
    x = torch.randn(2, 3, 129, 129, generator=torch.Generator().manual_seed(4)).to(DEV)
    monkeypatch.setenv("DFQ_Q8_FORK", "0")
    plain = EvalPlanQ8(model, tr, batch=2, in_shape=(3, 129, 129))
    want = plain(x).clone()
    monkeypatch.setenv("DFQ_Q8_FORK", "1")
    plan = EvalPlanQ8(model, tr, batch=2, in_shape=(3, 129, 129))
    assert not plain.forked and plan.forked and max(plan.forked) >= 3
    for _ in range(3):
        assert torch.equal(plan(x), want)
    replay = plan.capture()
    for _ in range(5):
        plan.input.buf.copy_(x)
        replay()
        torch.cuda.synchronize()
        assert torch.equal(plan.output.buf, want)
