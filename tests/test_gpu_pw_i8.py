"""GPU: the tcgen05 s8 x s8 -> s32 pointwise / linear kernel (csrc/pw_i8.cu, dfq_pw_i8) against (a) an exact integer
restatement -- codes from the oracle's fake-quant recipe (utils/quantize.py:72-76), the K-long sums in int64, the
affine terms in float64 -- and (b) the reference's own formulation, F.conv2d / F.linear in fp32 on the
fake-quantised tensors (utils/quantize.py:250-256, 358-363).  Tolerances: 2e-6 of the largest output against (a)
(one fp32 FMA chain per output), 2e-5 against (b) (the reference accumulates K terms in fp32)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from test_gpu_eval_kernels import _decode_stats, _slot

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _case(oracle, images, K, O, H, W, linear=False, relu=True, bias=True, wbits=8, abits=8, seed=0, o_cap=0):
    from data_free_quantization_b200 import pw_i8
    rs = np.random.default_rng(seed)
    P = 1 if linear else H * W
    x = (rs.standard_normal((images, K) if linear else (images, K, H, W)) * 1.5 + 0.3).astype(np.float32)
    lo, hi = float(np.float32(-2.1)), float(np.float32(3.4))         # clamps both tails
    w0 = (rs.standard_normal((O, K)) * 0.2).astype(np.float32)
    w = oracle.fake_quant(w0, wbits, float(w0.min()), float(w0.max()))    # what quantize_targ_layer stores
    b = (rs.standard_normal(O) * 0.5).astype(np.float32) if bias else None
    xq = oracle.fake_quant(x, abits, lo, hi)
    # (b) the reference's formulation
    if linear:
        want32 = F.linear(torch.from_numpy(xq), torch.from_numpy(w), None if b is None else torch.from_numpy(b))
    else:
        want32 = F.conv2d(torch.from_numpy(xq), torch.from_numpy(w).view(O, K, 1, 1),
                          None if b is None else torch.from_numpy(b))
    # (a) exact: integer sums, affine terms in float64
    slot = _slot(lo, hi, abits)
    sx = float(slot[2])
    qa = np.rint((xq.astype(np.float64) - lo) / sx).astype(np.int64)
    wd = torch.from_numpy(w).to(DEV)
    pk = pw_i8.pack_weight(wd, wbits, o_cap=o_cap)
    assert pk is not None, "quantised weight not recognised as on-grid"
    sw, mw, qb = pk.grid
    qb = qb.astype(np.int64)
    A = qa if linear else qa.transpose(0, 2, 3, 1).reshape(-1, K)
    acc = A @ qb.T
    exact = sx * sw * acc + sx * mw * A.sum(1, keepdims=True) + lo * sw * qb.sum(1)[None, :] + K * lo * mw
    if b is not None:
        exact = exact + b[None, :].astype(np.float64)
    exact = exact if linear else exact.reshape(images, H, W, O).transpose(0, 3, 1, 2)
    if relu:
        want32, exact = torch.relu(want32), np.maximum(exact, 0)
    xd = torch.from_numpy(x).to(DEV)
    y = torch.full(tuple(want32.shape), float("nan"), device=DEV)
    stats = torch.zeros(images, 2, dtype=torch.int32, device=DEV)
    bd = None if b is None else torch.from_numpy(b).to(DEV)
    d = pw_i8.make_desc(xd.data_ptr(), y.data_ptr(), pk, bd, slot.data_ptr(), stats.data_ptr(), images, P, relu,
                        abits, linear=linear)
    pw_i8.launch(d)
    torch.cuda.synchronize()
    got = y.cpu().numpy()
    assert not np.isnan(got).any()
    scale = np.abs(exact).max()
    err_a = np.abs(got - exact).max() / scale
    err_b = np.abs(got - want32.numpy()).max() / scale
    assert err_a <= 2e-6, ("exact", err_a)
    assert err_b <= 2e-5, ("fp32 reference", err_b)
    mn, mx = _decode_stats(stats)
    g = got.reshape(images, -1)
    assert np.array_equal(mn, g.min(1)) and np.array_equal(mx, g.max(1))


@pytest.mark.parametrize("images,K,O,H,W", [
    (3, 16, 96, 7, 7),          # one K chunk, ragged pixel tile crossing image boundaries
    (2, 24, 144, 56, 56),       # K not a multiple of 16
    (2, 32, 16, 28, 28),        # smallest N
    (1, 96, 24, 56, 56),        # two slabs
    (2, 144, 32, 14, 14),       # three slabs, last one ragged
    (3, 960, 320, 7, 7),        # 15 slabs, two passes of 160
    (2, 320, 1280, 7, 7),       # five passes of 256
    (1, 64, 384, 14, 14),       # two passes of 192
])
def test_pointwise_conv(oracle, images, K, O, H, W):
    _case(oracle, images, K, O, H, W, seed=K + O)


def test_pointwise_no_relu_no_bias(oracle):
    _case(oracle, 2, 192, 64, 14, 14, relu=False, bias=False)


def test_linear_classifier(oracle):
    _case(oracle, 5, 1280, 1000, 1, 1, linear=True, relu=False)


@pytest.mark.parametrize("o_cap", [16, 48, 128, 250])
def test_narrow_passes(oracle, o_cap):
    """dfq_pw_i8_plan_cap_host: the same layer packed with fewer output channels per pass (more CTAs for layers whose
    pixel tiles do not fill the SMs) -- the classifier (63 passes of 16 at o_cap = 16) and a 7 x 7 projection."""
    _case(oracle, 5, 1280, 1000, 1, 1, linear=True, relu=False, o_cap=o_cap)
    _case(oracle, 3, 960, 160, 7, 7, seed=o_cap, o_cap=o_cap)


def test_choose_o_cap():
    from data_free_quantization_b200.pw_i8 import choose_o_cap
    assert choose_o_cap(256 * 112 * 112, 96) == 0            # full grids keep 256 channels per pass
    assert choose_o_cap(256, 1000) == 16                     # classifier at batch 256: 2 tiles -> 63 passes
    assert choose_o_cap(256 * 49, 160) in (48, 64)           # 98 tiles -> 3-4 passes
    assert choose_o_cap(16 * 33 * 33, 256) == 96             # ASPP at 16 x 33 x 33: 137 tiles -> 3 passes


def test_linear_many_samples(oracle):
    _case(oracle, 300, 128, 40, 1, 1, linear=True, relu=False, seed=5)


@pytest.mark.parametrize("K", [130, 132, 200, 1280])
def test_linear_row_loads(oracle, K):
    """K-contiguous rows take four 128-bit loads per 16-float chunk when K % 4 == 0 (ragged last chunk scalar:
    K = 132, 200), the scalar form otherwise (K = 130)."""
    _case(oracle, 257, K, 48, 1, 1, linear=True, relu=False, seed=K)


@pytest.mark.parametrize("wbits,abits", [(4, 8), (6, 8), (8, 4)])
def test_low_bit_grids(oracle, wbits, abits):
    _case(oracle, 2, 64, 48, 14, 14, wbits=wbits, abits=abits, seed=wbits * 10 + abits)


def test_off_grid_weight_is_refused():
    from data_free_quantization_b200 import pw_i8
    w = torch.randn(32, 64, device=DEV)
    assert pw_i8.pack_weight(w, 8) is None


def test_role_split_persistent_variant():
    """The opt-in persistent kernel (pw_i8_ws_kernel: two worker groups, MMA issuer and weight streamer meeting at
    mbarriers, double-buffered TMEM accumulator) passes the same cases; the switch is read once per process, hence
    the subprocess."""
    import os
    import subprocess
    import sys
    if os.environ.get("DFQ_PW_WS") == "force":
        pytest.skip("already inside the forced run")
    env = dict(os.environ, DFQ_PW_WS="force")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-q", "-m", "gpu", "-x", "-k",
                        "not role_split"], env=env, capture_output=True, text=True, timeout=600,
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
