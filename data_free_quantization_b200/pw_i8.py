"""Host side of the tensor-core pointwise / linear kernel (csrc/pw_i8.cu, dfq_pw_i8): weight-code recovery and
packing at plan time, and the launch wrapper.  The arithmetic it replaces is F.conv2d / F.linear on fake-quantised
tensors in QuantNConv2d.forward / QuantNLinear.forward (utils/quantize.py:250-256, 358-363)."""
import ctypes

import numpy as np
import torch

from . import _lib as L


class PackedWeight:
    """Weights of one layer on an integer grid: w[o,k] = s_w * code[o,k] + min_w, code in [0, 2^bits - 1]."""

    def __init__(self, wpk, wr, s_w, c_w, O, K, codes, o_cap=0):
        self.wpk, self.wr, self.s_w, self.c_w, self.O, self.K, self.codes = wpk, wr, s_w, c_w, O, K, codes
        self.o_cap = int(o_cap)        # output channels per pass the layout was packed for (0 = 256)


def choose_o_cap(G, O, sms=148, fill=1.0):
    """Output channels per pass (dfq_pw_i8_plan_cap_host) for a layer of G pixels x O channels.  The default, 256,
    gives ceil(G / 128) * ceil(O / 256) CTAs; when that leaves SMs idle the passes are narrowed until about two CTAs
    per SM exist.  Every pass quantises its own copy of the A tile, so full grids keep the default; `fill` is the
    fraction of the SMs the default grid must cover to be left alone."""
    tiles = (int(G) + 127) // 128
    if tiles * ((O + 255) // 256) >= fill * sms:
        return 0
    n_pass = min((2 * sms + tiles - 1) // tiles, (O + 15) // 16)
    cap = ((O + n_pass - 1) // n_pass + 15) // 16 * 16
    return 0 if cap >= 256 else max(16, cap)


def pack_weight(w, bits=None, o_cap=0):
    """pack_weight_bits for the layer's weight width, or -- QuantN* layers do not record one -- for the widest grid
    of <= 8 bits the tensor sits on."""
    for b in ([bits] if bits else (8, 7, 6, 5, 4, 3, 2)):
        pk = pack_weight_bits(w, int(b), o_cap)
        if pk is not None:
            return pk
    return None


def pack_weight_bits(w, bits=8, o_cap=0):
    """Recover the integer codes of an already fake-quantised weight (what quantize_targ_layer stored,
    utils/layer_transform.py:288-305, or what QuantConv2d derives per forward, utils/quantize.py:227) and lay them
    out for the kernel.  Returns None when the tensor is not on a `bits`-bit grid (then the fp32 kernel applies)."""
    if not (1 <= bits <= 8) or not w.is_cuda:
        return None
    O = w.shape[0]
    w2 = w.detach().reshape(O, -1)
    K = w2.shape[1]
    lo, hi = float(w2.min()), float(w2.max())
    levels = float(2 ** bits - 1)
    s = max((hi - lo) / levels, 1e-8)
    s32 = float(np.float32(s))
    q = torch.round((w2.double() - lo) / s32).clamp_(0, levels)
    back = q * s32 + lo
    tol = 4 * 1.2e-7 * max(abs(lo), abs(hi)) + 1e-7 * (hi - lo)
    if float((back - w2.double()).abs().max()) > tol:
        return None
    # centred and scaled to the s8 range: b = (code - 2^(bits-1)) * 2^(8-bits), w = (s / 2^(8-bits)) * b + c_w
    up = 2 ** (8 - bits)
    codes = ((q - 2 ** (bits - 1)) * up).to(torch.int8).cpu().numpy()
    n_pass, opp, npass, slabs, nbytes = (ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), L.i64())
    lib = L.lib()
    L.check(lib.dfq_pw_i8_plan_cap_host(O, K, int(o_cap), ctypes.byref(n_pass), ctypes.byref(opp), ctypes.byref(npass),
                                        ctypes.byref(slabs), ctypes.byref(nbytes)), "dfq_pw_i8_plan_cap_host")
    packed = np.empty(nbytes.value, dtype=np.int8)
    codes = np.ascontiguousarray(codes)
    L.check(lib.dfq_pw_i8_pack_cap_host(codes.ctypes.data, O, K, int(o_cap), packed.ctypes.data),
            "dfq_pw_i8_pack_cap_host")
    s_eff = s32 / up
    wr = (codes.astype(np.int64).sum(1).astype(np.float64) * s_eff).astype(np.float32)
    dev = w.device
    pk = PackedWeight(torch.from_numpy(packed).to(dev), torch.from_numpy(wr).to(dev), s_eff,
                      float(np.float32(lo + 2 ** (bits - 1) * s32)), O, K, codes, o_cap)
    pk.grid = (s32, lo, q.to(torch.int16).cpu().numpy())       # w = s * code + lo, code in [0, 2^bits - 1]
    return pk


def make_desc(x_ptr, y_ptr, pk, bias, slot_ptr, stats_ptr, images, P, relu, bits, linear=False):
    """DfqPwI8Desc for a conv ([images, K, P] -> [images, O, P]) or a linear layer ([images, K] -> [images, O])."""
    d = L.DfqPwI8Desc()
    d.x, d.y, d.wpk, d.wr = x_ptr, y_ptr, pk.wpk.data_ptr(), pk.wr.data_ptr()
    d.bias = None if bias is None else L.dev_f32(bias, "bias")
    d.in_slot, d.out_stats = slot_ptr, stats_ptr
    d.K, d.O, d.bits, d.relu = pk.K, pk.O, int(bits), int(relu)
    d.s_w, d.c_w = pk.s_w, pk.c_w
    d.o_cap = pk.o_cap
    if linear:
        d.G, d.P = images, 1
        d.xs_n, d.xs_k, d.xs_p, d.ys_n, d.ys_o, d.ys_p = pk.K, 1, 0, pk.O, 1, 0
    else:
        d.G, d.P = images * P, P
        d.xs_n, d.xs_k, d.xs_p, d.ys_n, d.ys_o, d.ys_p = pk.K * P, P, 1, pk.O * P, P, 1
    return d


def launch(desc):
    L.check(L.lib().dfq_pw_i8(ctypes.byref(desc), L.stream()), "dfq_pw_i8")


# ---- dense k x k convolutions (csrc/pw_i8.cu, kConv; dfq_conv_i8) --------------------------------------------------
def pack_weight_conv(w, bits=None, o_cap=0):
    """pack_weight for a dense [O, C, kh, kw] weight (groups == 1): the reduction runs tap-major with the channels
    padded to a multiple of 16 (dfq_conv_i8_k_host), plus the per-tap code sums the padding correction needs.
    Returns None when the tensor is not on a <= 8-bit grid."""
    O, C, kh, kw = w.shape
    taps, Cp = kh * kw, (C + 15) // 16 * 16
    pk = None
    for b in ([bits] if bits else (8, 7, 6, 5, 4, 3, 2)):
        pk = _recover_codes(w, int(b))
        if pk is not None:
            break
    if pk is None:
        return None
    s32, lo, q, b = pk
    up = 2 ** (8 - b)
    codes = ((q - 2 ** (b - 1)) * up).to(torch.int8).cpu().numpy().reshape(O, C, taps)
    lay = np.zeros((O, taps, Cp), dtype=np.int8)
    lay[:, :, :C] = codes.transpose(0, 2, 1)
    K = taps * Cp
    lib = L.lib()
    assert K == lib.dfq_conv_i8_k_host(C, kh, kw)
    n_pass, opp, npass, slabs, nbytes = (ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), L.i64())
    L.check(lib.dfq_pw_i8_plan_cap_host(O, K, int(o_cap), ctypes.byref(n_pass), ctypes.byref(opp), ctypes.byref(npass),
                                        ctypes.byref(slabs), ctypes.byref(nbytes)), "dfq_pw_i8_plan_cap_host")
    packed = np.empty(nbytes.value, dtype=np.int8)
    lay = np.ascontiguousarray(lay.reshape(O, K))
    L.check(lib.dfq_pw_i8_pack_cap_host(lay.ctypes.data, O, K, int(o_cap), packed.ctypes.data),
            "dfq_pw_i8_pack_cap_host")
    s_eff = s32 / up
    tap_sums = codes.astype(np.int64).sum(1)                                   # [O, taps]
    wr = (tap_sums.sum(1).astype(np.float64) * s_eff).astype(np.float32)
    wrt = np.ascontiguousarray((tap_sums.T.astype(np.float64) * s_eff).astype(np.float32))   # [taps, O]
    dev = w.device
    out = PackedWeight(torch.from_numpy(packed).to(dev), torch.from_numpy(wr).to(dev), s_eff,
                       float(np.float32(lo + 2 ** (b - 1) * s32)), O, K, codes, o_cap)
    out.wrt = torch.from_numpy(wrt).to(dev)
    out.geom = (C, kh, kw)
    out.grid = (s32, lo, q.to(torch.int16).cpu().numpy().reshape(O, C, kh, kw))
    return out


def _recover_codes(w, bits):
    """(scale, min, codes as a double tensor [O, K], bits) of an on-grid weight, else None (pack_weight_bits)."""
    if not (1 <= bits <= 8) or not w.is_cuda:
        return None
    w2 = w.detach().reshape(w.shape[0], -1)
    lo, hi = float(w2.min()), float(w2.max())
    levels = float(2 ** bits - 1)
    s32 = float(np.float32(max((hi - lo) / levels, 1e-8)))
    q = torch.round((w2.double() - lo) / s32).clamp_(0, levels)
    tol = 4 * 1.2e-7 * max(abs(lo), abs(hi)) + 1e-7 * (hi - lo)
    if float((q * s32 + lo - w2.double()).abs().max()) > tol:
        return None
    return s32, lo, q, bits


def conv_scratch(x_shape, device):
    """Device buffer for the quantise-once form of dfq_conv_i8 (the input as s8 codes, NHWC, C padded to 16)."""
    N, C, H, W = x_shape
    return torch.empty(int(L.lib().dfq_conv_i8_scratch_bytes(N, C, H, W)), dtype=torch.int8, device=device)


def make_conv_desc(x, y, pk, bias, slot_ptr, stats_ptr, stride, pad, dil, relu, bits, scratch=None, halo=False):
    """DfqConvI8Desc for x [N, C, H, W] -> y [N, O, OH, OW] (both contiguous fp32 CUDA tensors).  With `scratch`
    (conv_scratch) the input is quantised once and the taps read 16-byte code words -- the form for k x k > 1."""
    C, kh, kw = pk.geom
    d = L.DfqConvI8Desc()
    d.x, d.y = L.dev_f32(x, "input"), L.dev_f32(y, "output")
    d.wpk, d.wr, d.wrt = pk.wpk.data_ptr(), pk.wr.data_ptr(), pk.wrt.data_ptr()
    d.bias = None if bias is None else L.dev_f32(bias, "bias")
    d.in_slot, d.out_stats = slot_ptr, stats_ptr
    d.N, d.C, d.H, d.W = x.shape
    d.O, d.OH, d.OW = y.shape[1], y.shape[2], y.shape[3]
    d.kh, d.kw, d.stride, d.pad, d.dil = kh, kw, int(stride), int(pad), int(dil)
    d.bits, d.relu = int(bits), int(relu)
    d.s_w, d.c_w = pk.s_w, pk.c_w
    d.scratch = None if scratch is None else scratch.data_ptr()
    d.o_cap = pk.o_cap
    d.halo = 1 if halo else 0      # pk packed by eval_q8.pack_conv_halo
    return d


def launch_conv(desc):
    L.check(L.lib().dfq_conv_i8(ctypes.byref(desc), L.stream()), "dfq_conv_i8")
