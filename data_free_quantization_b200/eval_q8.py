"""Static-range W8A8 eval plan: the quantised forward with every QuantMeasure in `update_stat == False` mode
(utils/quantize.py:119-126, set_update_stat), compiled into launches of csrc/q8.cu.

With static ranges a consumer's grid is known before its producer runs, so activations never exist as fp32 in HBM:
every producer requantises in its epilogue to the grid of each consumer's QuantMeasure (a tensor read by a
convolution AND a residual add is written twice, once per grid -- the reference quantises it twice, with the two
measures' ranges) and stores centred s8 codes, NHWC, channels padded to 16.  The 1 x 1 convolutions and the
classifier read those codes by tensor-map TMA straight into the tcgen05 operand layout.  13.9 MB per image of
activation traffic for MobileNetV2 instead of 55.8 MB with fp32 activations (eval_plan.py).

Supported graph: what MobileNetV2 traces to -- a dense 3 x 3 stem on the fp32 image, 1 x 1 convolutions, depthwise
3 x 3 (pad 1, stride 1 / 2), residual adds, the spatial mean, the linear classifier; ReLU fused, BatchNorm already
merged.  Anything else (DeepLab's pad / interpolate / cat / atrous convolutions) raises NotImplementedError: those run
on the dynamic-range plan.

Semantics per op are those of the reference with replace_op() in force (utils/layer_transform.py:18-125,
utils/quantize.py:250-256, 358-363); bits_activation is 8 (the reference's set_layer_bits never changes it).
"""
import ctypes

import numpy as np
import torch

from . import _lib as L
from . import pw_i8
from .eval_plan import EvalPlan


def _cp(c):
    return (int(c) + 15) // 16 * 16


class PackedQ8:
    """Weight codes of a 1 x 1 / linear layer in the layout dfq_pw_q8 streams, with a row of ones per pass."""

    def __init__(self, w, o_cap=0):
        rec = None
        for b in (8, 7, 6, 5, 4, 3, 2):
            rec = pw_i8._recover_codes(w, b)
            if rec is not None:
                break
        if rec is None:
            raise RuntimeError("q8 plan: weight is not on an integer grid of <= 8 bits (run quantize_targ_layer first)")
        s32, lo, q, bits = rec
        up = 2 ** (8 - bits)
        codes = np.ascontiguousarray(((q - 2 ** (bits - 1)) * up).to(torch.int8).cpu().numpy())
        self.O, self.K = codes.shape
        self.o_cap = int(o_cap)
        lib = L.lib()
        n_pass, opp, npass, slabs, nbytes = (ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), L.i64())
        L.check(lib.dfq_pw_q8_plan_host(self.O, self.K, self.o_cap, ctypes.byref(n_pass), ctypes.byref(opp),
                                        ctypes.byref(npass), ctypes.byref(slabs), ctypes.byref(nbytes)),
                "dfq_pw_q8_plan_host")
        packed = np.empty(nbytes.value, dtype=np.int8)
        L.check(lib.dfq_pw_q8_pack_host(codes.ctypes.data, self.O, self.K, self.o_cap, packed.ctypes.data),
                "dfq_pw_q8_pack_host")
        self.s_w = s32 / up
        self.c_w = float(np.float32(lo + 2 ** (bits - 1) * s32))
        wr = (codes.astype(np.int64).sum(1).astype(np.float64) * self.s_w).astype(np.float32)
        dev = w.device
        self.wpk, self.wr = torch.from_numpy(packed).to(dev), torch.from_numpy(wr).to(dev)
        self.codes = codes
        self.n_pass, self.o_per_pass = n_pass.value, opp.value


def pack_depthwise(w):
    """[C, 1, 3, 3] fake-quantised depthwise weight -> ([3][Cp] words (w(r,0), w(r,1), w(r,2), 0) of centred s8 codes,
    s_w, c_w) for dfq_dw_q8."""
    rec = None
    for b in (8, 7, 6, 5, 4, 3, 2):
        rec = pw_i8._recover_codes(w, b)
        if rec is not None:
            break
    if rec is None:
        raise RuntimeError("q8 plan: depthwise weight is not on an integer grid of <= 8 bits")
    s32, lo, q, bits = rec
    up = 2 ** (8 - bits)
    C = w.shape[0]
    codes = ((q - 2 ** (bits - 1)) * up).to(torch.int8).cpu().numpy().reshape(C, 3, 3).astype(np.uint8).astype(np.uint32)
    rows = np.zeros((3, _cp(C)), np.uint32)
    rows[:, :C] = (codes[:, :, 0] | (codes[:, :, 1] << 8) | (codes[:, :, 2] << 16)).T
    return torch.from_numpy(rows.view(np.int32)).to(w.device), s32 / up, float(np.float32(lo + 2 ** (bits - 1) * s32))


def choose_o_cap(G, O, sms=148):
    """Output channels per pass.  112 (+ the row of ones = 128 accumulator columns): two accumulators then take 256 of
    the 512 TMEM columns, so TWO CTAs -- 16 epilogue warps -- share an SM; the A tile a second pass re-reads is small
    next to the codes it writes.  Narrower when that still leaves most SMs without a tile (the classifier)."""
    tiles = (int(G) + 127) // 128
    if tiles * ((O + 111) // 112) >= sms:
        return 112
    want = max(1, (2 * sms) // tiles)
    return max(16, min(112, _cp((O + want - 1) // want)))


def to_codes_nhwc(x_codes_nchw):
    """uint8 codes [N, C, H, W] (or [N, C]) -> centred s8, NHWC, channels zero-padded to 16 (the q8 activation format)."""
    a = torch.as_tensor(x_codes_nchw)
    if a.dim() == 2:
        a = a[:, :, None, None]
    elif a.dim() == 3:
        a = a[..., None]
    N, C, H, W = a.shape
    out = torch.zeros(N, H, W, _cp(C), dtype=torch.int8)
    out[..., :C] = (a.permute(0, 2, 3, 1).to(torch.int16) - 128).to(torch.int8)
    return out


def from_codes_nhwc(buf, shape):
    """centred s8 NHWC buffer -> uint8 codes in the tensor's logical shape ([N, C, H, W] or [N, C])."""
    N, C = shape[0], shape[1]
    H, W = (shape[2], shape[3]) if len(shape) == 4 else (1, 1)
    a = buf.view(N, H, W, _cp(C))[..., :C].to(torch.int16) + 128
    a = a.permute(0, 3, 1, 2).to(torch.uint8)
    return a.reshape(shape)


class EvalPlanQ8(EvalPlan):
    def __init__(self, model, transformer, batch, in_shape=(3, 224, 224)):
        super().__init__(model, transformer, batch, in_shape)

    # ------------------------------------------------------------------------------------- buffers + launches
    def _assign(self, steps):
        dev, N = self.dev, self.N
        lib, stream = L.lib(), L.stream
        for qm, _idx, _t in self.measures:
            if qm.update_stat:
                raise RuntimeError("q8 plan: every QuantMeasure must be in static mode (set_update_stat(False), "
                                   "utils/quantize.py:125-126); the dynamic-range plan is eval_plan.EvalPlan")
            if int(qm.num_bits) != 8:
                raise NotImplementedError("q8 plan: 8-bit activations only")
        self.refresh_static()
        cons = {}
        for _qm, idx, t in self.measures:
            cons.setdefault(id(t), []).append(idx)
        slot_ptr = lambda i: self.slots[i].data_ptr()
        # code buffers, one per (tensor, consumer grid); reused once their single reader has run
        ins_of = lambda st: [st[k] for k in ("src", "a", "b") if k in st]
        slot_of_in = lambda st: [st[k] for k in ("slot", "slot_a", "slot_b") if st.get(k) is not None]
        last_read = {}
        for i, st in enumerate(steps):
            for idx in slot_of_in(st):
                last_read[idx] = i
        pool, self.codes, self.slot_key = {}, {}, {}
        self.act_bytes = 0

        def code_shape(t):
            return (t.shape[0], t.shape[2], t.shape[3], _cp(t.shape[1])) if len(t.shape) == 4 else (t.shape[0], _cp(t.shape[1]))

        def alloc(t, idx):
            shp = code_shape(t)
            n = int(np.prod(shp))
            free = pool.get(n)
            buf = free.pop() if free else torch.zeros(n, dtype=torch.int8, device=dev)
            self.codes[idx] = buf
            self.act_bytes += n
            return buf

        def outs_of(t, name):
            idxs = cons.get(id(t), [])
            if len(idxs) > 2:
                raise NotImplementedError(f"{name}: {len(idxs)} quantised consumers (the q8 kernels write two grids)")
            return idxs

        self.input.buf = torch.empty(*self.input.shape, dtype=torch.float32, device=dev)
        self.keep_q8 = []
        self.step_io = {}
        self.out_of = {}
        top_i = 0
        for i, st in enumerate(steps):
            kind, name, out = st["kind"], st["name"], st["out"]
            oidx = outs_of(out, name)
            obufs = [alloc(out, idx) for idx in oidx]
            y0 = obufs[0].data_ptr() if obufs else None
            y1 = obufs[1].data_ptr() if len(obufs) > 1 else None
            s0 = slot_ptr(oidx[0]) if oidx else None
            s1 = slot_ptr(oidx[1]) if len(oidx) > 1 else None
            if kind == "conv":
                m, src = st["m"], st["src"]
                kh, kw, s, p, d, g = st["geom"]
                if st["slot"] is None:
                    raise NotImplementedError(f"{name}: convolution without a QuantMeasure")
                self.slot_key[st["slot"]] = ("L", name)
                w = m.weight.data
                b = None if m.bias is None else m.bias.data
                Nn, C, H, W = src.shape
                if src is self.input:
                    if not (kh == 3 and kw == 3 and g == 1 and C == 3 and d == 1 and m.out_channels <= 32 and oidx):
                        raise NotImplementedError(f"{name}: only a dense 3 x 3 stem on 3 channels reads the image")
                    desc = L.DfqStemQ8Desc()
                    desc.x, desc.w = src.buf.data_ptr(), L.dev_f32(w.contiguous(), "weight")
                    desc.bias = None if b is None else L.dev_f32(b, "bias")
                    desc.in_slot, desc.y0, desc.out_slot0, desc.y1, desc.out_slot1 = slot_ptr(st["slot"]), y0, s0, y1, s1
                    desc.N, desc.H, desc.W, desc.O, desc.stride, desc.pad, desc.relu = Nn, H, W, m.out_channels, s, p, st["relu"]
                    self.keep_q8.append(desc)
                    fn = (lambda dd: lambda: L.check(lib.dfq_stem_q8(ctypes.byref(dd), stream()), "dfq_stem_q8"))(desc)
                    kind2 = "stem"
                elif kh == 1 and kw == 1 and g == 1 and s == 1 and p == 0:
                    fn = self._pw(name, w.reshape(w.shape[0], -1), b, self.codes[st["slot"]], slot_ptr(st["slot"]),
                                  Nn * H * W, y0, s0, y1, s1, st["relu"], None)
                    kind2 = "pw"
                elif kh == 3 and kw == 3 and g == C == m.out_channels and p == 1 and d == 1 and s in (1, 2) and oidx:
                    Cp = _cp(C)
                    wrow, dsw, dcw = pack_depthwise(w)
                    bp = torch.zeros(Cp, dtype=torch.float32, device=dev)
                    if b is not None:
                        bp[:C] = b
                    desc = L.DfqDwQ8Desc()
                    desc.x, desc.wrow, desc.bias = self.codes[st["slot"]].data_ptr(), wrow.data_ptr(), bp.data_ptr()
                    desc.s_w, desc.c_w = dsw, dcw
                    desc.in_slot, desc.y0, desc.out_slot0, desc.y1, desc.out_slot1 = slot_ptr(st["slot"]), y0, s0, y1, s1
                    desc.N, desc.H, desc.W, desc.C, desc.stride, desc.relu = Nn, H, W, C, s, st["relu"]
                    self.keep_q8 += [wrow, bp, desc]
                    fn = (lambda dd: lambda: L.check(lib.dfq_dw_q8(ctypes.byref(dd), stream()), "dfq_dw_q8"))(desc)
                    kind2 = "dw"
                else:
                    raise NotImplementedError(f"{name}: convolution geometry {st['geom']} is not covered by the q8 plan")
            elif kind == "linear":
                m, src = st["m"], st["src"]
                if st["slot"] is None:
                    raise NotImplementedError(f"{name}: linear layer without a QuantMeasure")
                self.slot_key[st["slot"]] = ("L", name)
                yf = None
                if out is self.output or not oidx:
                    out.buf = torch.empty(*out.shape, dtype=torch.float32, device=dev)
                    yf = out.buf
                fn = self._pw(name, m.weight.data, None if m.bias is None else m.bias.data, self.codes[st["slot"]],
                              slot_ptr(st["slot"]), src.shape[0], y0, s0, y1, s1, st["relu"], yf)
                kind2 = "linear"
            elif kind == "add":
                a_, b_ = st["a"], st["b"]
                self.slot_key[st["slot_a"]] = ("T", top_i, 0)
                self.slot_key[st["slot_b"]] = ("T", top_i, 1)
                top_i += 1
                if not oidx:
                    raise NotImplementedError(f"{name}: residual add without a quantised consumer")
                Nn, C, H, W = out.shape
                args = (self.codes[st["slot_a"]].data_ptr(), self.codes[st["slot_b"]].data_ptr(), slot_ptr(st["slot_a"]),
                        slot_ptr(st["slot_b"]), y0, s0, y1, s1, Nn * H * W, C)
                fn = (lambda a2: lambda: L.check(lib.dfq_add_q8(*a2, stream()), "dfq_add_q8"))(args)
                kind2 = "add"
            elif kind == "mean":
                src = st["src"]
                self.slot_key[st["slot"]] = ("T", top_i, 0)
                top_i += 1
                Nn, C, H, W = src.shape
                yf = None
                if not oidx:
                    out.buf = torch.empty(*out.shape, dtype=torch.float32, device=dev)
                    yf = out.buf.data_ptr()
                args = (self.codes[st["slot"]].data_ptr(), slot_ptr(st["slot"]), y0, s0, yf, Nn, C, H * W)
                fn = (lambda a2: lambda: L.check(lib.dfq_mean_q8(*a2, stream()), "dfq_mean_q8"))(args)
                kind2 = "mean"
            else:
                raise NotImplementedError(f"{name}: `{kind}` is not covered by the q8 plan (use eval_plan.EvalPlan)")
            self.ops.append((name, fn))
            self.out_of[name] = out
            self.step_io[name] = {"kind": kind2, "in_slots": slot_of_in(st), "srcs": ins_of(st), "out": out,
                                  "out_slots": oidx, "relu": st.get("relu", 0), "fn": fn}
            for idx in slot_of_in(st):                      # the codes this step read are dead now
                if last_read[idx] == i and idx in self.codes:
                    buf = self.codes[idx]
                    pool.setdefault(buf.numel(), []).append(buf)
        self.op_info = {n: {"kind": io["kind"]} for n, io in self.step_io.items()}
        self.hbm_bytes = self.act_bytes

    def _pw(self, name, w2d, bias, xbuf, in_slot, G, y0, s0, y1, s1, relu, yf):
        pk = PackedQ8(w2d, o_cap=choose_o_cap(G, w2d.shape[0]))
        d = L.DfqPwQ8Desc()
        d.x, d.wpk, d.wr = xbuf.data_ptr(), pk.wpk.data_ptr(), pk.wr.data_ptr()
        d.bias = None if bias is None else L.dev_f32(bias, "bias")
        d.in_slot, d.y0, d.out_slot0, d.y1, d.out_slot1 = in_slot, y0, s0, y1, s1
        if yf is not None:
            d.yf, d.yf_g, d.yf_o = yf.data_ptr(), pk.O, 1
        d.G, d.K, d.O, d.relu, d.o_cap, d.s_w, d.c_w = int(G), pk.K, pk.O, int(relu), pk.o_cap, pk.s_w, pk.c_w
        self.keep_q8 += [pk, d, bias]
        lib, stream = L.lib(), L.stream
        return lambda: L.check(lib.dfq_pw_q8(ctypes.byref(d), stream()), "dfq_pw_q8")

    # ------------------------------------------------------------------------------------------------- run
    def refresh_static(self):
        """(min, max, scale) of every measure from its running range (one device->host copy)."""
        if not self.measures:
            self.slots = torch.zeros(1, 4, dtype=torch.float32, device=self.dev)
            return
        rng = torch.stack([torch.cat([qm.running_min.detach().reshape(1), qm.running_max.detach().reshape(1)])
                           for qm, _i, _t in self.measures]).cpu().numpy()
        host = np.zeros((len(self.measures), 4), np.float32)
        p = L.DfqQuantParams()
        for i, (lo, hi) in enumerate(rng):
            L.lib().dfq_quant_params_host(float(lo), float(hi), 8, 0, ctypes.byref(p))
            host[i] = (lo, hi, p.scale, 0.0)
        if getattr(self, "slots", None) is not None and self.slots.shape[0] == host.shape[0]:
            self.slots.copy_(torch.from_numpy(host))
        else:
            self.slots = torch.from_numpy(host).to(self.dev)

    def launch(self):
        for _name, fn in self.ops:
            fn()

    def set_range(self, idx, lo, hi):
        p = L.DfqQuantParams()
        L.lib().dfq_quant_params_host(float(lo), float(hi), 8, 0, ctypes.byref(p))
        self.slots[idx].copy_(torch.tensor([float(lo), float(hi), p.scale, 0.0], dtype=torch.float32))

    def run_step(self, name, in_codes=None, image=None):
        """Run ONE op on given operand codes (uint8, logical NCHW / NC shape, one per operand) -- or the fp32 image for
        the stem -- and return {consumer slot: uint8 codes in the output's logical shape} (plus "fp32" for an op that
        writes fp32).  Ranges are whatever the slots hold (set_range)."""
        io = self.step_io[name]
        if image is not None:
            self.input.buf.copy_(image)
        if in_codes is not None:
            for idx, c in zip(io["in_slots"], in_codes):
                self.codes[idx].copy_(to_codes_nhwc(c).reshape(-1))
        io["fn"]()
        res = {idx: from_codes_nhwc(self.codes[idx], io["out"].shape).cpu() for idx in io["out_slots"]}
        if io["out"].buf is not None and io["out"].buf.dtype == torch.float32 and io["out"] is not self.input:
            res["fp32"] = io["out"].buf.clone().cpu()
        return res

    def debug_run(self, x):
        raise NotImplementedError("use run_step")

    def teacher_force(self, *a, **k):
        raise NotImplementedError("use run_step")
