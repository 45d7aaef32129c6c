"""Static-range W8A8 eval plan: the quantised forward with every QuantMeasure in `update_stat == False` mode
(utils/quantize.py:119-126, set_update_stat), compiled into launches of csrc/q8.cu.

With static ranges a consumer's grid is known before its producer runs, so activations never exist as fp32 in HBM:
every producer requantises in its epilogue to the grid of each consumer's QuantMeasure (a tensor read by a
convolution AND a residual add is written twice, once per grid -- the reference quantises it twice, with the two
measures' ranges) and stores centred s8 codes, NHWC, channels padded to 16.  The 1 x 1 convolutions and the
classifier read those codes by tensor-map TMA straight into the tcgen05 operand layout.  13.9 MB per image of
activation traffic for MobileNetV2 instead of 55.8 MB with fp32 activations (eval_plan.py).

Supported graph: what MobileNetV2 and DeepLabV3+ (MobileNetV2 backbone) trace to -- a dense 3 x 3 stem on the fp32
image, 1 x 1 convolutions, depthwise 3 x 3 (stride 1 / 2, any dilation; conv zero padding or F.pad ahead of the
measure, which is folded into the convolution), dense k x k convolutions (atrous ASPP branches, decoder) on tcgen05,
residual adds, the spatial mean, the global average pool, bilinear interpolation, channel concatenation, the linear
classifier; ReLU fused, BatchNorm already merged.  A tensor read by several QuantMeasures with bit-identical ranges is
written once.  Anything else raises NotImplementedError: it runs on the dynamic-range plan (eval_plan.EvalPlan).

Semantics per op are those of the reference with replace_op() in force (utils/layer_transform.py:18-125,
utils/quantize.py:250-256, 358-363); bits_activation is 8 (the reference's set_layer_bits never changes it).
"""
import ctypes
import os

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


class PackedConvQ8:
    """Weight codes of a dense k x k convolution in the layout of the TMA-fed kernel (dfq_conv_q8 with tma = 1):
    reduction tap-major, then 128-byte channel slabs; plus the per-tap code sums the zero-padding correction needs."""

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
        O, C, kh, kw = w.shape
        taps, Cp = kh * kw, _cp(C)
        codes = ((q - 2 ** (bits - 1)) * up).to(torch.int8).cpu().numpy().reshape(O, C, taps)
        lay = np.zeros((O, taps, Cp), dtype=np.int8)
        lay[:, :, :C] = codes.transpose(0, 2, 1)
        lib = L.lib()
        kq, nbytes = ctypes.c_int(), L.i64()
        self.o_cap = int(o_cap)
        L.check(lib.dfq_conv_q8_plan_host(O, C, taps, self.o_cap, ctypes.byref(kq), ctypes.byref(nbytes)),
                "dfq_conv_q8_plan_host")
        packed = np.empty(nbytes.value, dtype=np.int8)
        lay = np.ascontiguousarray(lay)
        L.check(lib.dfq_conv_q8_pack_host(lay.ctypes.data, O, C, taps, self.o_cap, packed.ctypes.data),
                "dfq_conv_q8_pack_host")
        self.s_w = s32 / up
        self.c_w = float(np.float32(lo + 2 ** (bits - 1) * s32))
        tap_sums = codes.astype(np.int64).sum(1)                                   # [O, taps]
        wr = (tap_sums.sum(1).astype(np.float64) * self.s_w).astype(np.float32)
        wrt = np.ascontiguousarray((tap_sums.T.astype(np.float64) * self.s_w).astype(np.float32))   # [taps, O]
        dev = w.device
        self.wpk, self.wr, self.wrt = (torch.from_numpy(packed).to(dev), torch.from_numpy(wr).to(dev),
                                       torch.from_numpy(wrt).to(dev))
        self.O = O


def pack_conv_halo(w):
    """Weights of a 3 x 3 convolution for the halo kernel (dfq_conv_q8 with tma = 2): pw_i8 packing over a reduction
    ordered [tap][16-byte channel chunk], the chunks of a tap zero-padded to a multiple of 4."""
    base = pw_i8.pack_weight_conv(w, None, o_cap=0)          # recovers the codes; wr / wrt / scale as for dfq_conv_i8
    if base is None:
        return None
    O, C, kh, kw = w.shape
    kq = L.lib().dfq_conv_q8_halo_k_host(C)
    per_tap = kq // 9
    lay = np.zeros((O, 9, per_tap), dtype=np.int8)
    lay[:, :, :C] = base.codes.transpose(0, 2, 1)            # codes: [O, C, taps] centred s8
    lay = np.ascontiguousarray(lay.reshape(O, kq))
    lib = L.lib()
    n_pass, opp, npass, slabs, nbytes = (ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), L.i64())
    L.check(lib.dfq_pw_i8_plan_cap_host(O, kq, 0, ctypes.byref(n_pass), ctypes.byref(opp), ctypes.byref(npass),
                                        ctypes.byref(slabs), ctypes.byref(nbytes)), "dfq_pw_i8_plan_cap_host")
    packed = np.empty(nbytes.value, dtype=np.int8)
    L.check(lib.dfq_pw_i8_pack_cap_host(lay.ctypes.data, O, kq, 0, packed.ctypes.data), "dfq_pw_i8_pack_cap_host")
    base.wpk = torch.from_numpy(packed).to(w.device)
    base.o_cap = 0
    return base


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


def pack_stem(w):
    """[O <= 32, 3, 3, 3] fake-quantised stem weight -> ([32][8] words of centred s8 codes (word k = taps 4k .. 4k + 3 in
    (c, kh, kw) order, zero beyond tap 26 and beyond row O), s_w, c_w) for dfq_stem_q8's integer kernel; None when the
    weight is not on a grid of <= 8 bits (the fp32 kernel takes it then)."""
    rec = None
    for b in (8, 7, 6, 5, 4, 3, 2):
        rec = pw_i8._recover_codes(w, b)
        if rec is not None:
            break
    if rec is None or w.shape[0] > 32 or tuple(w.shape[1:]) != (3, 3, 3):
        return None
    s32, lo, q, bits = rec
    up = 2 ** (8 - bits)
    codes = np.zeros((32, 32), np.int8)
    codes[:w.shape[0], :27] = ((q - 2 ** (bits - 1)) * up).to(torch.int8).cpu().numpy().reshape(w.shape[0], 27)
    words = np.ascontiguousarray(codes).view(np.int32).reshape(32, 8)
    return torch.from_numpy(words.copy()).to(w.device), s32 / up, float(np.float32(lo + 2 ** (bits - 1) * s32))


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
        dev = self.dev
        for qm, _idx, _t in self.measures:
            if qm.update_stat:
                raise RuntimeError("q8 plan: every QuantMeasure must be in static mode (set_update_stat(False), "
                                   "utils/quantize.py:125-126); the dynamic-range plan is eval_plan.EvalPlan")
            if int(qm.num_bits) != 8:
                raise NotImplementedError("q8 plan: 8-bit activations only")
        self.refresh_static()
        # F.pad is never quantised itself (utils/layer_transform.py ignore_op): the QuantMeasure of the convolution that
        # reads the padded tensor quantises the padded zeros with it.  So the producer of x writes x on that measure's
        # grid: into the interior of a pre-bordered padded tensor when it can place pixels (outs_of below), else plain,
        # and a pad op copies it into a padded code tensor whose border is code(0.0) on that grid.
        live = steps
        slots_in = lambda st: [st[k] for k in ("slot", "slot_a", "slot_b") if st.get(k) is not None] + \
            list(st.get("slots", ()))
        ins_of = lambda st: [st[k] for k in ("src", "a", "b") if k in st] + list(st.get("srcs", ()))
        direct, raw, via_pad = {}, set(), {}
        for _qm, idx, t in self.measures:
            direct.setdefault(id(t), []).append(idx)
        padded = {id(st["out"]) for st in live if st["kind"] == "pad"}
        for st in live:
            if st["kind"] == "pad":
                if id(st["src"]) in padded:
                    raise NotImplementedError(f"{st['name']}: F.pad of a padded tensor")
                via_pad.setdefault(id(st["src"]), []).extend(direct.get(id(st["out"]), []))
            elif st["kind"] == "pool":
                raw.add(id(st["src"]))
        # consumers with bit-identical ranges share one code buffer (ASPP: five branches read one tensor on one grid)
        host = self.slots_host
        same = lambda i, j: host[i, :2].tobytes() == host[j, :2].tobytes()

        def group(idxs):
            reps = []
            for idx in idxs:
                if not any(same(r, idx) for r in reps):
                    reps.append(idx)
            return reps
        rep_in = lambda reps, idx: next(k for k, r in enumerate(reps) if same(r, idx))
        step_of_slot = {}
        for i, st in enumerate(live):
            for idx in slots_in(st):
                step_of_slot[idx] = i
        slot_ptr = lambda i: self.slots[i].data_ptr()
        pool, self.codes, self.pre, self.slot_key = {}, {}, {}, {}
        self.act_bytes = 0
        self.f32_of = {}

        def code_shape(t):
            return (t.shape[0], t.shape[2], t.shape[3], _cp(t.shape[1])) if len(t.shape) == 4 else (t.shape[0], _cp(t.shape[1]))

        alive = []                       # [buffer, steps that still have to read it]

        def take(n):
            free = pool.get(n)
            buf = free.pop() if free else torch.zeros(n, dtype=torch.int8, device=dev)
            self.act_bytes += n
            return buf

        def retire(i):
            for ent in alive:
                ent[1].discard(i)
            for ent in alive:
                if not ent[1]:
                    pool.setdefault(ent[0].numel(), []).append(ent[0])
            alive[:] = [e for e in alive if e[1]]

        self.shape_of, self.padded_bufs = {}, []
        pads_of = {}                      # id(x) -> pad steps reading x
        for st in live:
            if st["kind"] == "pad":
                pads_of.setdefault(id(st["src"]), []).append(st)

        def outs_of(t, name, can_map=False):
            """Code buffers for the consumers of tensor t -> (y0, s0, y1, s1, slots served, map0, map1).
            One plain buffer per distinct grid of the measures on t itself.  Measures that see F.pad(t): when the
            producer can place pixels (can_map: stem, 1 x 1 convolution, residual add) it writes the INTERIOR of a
            dedicated padded tensor whose border was filled once with code(0.0) on that grid -- no pad launch; otherwise
            t is written plain on that grid (self.pre) and the pad op copies it."""
            mine, through = direct.get(id(t), []), via_pad.get(id(t), [])
            n = int(np.prod(code_shape(t)))
            outs = []                     # [buffer, slot idx, map]
            if can_map:
                reps = group(mine)
                ents = [[take(n), set()] for _ in reps]
                for idx in mine:
                    k = rep_in(reps, idx)
                    self.codes[idx], self.shape_of[idx] = ents[k][0], tuple(t.shape)
                    ents[k][1].add(step_of_slot[idx])
                alive.extend(ents)
                outs += [[e[0], r, None] for e, r in zip(ents, reps)]
                for pst in pads_of.get(id(t), []):
                    pt_, (pl, pr, ptop, pb) = pst["out"], pst["pad"]
                    idxs = direct.get(id(pt_), [])
                    for r in group(idxs):
                        buf = torch.zeros(int(np.prod(code_shape(pt_))), dtype=torch.int8, device=dev)   # never recycled
                        self.act_bytes += buf.numel()
                        m = L.DfqQ8OutMap()
                        m.OH, m.OW, m.pl, m.pr, m.pt, m.pb = t.shape[2], t.shape[3], pl, pr, ptop, pb
                        for idx in idxs:
                            if same(r, idx):
                                self.codes[idx], self.shape_of[idx] = buf, tuple(pt_.shape)
                        self.padded_bufs.append((buf, code_shape(pt_), pt_.shape[1], (pl, pr, ptop, pb), r))
                        outs.append([buf, r, m])
                served = list(mine) + list(through)
            else:
                reps = group(mine + through)
                ents = [[take(n), set()] for _ in reps]
                for idx in mine:
                    k = rep_in(reps, idx)
                    self.codes[idx], self.shape_of[idx] = ents[k][0], tuple(t.shape)
                    ents[k][1].add(step_of_slot[idx])
                for idx in through:                   # read by the pad op, which writes what the measure's layer reads
                    k = rep_in(reps, idx)
                    self.pre[idx] = ents[k][0]
                    ents[k][1].add(next(i for i, st in enumerate(live) if st["kind"] == "pad" and st["src"] is t
                                        and idx in direct.get(id(st["out"]), [])))
                alive.extend(ents)
                outs += [[e[0], r, None] for e, r in zip(ents, reps)]
                served = list(mine)
            if len(outs) > 2:
                raise NotImplementedError(f"{name}: {len(outs)} distinct consumer tensors (the q8 kernels write two)")
            y0 = outs[0][0].data_ptr() if outs else None
            y1 = outs[1][0].data_ptr() if len(outs) > 1 else None
            return (y0, slot_ptr(outs[0][1]) if outs else None, y1, slot_ptr(outs[1][1]) if len(outs) > 1 else None,
                    served, outs[0][2] if outs else None, outs[1][2] if len(outs) > 1 else None)

        def f32_out(t, shape):
            buf = torch.empty(*shape, dtype=torch.float32, device=dev)
            self.f32_of[id(t)] = buf
            return buf

        self.input.buf = torch.empty(*self.input.shape, dtype=torch.float32, device=dev)
        self.keep_q8 = []
        self.step_io = {}
        self.out_of = {}
        lib, stream = L.lib(), L.stream
        top_i = 0
        for i, st in enumerate(live):
            kind, name, out = st["kind"], st["name"], st["out"]
            is_final = out is self.output
            io_extra = {}
            if kind == "pad":
                src = st["src"]
                pl, pr, pt, pb = st["pad"]
                Nn, C, H, W = src.shape
                oidx = [idx for idx in direct.get(id(out), []) if idx not in self.codes]
                if not oidx:                 # the producer wrote the interior of the padded tensor itself
                    continue
                reps = group(oidx)
                n = int(np.prod(code_shape(out)))
                calls = []
                for r in reps:
                    ent = [take(n), set()]
                    for idx in oidx:
                        if same(r, idx):
                            self.codes[idx], self.shape_of[idx] = ent[0], tuple(out.shape)
                            ent[1].add(step_of_slot[idx])
                    alive.append(ent)
                    calls.append((self.pre[r].data_ptr(), slot_ptr(r), Nn, C, H, W, pl, pr, pt, pb, ent[0].data_ptr()))

                def fn(calls=calls):
                    for a2 in calls:
                        L.check(lib.dfq_pad_q8(*a2, stream()), "dfq_pad_q8")
                kind2 = "pad"
                io_extra = {"pre_slots": reps, "pad": (pl, pr, pt, pb)}
                self.ops.append((name, fn))
                self.out_of[name] = out
                self.step_io[name] = {"kind": kind2, "in_slots": [], "srcs": [src], "out": out, "out_slots": oidx,
                                      "relu": 0, "fn": fn, **io_extra}
                retire(i)
                continue
            can_map = kind == "add"
            fork_info = None              # (source tensor, buffers read, buffers written) of a 1 x 1 / dense convolution
            if kind == "conv":
                kh_, kw_, s_, p_, d_, g_ = st["geom"]
                can_map = (st["src"] is self.input) or (kh_ == 1 and kw_ == 1 and g_ == 1 and s_ == 1 and p_ == 0)
            if os.environ.get("DFQ_Q8_PAD_OP", "0") == "1":      # A/B switch: F.pad as its own launch (dfq_pad_q8)
                can_map = False
            y0, s0, y1, s1, oidx, m0, m1 = outs_of(out, name, can_map)
            if kind == "conv":
                m, src = st["m"], st["src"]
                kh, kw, s, p, d, g = st["geom"]
                if st["slot"] is None:
                    raise NotImplementedError(f"{name}: convolution without a QuantMeasure")
                self.slot_key[st["slot"]] = ("L", name)
                w = m.weight.data
                b = None if m.bias is None else m.bias.data
                Nn, C, H, W = src.shape
                xcodes = self.codes.get(st["slot"])
                if src is self.input:
                    if not (kh == 3 and kw == 3 and g == 1 and C == 3 and d == 1 and m.out_channels <= 32 and y0 is not None):
                        raise NotImplementedError(f"{name}: only a dense 3 x 3 stem on 3 channels reads the image")
                    desc = L.DfqStemQ8Desc()
                    desc.x, desc.w = src.buf.data_ptr(), L.dev_f32(w.contiguous(), "weight")
                    desc.bias = None if b is None else L.dev_f32(b, "bias")
                    desc.in_slot, desc.y0, desc.out_slot0, desc.y1, desc.out_slot1 = slot_ptr(st["slot"]), y0, s0, y1, s1
                    desc.N, desc.H, desc.W, desc.O, desc.stride, desc.pad, desc.relu = Nn, H, W, m.out_channels, s, p, st["relu"]
                    if m0 is not None:
                        desc.map0 = m0
                    if m1 is not None:
                        desc.map1 = m1
                    pk = pack_stem(w) if os.environ.get("DFQ_Q8_STEM_I8", "1") != "0" else None
                    if pk is not None:                  # weights on a grid: exact integer sums (dp4a)
                        desc.wq, desc.s_w, desc.c_w = pk[0].data_ptr(), pk[1], pk[2]
                        self.keep_q8.append(pk[0])
                    self.keep_q8.append(desc)
                    fn = (lambda dd: lambda: L.check(lib.dfq_stem_q8(ctypes.byref(dd), stream()), "dfq_stem_q8"))(desc)
                    kind2 = "stem"
                elif kh == 1 and kw == 1 and g == 1 and s == 1 and p == 0:
                    yf = None
                    if is_final or y0 is None:
                        raise NotImplementedError(f"{name}: a 1 x 1 convolution as the network's output")
                    if id(out) in raw:
                        yf = f32_out(out, (Nn * H * W, m.out_channels))      # fp32 NHWC rows next to the codes
                    fn = self._pw(name, w.reshape(w.shape[0], -1), b, xcodes, slot_ptr(st["slot"]),
                                  Nn * H * W, y0, s0, y1, s1, st["relu"], yf, m0, m1)
                    kind2 = "pw"
                    fork_info = (id(src), {xcodes.data_ptr()}, {y0, y1, None if yf is None else yf.data_ptr()} - {None})
                elif kh == 3 and kw == 3 and g == C == m.out_channels and s in (1, 2) and y0 is not None:
                    if p > d:
                        raise NotImplementedError(f"{name}: depthwise padding {p} with dilation {d}")
                    pad_each = p
                    Cp = _cp(C)
                    wrow, dsw, dcw = pack_depthwise(w)
                    bp = torch.zeros(Cp, dtype=torch.float32, device=dev)
                    if b is not None:
                        bp[:C] = b
                    desc = L.DfqDwQ8Desc()
                    desc.x, desc.wrow, desc.bias = xcodes.data_ptr(), wrow.data_ptr(), bp.data_ptr()
                    desc.s_w, desc.c_w = dsw, dcw
                    desc.in_slot, desc.y0, desc.out_slot0, desc.y1, desc.out_slot1 = slot_ptr(st["slot"]), y0, s0, y1, s1
                    desc.N, desc.H, desc.W, desc.C, desc.stride, desc.relu = Nn, H, W, C, s, st["relu"]
                    desc.pad, desc.dil = pad_each, d
                    self.keep_q8 += [wrow, bp, desc]
                    fn = (lambda dd: lambda: L.check(lib.dfq_dw_q8(ctypes.byref(dd), stream()), "dfq_dw_q8"))(desc)
                    kind2 = "dw"
                elif g == 1 and y0 is not None:        # dense k x k (ASPP's atrous 3 x 3, the decoder's 3 x 3)
                    G_out = out.shape[0] * out.shape[2] * out.shape[3]
                    # DFQ_CONV_Q8_TMA=1 selects the TMA-fed, warp-specialised form (stride 1).  Same codes out; measured
                    # SLOWER on the decoder's 304 / 256 -> 256 layers at 129 x 129 (0.71 / 0.53 ms against 0.35 / 0.33 ms):
                    # a 128-pixel tile re-streams the whole weight tensor through L2 -> shared memory (~8.9 TB/s
                    # measured), and the row-of-ones trick that feeds SUM_k a costs a third pass at 256 output
                    # channels, while the register-gather kernel forms the code sums with dp4a and runs ONE pass of
                    # N = 256.  Kept selectable (and tested) as the base for a cluster / multicast version.
                    tma = 1 if s == 1 and os.environ.get("DFQ_CONV_Q8_TMA", "0") == "1" else 0
                    # 3 x 3, stride 1, pad 1, dilation 1 (the decoder): halo tile in shared memory, no gather
                    if (kh, kw, s, p, d) == (3, 3, 1, 1, 1) and os.environ.get("DFQ_CONV_Q8_HALO", "1") != "0" and not tma:
                        tma = 2
                    if tma == 2:
                        pk = pack_conv_halo(w)
                    elif tma:
                        pk = PackedConvQ8(w, o_cap=choose_o_cap(G_out, w.shape[0]))
                    else:
                        pk = pw_i8.pack_weight_conv(w, None, o_cap=pw_i8.choose_o_cap(G_out, w.shape[0]))
                    if pk is None:
                        raise RuntimeError(f"{name}: weight is not on an integer grid of <= 8 bits")
                    desc = L.DfqConvQ8Desc()
                    desc.x, desc.wpk, desc.wr, desc.wrt = xcodes.data_ptr(), pk.wpk.data_ptr(), pk.wr.data_ptr(), pk.wrt.data_ptr()
                    desc.bias = None if b is None else L.dev_f32(b, "bias")
                    desc.in_slot, desc.y0, desc.out_slot0, desc.y1, desc.out_slot1 = slot_ptr(st["slot"]), y0, s0, y1, s1
                    desc.N, desc.C, desc.H, desc.W = Nn, C, H, W
                    desc.O, desc.OH, desc.OW = out.shape[1], out.shape[2], out.shape[3]
                    desc.kh, desc.kw, desc.stride, desc.pad, desc.dil, desc.relu = kh, kw, s, p, d, st["relu"]
                    desc.s_w, desc.c_w, desc.o_cap, desc.tma = pk.s_w, pk.c_w, pk.o_cap, tma
                    self.keep_q8 += [pk, desc, b]
                    fork_info = (id(src), {xcodes.data_ptr()}, {y0, y1} - {None})
                    fn = (lambda dd: lambda: L.check(lib.dfq_conv_q8(ctypes.byref(dd), stream()), "dfq_conv_q8"))(desc)
                    kind2 = "dense"
                else:
                    raise NotImplementedError(f"{name}: convolution geometry {st['geom']} is not covered by the q8 plan")
            elif kind == "linear":
                m, src = st["m"], st["src"]
                if st["slot"] is None:
                    raise NotImplementedError(f"{name}: linear layer without a QuantMeasure")
                self.slot_key[st["slot"]] = ("L", name)
                yf = None
                if is_final or y0 is None:
                    yf = f32_out(out, out.shape)
                    out.buf = yf
                fn = self._pw(name, m.weight.data, None if m.bias is None else m.bias.data, self.codes[st["slot"]],
                              slot_ptr(st["slot"]), src.shape[0], y0, s0, y1, s1, st["relu"], yf)
                kind2 = "linear"
            elif kind == "add":
                self.slot_key[st["slot_a"]] = ("T", top_i, 0)
                self.slot_key[st["slot_b"]] = ("T", top_i, 1)
                top_i += 1
                if y0 is None:
                    raise NotImplementedError(f"{name}: residual add without a quantised consumer")
                Nn, C, H, W = out.shape
                args = (self.codes[st["slot_a"]].data_ptr(), self.codes[st["slot_b"]].data_ptr(), slot_ptr(st["slot_a"]),
                        slot_ptr(st["slot_b"]), y0, s0, y1, s1, Nn * H * W, C,
                        None if m0 is None else ctypes.pointer(m0), None if m1 is None else ctypes.pointer(m1))
                self.keep_q8 += [m0, m1]
                fn = (lambda a2: lambda: L.check(lib.dfq_add_q8(*a2, stream()), "dfq_add_q8"))(args)
                kind2 = "add"
            elif kind == "mean":
                src = st["src"]
                self.slot_key[st["slot"]] = ("T", top_i, 0)
                top_i += 1
                Nn, C, H, W = src.shape
                yf = None
                if y0 is None:
                    out.buf = f32_out(out, out.shape)
                    yf = out.buf.data_ptr()
                args = (self.codes[st["slot"]].data_ptr(), slot_ptr(st["slot"]), y0, s0, yf, Nn, C, H * W)
                fn = (lambda a2: lambda: L.check(lib.dfq_mean_q8(*a2, stream()), "dfq_mean_q8"))(args)
                kind2 = "mean"
            elif kind == "pool":
                src = st["src"]
                x32 = self.f32_of.get(id(src))
                if x32 is None:
                    raise NotImplementedError(f"{name}: the pooled tensor must come from a 1 x 1 convolution of the plan")
                if y0 is None:
                    raise NotImplementedError(f"{name}: global pool without a quantised consumer")
                Nn, C, H, W = src.shape
                args = (x32.data_ptr(), C, Nn, H * W, C, y0, s0, y1, s1)
                fn = (lambda a2: lambda: L.check(lib.dfq_pool_q8(*a2, stream()), "dfq_pool_q8"))(args)
                kind2 = "pool"
            elif kind == "interp":
                src = st["src"]
                self.slot_key[st["slots"][0]] = ("T", top_i, 0)
                top_i += 1
                Nn, C, H, W = src.shape
                OH, OW = st["size"]
                yf = None
                if is_final or y0 is None:
                    out.buf = f32_out(out, out.shape)               # fp32 NCHW: the network's output
                    yf = out.buf.data_ptr()
                args = (self.codes[st["slots"][0]].data_ptr(), slot_ptr(st["slots"][0]), Nn, C, H, W, OH, OW,
                        1 if st["align"] else 0, y0, s0, y1, s1, yf)
                fn = (lambda a2: lambda: L.check(lib.dfq_interp_q8(*a2, stream()), "dfq_interp_q8"))(args)
                kind2 = "interp"
            elif kind == "cat":
                if y0 is None:
                    raise NotImplementedError(f"{name}: concatenation without a quantised consumer")
                Nn, Ct, H, W = out.shape
                calls, c_off = [], 0
                for j, (t, sl) in enumerate(zip(st["srcs"], st["slots"])):
                    self.slot_key[sl] = ("T", top_i, j)
                    if c_off % 4:
                        raise NotImplementedError(f"{name}: operand {j} starts at channel {c_off} (not a multiple of 4)")
                    calls.append((self.codes[sl].data_ptr(), slot_ptr(sl), Nn * H * W, t.shape[1], y0, s0, y1, s1,
                                  _cp(Ct), c_off, 1 if j == len(st["srcs"]) - 1 else 0))
                    c_off += t.shape[1]
                top_i += 1

                def fn(calls=calls):
                    for a2 in calls:
                        L.check(lib.dfq_regrid_q8(*a2, stream()), "dfq_regrid_q8")
                kind2 = "cat"
            else:
                raise NotImplementedError(f"{name}: `{kind}` is not covered by the q8 plan (use eval_plan.EvalPlan)")
            self.ops.append((name, fn))
            self.out_of[name] = out
            self.step_io[name] = {"kind": kind2, "in_slots": slots_in(st), "srcs": ins_of(st), "out": out,
                                  "out_slots": oidx, "relu": st.get("relu", 0), "fn": fn, "fork": fork_info}
            retire(i)                                       # code buffers whose last reader this step was
        if self.output.buf is None:
            raise NotImplementedError("q8 plan: the network's output is not produced as fp32")
        self.op_info = {n: {"kind": io["kind"]} for n, io in self.step_io.items()}
        self.hbm_bytes = self.act_bytes
        self._fill_borders()
        self._schedule()

    def _schedule(self):
        """Sibling convolutions -- consecutive 1 x 1 / dense convolutions of ONE tensor, e.g. the four ASPP branches,
        each a latency-bound kernel of ~140 CTAs at 33 x 33 -- are launched side by side on forked streams (joined
        before the next op; captured into the CUDA graph as parallel branches).  A group forms only when no member
        writes a buffer another member reads or writes (the code-buffer pool recycles by size).  DFQ_Q8_FORK=0: one
        stream."""
        self.schedule, group = [], []
        use = os.environ.get("DFQ_Q8_FORK", "1") != "0"

        def flush():
            if group:
                self.schedule.append([fn for _name, fn, _f in group])
                group.clear()
        for name, fn in self.ops:
            f = self.step_io.get(name, {}).get("fork") if use else None
            ok = f is not None and len(group) < 4 and bool(group) and group[-1][2][0] == f[0] and all(
                not (f[2] & (g[2][1] | g[2][2])) and not (f[1] & g[2][2]) for g in group)
            if f is None:
                flush()
                self.schedule.append([fn])
                continue
            if not ok:
                flush()
            group.append((name, fn, f))
        flush()
        self.forked = [len(g) for g in self.schedule if len(g) > 1]
        self.side = [torch.cuda.Stream(device=self.dev) for _ in range(3)] if self.forked else []

    def _fill_borders(self):
        """Border of every producer-placed padded tensor: code(0.0) on its consumer's grid, i.e. what the reference's
        QuantMeasure makes of the zeros F.pad put there (utils/quantize.py:72-74 with x = 0), as centred s8; the tensor's
        padding channels stay s8 zero.  The interior is rewritten by the producer on every forward."""
        for buf, (N, Hp, Wp, Cp), C, (pl, pr, pt, pb), idx in getattr(self, "padded_bufs", ()):
            lo, _hi, scale = (np.float32(v) for v in self.slots_host[idx, :3])
            q = np.rint(np.float32(np.float32(0.0) + np.float32(-lo)) / scale)
            code = int(min(max(q, 0.0), 255.0)) - 128
            v = buf.view(N, Hp, Wp, Cp)
            for sl in ((slice(None), slice(0, pt)), (slice(None), slice(Hp - pb, Hp)),
                       (slice(None), slice(None), slice(0, pl)), (slice(None), slice(None), slice(Wp - pr, Wp))):
                v[sl][..., :C] = code

    def _pw(self, name, w2d, bias, xbuf, in_slot, G, y0, s0, y1, s1, relu, yf, m0=None, m1=None):
        pk = PackedQ8(w2d, o_cap=choose_o_cap(G, w2d.shape[0]))
        d = L.DfqPwQ8Desc()
        d.x, d.wpk, d.wr = xbuf.data_ptr(), pk.wpk.data_ptr(), pk.wr.data_ptr()
        d.bias = None if bias is None else L.dev_f32(bias, "bias")
        d.in_slot, d.y0, d.out_slot0, d.y1, d.out_slot1 = in_slot, y0, s0, y1, s1
        if yf is not None:
            d.yf, d.yf_g, d.yf_o = yf.data_ptr(), pk.O, 1
        d.G, d.K, d.O, d.relu, d.o_cap, d.s_w, d.c_w = int(G), pk.K, pk.O, int(relu), pk.o_cap, pk.s_w, pk.c_w
        if m0 is not None:
            d.map0 = m0
        if m1 is not None:
            d.map1 = m1
        self.keep_q8 += [pk, d, bias]
        lib, stream = L.lib(), L.stream
        return lambda: L.check(lib.dfq_pw_q8(ctypes.byref(d), stream()), "dfq_pw_q8")

    # ------------------------------------------------------------------------------------------------- run
    def refresh_static(self):
        """(min, max, scale) of every measure from its running range (one device->host copy)."""
        if not self.measures:
            self.slots = torch.zeros(1, 4, dtype=torch.float32, device=self.dev)
            self.slots_host = np.zeros((1, 4), np.float32)
            return
        rng = torch.stack([torch.cat([qm.running_min.detach().reshape(1), qm.running_max.detach().reshape(1)])
                           for qm, _i, _t in self.measures]).cpu().numpy()
        host = np.zeros((len(self.measures), 4), np.float32)
        p = L.DfqQuantParams()
        for i, (lo, hi) in enumerate(rng):
            L.lib().dfq_quant_params_host(float(lo), float(hi), 8, 0, ctypes.byref(p))
            host[i] = (lo, hi, p.scale, 0.0)
        self.slots_host = host
        if getattr(self, "slots", None) is not None and self.slots.shape[0] == host.shape[0]:
            self.slots.copy_(torch.from_numpy(host))
        else:
            self.slots = torch.from_numpy(host).to(self.dev)
        self._fill_borders()

    def launch(self):
        for grp in self.schedule:
            if len(grp) == 1:
                grp[0]()
                continue
            main = torch.cuda.current_stream()
            start = torch.cuda.Event()
            start.record(main)
            grp[0]()
            done = []
            for side, fn in zip(self.side, grp[1:]):
                side.wait_event(start)
                with torch.cuda.stream(side):
                    fn()
                ev = torch.cuda.Event()
                ev.record(side)
                done.append(ev)
            for ev in done:
                main.wait_event(ev)

    def set_range(self, idx, lo, hi):
        p = L.DfqQuantParams()
        L.lib().dfq_quant_params_host(float(lo), float(hi), 8, 0, ctypes.byref(p))
        self.slots[idx].copy_(torch.tensor([float(lo), float(hi), p.scale, 0.0], dtype=torch.float32))
        self.slots_host[idx] = (np.float32(lo), np.float32(hi), np.float32(p.scale), 0.0)
        if any(r == idx for *_rest, r in self.padded_bufs):
            self._fill_borders()

    def run_step(self, name, in_codes=None, image=None):
        """Run ONE op on given operand codes (uint8, logical NCHW / NC shape, one per operand) -- or the fp32 image for
        the stem -- and return {consumer slot: uint8 codes in the output's logical shape} (plus "fp32" for an op that
        writes fp32).  Ranges are whatever the slots hold (set_range)."""
        io = self.step_io[name]
        if image is not None:
            self.input.buf.copy_(image)
        if in_codes is not None:
            # a pad op reads what its producer wrote for the padded tensor's consumers (self.pre), any other op its
            # measures' code buffers
            bufs = [self.pre[i] for i in io["pre_slots"]] if io["kind"] == "pad" else [self.codes[i] for i in io["in_slots"]]
            for buf, c in zip(bufs, in_codes):
                buf.copy_(to_codes_nhwc(c).reshape(-1))
        io["fn"]()
        res = {idx: from_codes_nhwc(self.codes[idx], self.shape_of.get(idx, io["out"].shape)).cpu()
               for idx in io["out_slots"]}
        if io["out"].buf is not None and io["out"].buf.dtype == torch.float32 and io["out"] is not self.input:
            res["fp32"] = io["out"].buf.clone().cpu()
        return res

    def debug_run(self, x):
        raise NotImplementedError("use run_step")

    def teacher_force(self, *a, **k):
        raise NotImplementedError("use run_step")
