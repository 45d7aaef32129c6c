"""Static plan of the quantised eval forward: the traced graph compiled once into a flat list of
libdfq_b200 launches with preallocated HBM buffers -- no Python graph interpretation, no inspect.stack(), no
host synchronisation per layer, capturable in a CUDA graph.

Semantics are those of the reference's eval path (utils/quantize.py:106-123, 225-256, 336-363 and the patched
tensor ops of utils/layer_transform.py:18-125) with replace_op() in force:
  * every target layer fake-quantises its input with its own QuantMeasure (running range widened by the batch
    statistics when `update_stat` is truthy -- the set_layer_bits bug);
  * both operands of a residual add and the operand of the spatial mean go through the CustomTensorOP measures;
  * identity BatchNorms (after merge_batchnorm) vanish, ReLU is fused into the producing kernel.
Quantisation happens on load inside the consuming kernel; producers accumulate the per-sample extrema the next
measure needs in their epilogue.
"""
import ctypes
import operator
import os

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib as L
from .graph import _ADD

_PASS_METHODS = {"view", "reshape", "contiguous", "flatten"}


class _T:
    """A tensor of the plan: shape, buffer, optional per-sample stats table."""

    def __init__(self, shape):
        self.shape = tuple(shape)
        self.buf = None
        self.stats = None
        self.stats_encoded = True
        self.relu_of = None
        self.relu_done = False
        self.need_stats = False
        self.by_torch = False             # written by a torch op: its per-sample extrema need their own pass

    @property
    def numel(self):
        n = 1
        for s in self.shape:
            n *= s
        return n


class EvalPlan:
    def __init__(self, model, transformer, batch, in_shape=(3, 224, 224), pw_impl="i8", dense_impl="i8",
                 move_impl="native", sync_group=None):
        """pw_impl: "i8" = pointwise convs / linear layers on the tcgen05 s8 kernel (csrc/pw_i8.cu) wherever the
        layer's weights sit on an integer grid of <= 8 bits, "fp32" = the CUDA-core comparator everywhere.
        move_impl: "native" = pad / pooling / interpolate / cat as one-pass libdfq_b200 kernels (csrc/move.cu) that
        quantise on load and reduce the output statistics themselves; "torch" = the comparator (our fake-quant into a
        temporary, the torch op, a separate statistics pass)."""
        self.move_impl = move_impl
        # sync_stats mode: `batch` is this rank's SHARD of a batch split across `sync_group`; before every
        # running-range update the per-sample extrema of all shards are all-gathered (in rank order = sample order), so
        # mean_n(max_chw x) of utils/quantize.py:110-111 is taken over the UN-SHARDED batch and every rank reproduces
        # the single-GPU ranges bit for bit (74 tiny collectives per forward: a parity mode, not a fast one).
        self.sync_group = sync_group
        # Narrower tensor-core passes (pw_i8.choose_o_cap) for layers whose default grid leaves SMs idle.  Measured on
        # B200: every pass re-quantises its A tile, and the producer, not the accumulator width, sets a CTA's time --
        # narrowing EVERY under-filled layer made 576 -> 160 at 256 x 7 x 7 0.050 -> 0.074 ms (4 passes of 48) and
        # DeepLab's pointwise layers 2.10 -> 2.27 ms; a layer of a handful of CTAs gains (the classifier, 2 tiles x 4
        # passes: 0.100 -> 0.083 ms with 63 passes of 16).
        self.narrow_passes = float(os.environ.get("DFQ_NARROW_PASSES", "0.25"))
        self.model, self.fxg = model, transformer.fx_graph
        self.dev = next(model.parameters()).device
        self.N = int(batch)
        self.pw_impl = pw_impl
        self.dense_impl = dense_impl     # "i8": dense k x k on tcgen05 (dfq_conv_i8); "direct": CUDA-core direct form
        self.ops = []                # (name, callable)
        self.measures = []           # (qm, slot index, input tensor)
        self.keep = []               # ctypes objects / tensors that must outlive the plan
        self.i8_layers = []          # layers running on the tensor-core kernel
        self._build(in_shape)

    # ------------------------------------------------------------------------------------------------ build
    def _slot_for(self, qm, t):
        """Slot of QuantMeasure `qm` applied to tensor `t`; schedules the measure kernel before the consumer."""
        idx = len(self.measures)
        self.measures.append((qm, idx, t))
        if qm.update_stat:
            t.need_stats = True
        return idx

    def _build(self, in_shape):
        mods = dict(self.model.named_modules())
        top = getattr(self.model, "custom_tensor_op", None)
        top_qms = list(top.children()) if top is not None else []
        top_i = 0
        env = {}
        steps = []                   # symbolic ops, resolved to launches after buffer assignment
        for node in self.fxg.nodes:
            if node.op == "placeholder":
                t = _T((self.N,) + tuple(in_shape))
                t.stats_encoded = False
                env[node] = t
                self.input = t
            elif node.op == "call_module":
                m = mods[node.target]
                src = env[node.args[0]]
                if isinstance(m, nn.BatchNorm2d):
                    if not hasattr(m, "fake_weight"):
                        raise NotImplementedError(f"{node.target}: BatchNorm not merged (run merge_batchnorm first)")
                    env[node] = src                      # identity after merge_batchnorm
                elif isinstance(m, (nn.ReLU, nn.ReLU6)):
                    if isinstance(m, nn.ReLU6) or src.relu_of is None or src.relu_done:
                        raise NotImplementedError(f"{node.target}: stand-alone activation is not supported by the plan")
                    src.relu_of["relu"] = 1              # fuse into the producing conv
                    src.relu_done = True
                    env[node] = src
                elif isinstance(m, nn.Conv2d):
                    N, C, H, W = src.shape
                    kh, kw = m.kernel_size
                    s, p, d = m.stride[0], m.padding[0], m.dilation[0]
                    OH = (H + 2 * p - d * (kh - 1) - 1) // s + 1
                    OW = (W + 2 * p - d * (kw - 1) - 1) // s + 1
                    out = _T((N, m.out_channels, OH, OW))
                    step = {"kind": "conv", "m": m, "src": src, "out": out, "relu": 0, "name": node.target,
                            "geom": (kh, kw, s, p, d, m.groups)}
                    step["slot"] = self._slot_for(m.quant, src) if hasattr(m, "quant") else None
                    out.relu_of, out.relu_done = step, False
                    steps.append(step)
                    env[node] = out
                elif isinstance(m, nn.Linear):
                    out = _T((src.shape[0], m.out_features))
                    step = {"kind": "linear", "m": m, "src": src, "out": out, "relu": 0, "name": node.target}
                    step["slot"] = self._slot_for(m.quant, src) if hasattr(m, "quant") else None
                    out.relu_of, out.relu_done = step, False
                    steps.append(step)
                    env[node] = out
                elif isinstance(m, (nn.Dropout, nn.Identity)):
                    env[node] = src
                elif isinstance(m, nn.AdaptiveAvgPool2d) and m.output_size in (1, (1, 1)):
                    out = _T(src.shape[:2] + (1, 1))
                    out.by_torch = self.move_impl == "torch"; out.stats_encoded = not out.by_torch
                    steps.append({"kind": "pool", "src": src, "out": out, "name": node.target})
                    env[node] = out
                else:
                    raise NotImplementedError(f"{node.target}: {type(m).__name__} is not supported by the eval plan")
            elif node.op in ("call_function", "call_method"):
                key = node.meta.get("dfq_key")
                if key is not None and (node.target in _ADD or node.target in ("add", "add_")):
                    a, b = (env[x] for x in node.meta["dfq_operands"])
                    out = _T(a.shape)
                    qa, qb = top_qms[top_i], top_qms[top_i + 1]
                    top_i += 2
                    steps.append({"kind": "add", "a": a, "b": b, "out": out, "name": key,
                                  "slot_a": self._slot_for(qa, a), "slot_b": self._slot_for(qb, b),
                                  "bits": (qa.num_bits, qb.num_bits)})
                    env[node] = out
                elif key is not None and key.startswith("torch.mean"):
                    src = env[node.meta["dfq_operands"][0]]
                    N, C = src.shape[0], src.shape[1]
                    out = _T((N, C))
                    qm = top_qms[top_i]
                    top_i += 1
                    steps.append({"kind": "mean", "src": src, "out": out, "name": key, "bits": qm.num_bits,
                                  "slot": self._slot_for(qm, src)})
                    env[node] = out
                elif key is not None and key.startswith("F.pad"):
                    # utils/layer_transform.py ignore_op=['pad']: recorded in the graph, never quantised
                    src = env[node.meta["dfq_operands"][0]]
                    pad = tuple(int(v) for v in (node.args[1] if len(node.args) > 1 else node.kwargs["pad"]))
                    if len(pad) != 4 or node.kwargs.get("mode", "constant") != "constant" or \
                            (node.args[3] if len(node.args) > 3 else node.kwargs.get("value", 0)) not in (0, 0.0, None):
                        raise NotImplementedError(f"{key}: only constant zero padding of H and W")
                    N, C, H, W = src.shape
                    out = _T((N, C, H + pad[2] + pad[3], W + pad[0] + pad[1]))
                    out.by_torch = self.move_impl == "torch"; out.stats_encoded = not out.by_torch
                    steps.append({"kind": "pad", "src": src, "out": out, "name": key, "pad": pad})
                    env[node] = out
                elif key is not None and key.startswith("F.interpolate"):
                    src = env[node.meta["dfq_operands"][0]]
                    kw = {k: (env[v] if isinstance(v, torch.fx.Node) else v) for k, v in node.kwargs.items()}
                    size = kw.get("size", env[node.args[1]] if len(node.args) > 1 and
                                  isinstance(node.args[1], torch.fx.Node) else (node.args[1] if len(node.args) > 1 else None))
                    if size is None or kw.get("mode", "nearest") != "bilinear":
                        raise NotImplementedError(f"{key}: only bilinear interpolation to a given size")
                    size = tuple(int(v) for v in size)
                    qm = top_qms[top_i]
                    top_i += 1
                    out = _T(src.shape[:2] + size)
                    out.by_torch = self.move_impl == "torch"; out.stats_encoded = not out.by_torch
                    steps.append({"kind": "interp", "src": src, "out": out, "name": key, "size": size,
                                  "align": bool(kw.get("align_corners", False)), "bits": [qm.num_bits],
                                  "slots": [self._slot_for(qm, src)]})
                    env[node] = out
                elif key is not None and key.startswith("torch.cat"):
                    srcs = [env[x] for x in node.meta["dfq_operands"]]
                    dim = node.kwargs.get("dim", node.args[1] if len(node.args) > 1 else 0)
                    if dim != 1 or any(t.shape[0] != srcs[0].shape[0] or t.shape[2:] != srcs[0].shape[2:] for t in srcs):
                        raise NotImplementedError(f"{key}: only channel concatenation")
                    qms = top_qms[top_i:top_i + len(srcs)]
                    top_i += len(srcs)
                    out = _T((srcs[0].shape[0], sum(t.shape[1] for t in srcs)) + srcs[0].shape[2:])
                    out.by_torch = self.move_impl == "torch"; out.stats_encoded = not out.by_torch
                    steps.append({"kind": "cat", "srcs": srcs, "out": out, "name": key,
                                  "bits": [q.num_bits for q in qms],
                                  "slots": [self._slot_for(q, t) for q, t in zip(qms, srcs)]})
                    env[node] = out
                elif key is not None:
                    raise NotImplementedError(f"{key}: tensor op not supported by the eval plan yet")
                elif node.op == "call_method" and node.target == "size":
                    t = env.get(node.args[0])
                    shape = torch.Size(t.shape) if isinstance(t, _T) else None
                    env[node] = shape if len(node.args) == 1 or shape is None else shape[node.args[1]]
                elif node.op == "call_method" and node.target == "dim":
                    env[node] = None
                elif node.target is operator.getitem and not isinstance(env.get(node.args[0]), _T) \
                        and env.get(node.args[0]) is not None:
                    env[node] = env[node.args[0]][node.args[1]]            # indexing a folded shape
                elif (node.op == "call_method" and node.target in _PASS_METHODS) or node.target is operator.getitem:
                    first = node.args[0]
                    env[node] = env.get(first) if isinstance(first, torch.fx.Node) else None
                else:
                    raise NotImplementedError(f"{node.target}: not supported by the eval plan")
            elif node.op == "output":
                self.output = env[node.args[0]]
        self._assign(steps)

    # ------------------------------------------------------------------------------------- buffers + launches
    def _assign(self, steps):
        dev, N = self.dev, self.N
        lib = L.lib()
        f32 = lambda *shape: torch.empty(*shape, dtype=torch.float32, device=dev)
        self.slots = torch.zeros(max(1, len(self.measures)), 4, dtype=torch.float32, device=dev)
        # last use of every tensor -> simple free-list reuse of activation buffers
        last = {}
        ins_of = lambda st: [st[k] for k in ("src", "a", "b") if k in st] + list(st.get("srcs", ()))
        for i, st in enumerate(steps):
            for t in ins_of(st):
                last[id(t)] = i
        last[id(self.output)] = len(steps)
        pool = {}
        self.input.buf = f32(*self.input.shape)
        tensors = [self.input]
        for i, st in enumerate(steps):
            out = st["out"]
            free = pool.get(out.numel)
            out.buf = free.pop() if free else f32(*out.shape)
            tensors.append(out)
            for t in {id(t): t for t in ins_of(st)}.values():
                if last[id(t)] == i and t is not self.input and t.buf is not None:
                    pool.setdefault(t.numel, []).append(t.buf)
        need = [t for t in tensors if t.need_stats]
        self.stats = torch.zeros(max(1, len(need)), N, 2, dtype=torch.int32, device=dev)
        for i, t in enumerate(need):
            t.stats = self.stats[i]
        self.hbm_bytes = sum(b.numel() * 4 for b in {id(t.buf): t.buf for t in tensors}.values())
        stream = L.stream
        slot_ptr = lambda i: None if i is None else self.slots[i].data_ptr()
        stats_ptr = lambda t: None if t.stats is None else t.stats.data_ptr()
        done = set()

        def measure(idx):
            """Launch closure of the measure kernel feeding slot `idx` (dynamic ranges run every forward)."""
            qm, _, t = self.measures[idx]
            upd = 1 if qm.update_stat else 0
            self.keep.append(qm)
            if upd and self.sync_group is not None:
                import torch.distributed as dist
                world = dist.get_world_size(self.sync_group)
                full = torch.empty(world * N, 2, dtype=t.stats.dtype, device=dev)
                self.keep.append(full)
                sargs = (full.data_ptr(), 1 if t.stats_encoded else 0, world * N, qm.running_min.data_ptr(),
                         qm.running_max.data_ptr(), slot_ptr(idx), int(qm.num_bits), 0, upd)
                group, local = self.sync_group, t.stats

                def synced():
                    dist.all_gather_into_tensor(full, local.view(N, 2), group=group)
                    L.check(lib.dfq_measure_from_stats(*sargs, stream()), "dfq_measure_from_stats")
                return synced
            args = (stats_ptr(t) if upd else None, 1 if t.stats_encoded else 0, N, qm.running_min.data_ptr(),
                    qm.running_max.data_ptr(), slot_ptr(idx), int(qm.num_bits), 0, upd)
            return lambda: L.check(lib.dfq_measure_from_stats(*args, stream()), "dfq_measure_from_stats")

        self.static_measures = []
        if self.input.need_stats:       # per-sample extrema of the network input (utils/quantize.py:110-111)
            in_args = (self.input.buf.data_ptr(), N, self.input.numel // N, self.input.stats.data_ptr())
            self.ops.append(("input_stats", lambda: L.check(
                lib.dfq_minmax_rows_f32(*in_args, stream()), "dfq_minmax_rows_f32")))
        self.out_of = {st["name"]: st["out"] for st in steps}
        self.op_info = {}
        self.step_io = {}            # name -> operands / range slots / output / launch closure (teacher_force)
        for st in steps:
            kind = st["kind"]
            if kind == "conv":
                kh, _, _, _, _, g = st["geom"]
                kind = "pw" if (kh == 1 and g == 1) else ("dw" if g > 1 else "dense")
            ins = ins_of(st)
            self.op_info[st["name"]] = {"kind": kind, "bytes": 4 * (sum(t.numel for t in ins) + st["out"].numel),
                                        "shape": tuple(ins[0].shape[1:]) + ("->",) + tuple(st["out"].shape[1:])}
        for st in steps:
            for idx in [st.get(key) for key in ("slot", "slot_a", "slot_b")] + list(st.get("slots", ())):
                if idx is None or idx in done:
                    continue
                done.add(idx)
                fn = measure(idx)
                if self.measures[idx][0].update_stat:
                    self.ops.append((f"measure[{st['name']}]", fn))
                else:
                    self.static_measures.append(fn)
            if st["kind"] == "conv":
                m, src, out = st["m"], st["src"], st["out"]
                kh, kw, s, p, d, g = st["geom"]
                desc = L.DfqConvDesc()
                wq, bq = self._params(m)
                desc.x, desc.y, desc.w = src.buf.data_ptr(), out.buf.data_ptr(), L.dev_f32(wq, "weight")
                desc.bias = None if bq is None else L.dev_f32(bq, "bias")
                desc.in_slot, desc.out_stats = slot_ptr(st["slot"]), stats_ptr(out)
                (desc.N, desc.C_in, desc.H, desc.W), (_, desc.C_out, desc.OH, desc.OW) = src.shape, out.shape
                desc.kh, desc.kw, desc.stride, desc.pad, desc.dil, desc.groups = kh, kw, s, p, d, g
                desc.relu, desc.bits = st["relu"], int(m.quant.num_bits) if hasattr(m, "quant") else 8
                self.keep.append(desc)
                i8 = None
                if (self.pw_impl == "i8" and kh == 1 and kw == 1 and g == 1 and s == 1 and p == 0
                        and st["slot"] is not None):
                    i8 = self._i8_launch(m, wq, bq, src, out, st, N, src.shape[2] * src.shape[3], linear=False)
                elif (self.pw_impl == "i8" and g == 1 and kh * kw > 1 and st["slot"] is not None
                        and self.dense_impl == "i8" and m.in_channels >= 16):
                    # (the 3-channel stem stays on the register-blocked direct kernel: 0.43 ms against 1.41 ms as
                    # an implicit GEMM whose 16-byte K chunks would be 13/16 padding)
                    i8 = self._conv_i8_launch(m, wq, bq, src, out, st, s, p, d)
                elif (g == 1 and kh == 3 and kw == 3 and d == 1 and src.shape[1] == 3 and m.out_channels <= 32
                        and st["slot"] is not None and os.environ.get("DFQ_STEM_I8", "1") != "0"):
                    i8 = self._stem_i8_launch(m, wq, bq, src, out, st, s, p)
                self.ops.append((st["name"], i8 or self._conv_launch(desc, st)))
            elif st["kind"] == "linear":
                m, src, out = st["m"], st["src"], st["out"]
                wq, bq = self._params(m)
                args = (src.buf.data_ptr(), out.buf.data_ptr(), L.dev_f32(wq, "weight"),
                        None if bq is None else L.dev_f32(bq, "bias"), src.shape[0], m.in_features,
                        m.out_features, slot_ptr(st["slot"]), int(m.quant.num_bits) if hasattr(m, "quant") else 8,
                        st["relu"], stats_ptr(out))
                i8 = None
                if self.pw_impl == "i8" and st["slot"] is not None:
                    i8 = self._i8_launch(m, wq, bq, src, out, st, src.shape[0], 1, linear=True)
                self.ops.append((st["name"], i8 or (lambda a: lambda: L.check(lib.dfq_qlinear_f32(*a, stream()),
                                                                             "dfq_qlinear_f32"))(args)))
            elif st["kind"] == "add":
                a, b, out = st["a"], st["b"], st["out"]
                args = (a.buf.data_ptr(), b.buf.data_ptr(), out.buf.data_ptr(), N, out.numel // N,
                        slot_ptr(st["slot_a"]), slot_ptr(st["slot_b"]), int(st["bits"][0]), int(st["bits"][1]),
                        stats_ptr(out))
                self.ops.append((st["name"], (lambda a_: lambda: L.check(lib.dfq_qadd_f32(*a_, stream()),
                                                                        "dfq_qadd_f32"))(args)))
            elif st["kind"] == "mean":
                src, out = st["src"], st["out"]
                args = (src.buf.data_ptr(), out.buf.data_ptr(), N, src.shape[1], src.shape[2] * src.shape[3],
                        slot_ptr(st["slot"]), int(st["bits"]), stats_ptr(out))
                self.ops.append((st["name"], (lambda a_: lambda: L.check(lib.dfq_qmean_f32(*a_, stream()),
                                                                        "dfq_qmean_f32"))(args)))
            elif st["kind"] in ("pad", "pool", "interp", "cat"):
                op = self._torch_op if self.move_impl == "torch" else self._move_op
                self.ops.append((st["name"], op(st, f32, slot_ptr)))
            slots_of = [st[k] for k in ("slot", "slot_a", "slot_b") if st.get(k) is not None] + list(st.get("slots", ()))
            bits_of = [int(self.measures[i][0].num_bits) for i in slots_of]
            self.step_io[st["name"]] = {"kind": st["kind"], "srcs": ins_of(st), "slots": slots_of, "bits": bits_of,
                                        "out": st["out"], "relu": st.get("relu", 0), "fn": self.ops[-1][1]}
            out = st["out"]
            if out.by_torch and out.need_stats:       # extrema of what the torch op wrote (plain floats)
                sargs = (out.buf.data_ptr(), N, out.numel // N, out.stats.data_ptr())
                self.ops.append((f"stats[{st['name']}]", (lambda a_: lambda: L.check(
                    lib.dfq_minmax_rows_f32(*a_, stream()), "dfq_minmax_rows_f32"))(sargs)))
        self.refresh_static()

    def _move_op(self, st, f32, slot_ptr):
        """Launch closure of a data-movement op on the one-pass kernels of csrc/move.cu: operand quantised on load
        with its QuantMeasure's range, destination written once, per-sample extrema of the result reduced in the
        same pass (so no temporary and no statistics pass)."""
        lib, stream = L.lib(), L.stream
        out, kind, N = st["out"], st["kind"], self.N
        stats = None if out.stats is None else out.stats.data_ptr()
        if kind == "pad":
            l, r, t, b = st["pad"]
            _, C, H, W = st["src"].shape
            args = (st["src"].buf.data_ptr(), out.buf.data_ptr(), N, C, H, W, l, r, t, b, stats)
            return lambda: L.check(lib.dfq_pad2d_f32(*args, stream()), "dfq_pad2d_f32")
        if kind == "pool":          # mean over H x W of the raw operand (nn.AdaptiveAvgPool2d is not a patched op)
            _, C, H, W = st["src"].shape
            args = (st["src"].buf.data_ptr(), out.buf.data_ptr(), N, C, H * W, None, 8, stats)
            return lambda: L.check(lib.dfq_qmean_f32(*args, stream()), "dfq_qmean_f32")
        if kind == "interp":
            _, C, H, W = st["src"].shape
            args = (st["src"].buf.data_ptr(), out.buf.data_ptr(), N, C, H, W, st["size"][0], st["size"][1],
                    1 if st["align"] else 0, slot_ptr(st["slots"][0]), int(st["bits"][0]), stats)
            return lambda: L.check(lib.dfq_qinterp_bilinear_f32(*args, stream()), "dfq_qinterp_bilinear_f32")
        launches, c0 = [], 0
        hw = out.shape[2] * out.shape[3]
        for t, slot, bits in zip(st["srcs"], st["slots"], st["bits"]):
            count = t.shape[1] * hw
            launches.append((t.buf.data_ptr(), out.buf.data_ptr() + 4 * c0 * hw, N, count, count, out.numel // N,
                             slot_ptr(slot), int(bits), stats))
            c0 += t.shape[1]

        def run():
            for a in launches:
                L.check(lib.dfq_qcopy_f32(*a, stream()), "dfq_qcopy_f32")
        return run

    def _torch_op(self, st, f32, slot_ptr):
        """Launch closure of a data-movement op of the graph (F.pad, adaptive average pooling, bilinear
        interpolation, channel cat: utils/layer_transform.py:18-125 patches the last two so that their operands go
        through QuantMeasures first).  The operands are fake-quantised with our kernel into a scratch buffer; the
        movement itself is the torch op the reference calls, writing into the plan's preallocated buffer (capturable
        in the CUDA graph)."""
        lib, stream = L.lib(), L.stream
        out = st["out"]
        dst = out.buf.view(out.shape)
        kind = st["kind"]

        def quantised(t, slot, bits):
            tmp = f32(*t.shape)
            self.keep.append(tmp)
            args = (t.buf.data_ptr(), tmp.data_ptr(), None, t.numel, int(bits), 0, slot_ptr(slot))
            return tmp, lambda: L.check(lib.dfq_fake_quant_apply_slot_f32(*args, stream()),
                                        "dfq_fake_quant_apply_slot_f32")
        if kind == "pad":
            src = st["src"].buf.view(st["src"].shape)
            l, r, t, b = st["pad"]
            H, W = st["src"].shape[2:]
            inner = dst[:, :, t:t + H, l:l + W]

            def run():
                dst.zero_()
                inner.copy_(src)
            return run
        if kind == "pool":
            src = st["src"].buf.view(st["src"].shape)
            return lambda: dst.copy_(F.adaptive_avg_pool2d(src, (1, 1)))
        if kind == "interp":
            tmp, q = quantised(st["src"], st["slots"][0], st["bits"][0])
            size, align = st["size"], st["align"]

            up = torch.ops.aten.upsample_bilinear2d.out       # the kernel F.interpolate dispatches to, no temporary

            def run():
                q()
                up(tmp, list(size), align, None, None, out=dst)
            return run
        parts, c0 = [], 0
        for t, slot, bits in zip(st["srcs"], st["slots"], st["bits"]):
            tmp, q = quantised(t, slot, bits)
            parts.append((q, dst[:, c0:c0 + t.shape[1]], tmp))
            c0 += t.shape[1]

        def run():
            for q, view, tmp in parts:
                q()
                view.copy_(tmp)
        return run

    def _params(self, m):
        """Weight / bias the layer's forward uses.  QuantN* layers use the stored parameters (quantised offline by
        quantize_targ_layer); Quant* layers re-quantise them on every forward (utils/quantize.py:227-231) -- the
        parameters are constant during eval, so that is done once here."""
        from .utils.quantize import _quant_weight, quantize
        w, b = m.weight.data, (None if m.bias is None else m.bias.data)
        if hasattr(m, "num_bits_bias"):
            with torch.no_grad():
                w = _quant_weight(w, m.num_bits)
                b = None if b is None else quantize(b, num_bits=m.num_bits_bias)
            self.keep += [w, b]
        return w, b

    def _i8_launch(self, m, wq, bq, src, out, st, images, P, linear):
        """Launch closure of the tensor-core kernel for this layer, or None when its weights are not on a grid."""
        from . import pw_i8
        if not 1 <= int(m.quant.num_bits) <= 8:
            return None                 # the s8 kernels cover activation grids of <= 8 bits: fp32 kernel instead
        pk = pw_i8.pack_weight(wq, getattr(m, "num_bits", None),
                               o_cap=pw_i8.choose_o_cap(images * P, wq.shape[0], fill=self.narrow_passes)
                               if self.narrow_passes > 0 else 0)
        if pk is None:
            return None
        stats = None if out.stats is None else out.stats.data_ptr()
        desc = pw_i8.make_desc(src.buf.data_ptr(), out.buf.data_ptr(), pk, bq, self.slots[st["slot"]].data_ptr(),
                               stats, images, P, st["relu"], int(m.quant.num_bits), linear=linear)
        self.keep += [pk, desc]
        self.i8_layers.append(st["name"])
        return lambda: pw_i8.launch(desc)

    def _conv_i8_launch(self, m, wq, bq, src, out, st, stride, pad, dil):
        """Dense k x k convolution as tcgen05 implicit GEMM (dfq_conv_i8), or None when the weights are off-grid."""
        from . import pw_i8
        if not 1 <= int(m.quant.num_bits) <= 8:
            return None
        # 3 x 3, stride 1, pad 1 on 8-bit activations (DeepLab's decoder): the halo kernel (csrc/conv_halo.cu)
        halo = (tuple(wq.shape[2:]) == (3, 3) and (stride, pad, dil) == (1, 1, 1) and int(m.quant.num_bits) == 8
                and os.environ.get("DFQ_CONV_HALO", "1") != "0")
        if halo:
            from .eval_q8 import pack_conv_halo
            pk = pack_conv_halo(wq)
        else:
            pk = pw_i8.pack_weight_conv(wq, getattr(m, "num_bits", None),
                                        o_cap=pw_i8.choose_o_cap(out.numel // out.shape[1], wq.shape[0],
                                                                 fill=self.narrow_passes)
                                        if self.narrow_passes > 0 else 0)
        if pk is None:
            return None
        stats = None if out.stats is None else out.stats.data_ptr()
        scratch = pw_i8.conv_scratch(src.shape, self.dev)        # the input quantised once, s8 NHWC
        desc = pw_i8.make_conv_desc(src.buf.view(src.shape), out.buf.view(out.shape), pk, bq,
                                    self.slots[st["slot"]].data_ptr(), stats, stride, pad, dil, st["relu"],
                                    int(m.quant.num_bits), scratch=scratch, halo=halo)
        self.keep += [pk, desc, scratch]
        self.i8_layers.append(st["name"])
        return lambda: pw_i8.launch_conv(desc)

    def _stem_i8_launch(self, m, wq, bq, src, out, st, stride, pad):
        """The 3-channel 3 x 3 stem on exact integer sums (dfq_stem_q8 with fp32 output: seven dp4a per output channel
        instead of 27 FMAs), or None when the weights are off-grid / the activations are not 8-bit."""
        from .eval_q8 import pack_stem
        if int(m.quant.num_bits) != 8:
            return None
        pk = pack_stem(wq)
        if pk is None:
            return None
        lib, stream = L.lib(), L.stream
        d = L.DfqStemQ8Desc()
        d.x, d.w = src.buf.data_ptr(), L.dev_f32(wq.contiguous(), "weight")
        d.bias = None if bq is None else L.dev_f32(bq, "bias")
        d.in_slot = self.slots[st["slot"]].data_ptr()
        d.N, d.H, d.W, d.O, d.stride, d.pad, d.relu = src.shape[0], src.shape[2], src.shape[3], m.out_channels, stride, pad, st["relu"]
        d.wq, d.s_w, d.c_w = pk[0].data_ptr(), pk[1], pk[2]
        d.yf = out.buf.data_ptr()
        d.out_stats = None if out.stats is None else out.stats.data_ptr()
        self.keep += [pk[0], d]
        self.i8_layers.append(st["name"])
        return lambda: L.check(lib.dfq_stem_q8(ctypes.byref(d), stream()), "dfq_stem_q8")

    def _conv_launch(self, desc, st):
        lib, stream = L.lib(), L.stream
        return lambda: L.check(lib.dfq_qconv2d_f32(ctypes.byref(desc), stream()), "dfq_qconv2d_f32")

    # ------------------------------------------------------------------------------------------------- run
    def refresh_static(self):
        """Recompute the slots of measures with update_stat == False (call after their ranges were changed)."""
        for fn in self.static_measures:
            fn()

    def launch(self):
        """All launches of one forward on the current stream (inputs already in self.input.buf)."""
        self.stats.zero_()
        for _name, fn in self.ops:
            fn()

    def teacher_force(self, name, inputs, ranges):
        """Run ONE op of the plan on given operands and given (min, max) ranges -- the reference's input activation
        and the range its QuantMeasure quantised with (tests/golden/eval_tf_*.npz) -- and return the op's output.
        The range slots are written directly ({min, max, scale, 0}, scale as utils/quantize.py:51-68 forms it), no
        measure kernel runs."""
        io = self.step_io[name]
        assert len(inputs) == len(io["srcs"]) and len(ranges) == len(io["slots"]), name
        for t, x in zip(io["srcs"], inputs):
            t.buf.copy_(x.reshape(t.buf.shape))
        for idx, bits, (lo, hi) in zip(io["slots"], io["bits"], ranges):
            p = L.DfqQuantParams()
            L.lib().dfq_quant_params_host(float(lo), float(hi), int(bits), 0, ctypes.byref(p))
            self.slots[idx].copy_(torch.tensor([float(lo), float(hi), p.scale, 0.0], dtype=torch.float32))
        self.stats.zero_()
        io["fn"]()
        return io["out"].buf.clone()

    def debug_run(self, x):
        """Forward that also returns a clone of every op's output (diagnostics; buffers are reused in the plan)."""
        self.input.buf.copy_(x)
        self.stats.zero_()
        outs = {}
        for name, fn in self.ops:
            fn()
            if name in self.out_of:
                outs[name] = self.out_of[name].buf.clone()
        return outs

    def __call__(self, x):
        self.input.buf.copy_(x, non_blocking=True)
        self.launch()
        return self.output.buf

    def capture(self):
        """Capture launch() into a CUDA graph; returns a callable replaying it.  The warm-up launch outside the capture
        runs on whatever self.input.buf holds; with update_stat measures it would widen the running ranges with that
        data, so the ranges are restored afterwards."""
        saved = [(qm, qm.running_min.clone(), qm.running_max.clone()) for qm, _i, _t in self.measures]
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            self.launch()                       # warm-up outside capture
        torch.cuda.current_stream().wait_stream(s)
        for qm, lo, hi in saved:
            qm.running_min.copy_(lo)
            qm.running_max.copy_(hi)
        with torch.cuda.graph(g):
            self.launch()
        self.graph = g
        return g.replay
