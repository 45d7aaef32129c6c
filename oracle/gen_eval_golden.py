"""Generate the eval-parity goldens from the UNMODIFIED reference (build container only; needs /root/reference).

TEST INFRASTRUCTURE.  Usage:
    python -m oracle.gen_eval_golden --model mobilenetv2 [--what tf,top1,static] [--cle-cache /tmp/ref_cle_mobilenetv2.pt]
    python -m oracle.gen_eval_golden --model deeplab     [--what tf,top1]

The reference model is taken through main_dfq.py:188-231 exactly as oracle/gen_golden.py does (same seeds, the
same IEEE-sqrt shim), then its quantised eval forward (utils/quantize.py:106-123, 250-256, 358-363 with
replace_op() in force, main_dfq.py:96-103) is run with hooks on every target layer, every layer QuantMeasure and
every CustomTensorOP QuantMeasure (utils/layer_transform.py:186, 200-237).  Written:

  tests/golden/eval_tf_<model>.npz  (+ .json)   "teacher-forced" records of ONE forward:
      L/<layer>/xq     uint8   codes the reference's QuantMeasure produced for the layer's input (NCHW)
      L/<layer>/range  f32[2]  the (running_min, running_max) it quantised with (after the update_stat widening)
      L/<layer>/y      f32     the layer's output, every STRIDE-th element of the flattened tensor
      T/<i>/...                the same for the i-th patched tensor op (add / mean / interpolate / cat):
                               xq<j>, range<j> per operand, y = strided sample of the op's result
      (--tf-full: x = the raw fp32 input and the full output instead, for a small image)
  tests/golden/eval_top1_<model>.npz (+ .json)  argmax of the reference over >= 1024 images (4 x 513^2 for
      DeepLab), in batches as main_dfq.py:96-103 feeds them, the top-1 / top-2 margin, and the reference's own
      agreement with itself when the input is scaled by (1 + 1e-7) -- the noise floor of this random-init network.
  tests/golden/eval_static_<model>.npz          logits and argmax of the same model with set_update_stat(False)
      on every QuantMeasure (utils/quantize.py:125-126): the static-range W8A8 mode.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_harness as RH                               # noqa: E402
from data_free_quantization_b200.fixtures import synthetic_state_dict   # noqa: E402

STRIDE = 61


def build_run(ns, model_name, cle_cache):
    torch.manual_seed(0)
    if model_name == "deeplab":
        from modeling.segmentation.deeplab import DeepLab as RefDeepLab
        from data_free_quantization_b200.modeling.deeplab import DeepLab
        sd = synthetic_state_dict(DeepLab(), seed=0, bn_seed=1)
        model = RefDeepLab(sync_bn=False)
        missing, unexpected = model.load_state_dict(sd, strict=False)
        assert not unexpected and all("level_features" in k for k in missing)
        run = RH.RefRun(ns, model, torch.ones(1, 3, 129, 129))
    else:
        model = ns.mobilenet_v2(None)
        sd = synthetic_state_dict(model, seed=0, bn_seed=1)
        model.load_state_dict(sd)
        run = RH.RefRun(ns, model, torch.ones(4, 3, 224, 224))
    with RH.ieee_sqrt():
        run.merge_batchnorm()
    run.create_relation()
    if cle_cache and os.path.exists(cle_cache):
        cache = torch.load(cle_cache)
        with torch.no_grad():
            for k, m in run.graph.items():
                if isinstance(m, str):
                    continue
                n = run.names[k]
                for attr in ("weight", "bias", "fake_weight", "fake_bias"):
                    if f"{n}.{attr}" in cache["snap"]:
                        if attr == "bias" and m.bias is None:
                            m.bias = torch.nn.Parameter(cache["snap"][f"{n}.{attr}"].clone(), requires_grad=False)
                        else:
                            getattr(m, attr).copy_(cache["snap"][f"{n}.{attr}"])
        for r, S in zip(run.relations, cache["S"]):
            r.S = S
        print("CLE state from cache", cle_cache, flush=True)
    else:
        t = time.perf_counter()
        with RH.ieee_sqrt():
            run.equalize()
        dt = time.perf_counter() - t
        print(f"reference CLE {dt:.1f} s", flush=True)
        if cle_cache:
            torch.save({"init": "n/a", "snap": run.snapshot(), "S": [r.get_scale_vec() for r in run.relations],
                        "time": dt}, cle_cache)
    run.absorb(3)
    run.set_layer_bits(8, 8, 8)
    run.merge_batchnorm()
    run.quantize(8, 8)
    run.set_quant_minmax()
    run.clip(-15, 15)
    run.bias_correction_as_wired(8)
    return run


def measures_of(run):
    return [m for m in run.model.modules() if isinstance(m, run.ns.q.QuantMeasure)]


def save_ranges(run):
    return [(m.running_min.clone(), m.running_max.clone(), m.update_stat) for m in measures_of(run)]


def load_ranges(run, saved):
    for m, (lo, hi, us) in zip(measures_of(run), saved):
        m.running_min = lo.clone()
        m.running_max = hi.clone()
        m.update_stat = us


def reset_tensor_op_index(run):
    top = run.model.custom_tensor_op
    top.idx_tensor_op = 0
    top.idx_name_tensor_op = 0


def codes_of(xq, lo, hi):
    """uint8 codes of a fake-quantised tensor; asserts that code * scale + lo reproduces it bit for bit
    (utils/quantize.py:62-76: scale in double from the Python floats, cast to fp32; separate multiply and add)."""
    scale = np.float32(max((float(hi) - float(lo)) / 255.0, 1e-8))
    a = xq.numpy()
    q = np.rint((a.astype(np.float64) - float(lo)) / float(scale))
    assert q.min() >= 0 and q.max() <= 255
    back = (q.astype(np.float32) * scale).astype(np.float32) + np.float32(lo)
    assert back.tobytes() == a.tobytes(), "codes do not reproduce the reference's quantised tensor"
    return q.astype(np.uint8)


def teacher_forced(run, x, full=False):
    """One forward with hooks.  Returns (records dict for np.savez, meta, logits)."""
    ns = run.ns
    rec, meta = {}, {"layers": [], "tensor_ops": [], "stride": 1 if full else STRIDE}
    hooks = []
    top_qms = set(id(m) for m in run.model.custom_tensor_op.children())
    pending = []
    events = []

    def layer_hook(name):
        def fn(mod, inp, out):
            qm = mod.quant
            lo, hi = float(qm.running_min), float(qm.running_max)
            rec[f"L/{name}/range"] = np.array([lo, hi], np.float32)
            xq = qm._last_out
            if full:
                rec[f"L/{name}/x"] = inp[0].detach().numpy().copy()
                rec[f"L/{name}/y"] = out.detach().numpy().copy()
            else:
                rec[f"L/{name}/xq"] = codes_of(xq, lo, hi)
                rec[f"L/{name}/y"] = out.detach().numpy().ravel()[::STRIDE].copy()
            meta["layers"].append({"name": name, "in_shape": list(inp[0].shape), "out_shape": list(out.shape)})
        return fn

    def qm_hook(mod, inp, out):
        mod._last_out = out.detach()
        if id(mod) in top_qms:
            pending.append((inp[0].detach().clone(), out.detach().clone(), float(mod.running_min),
                            float(mod.running_max)))

    for n, m in run.targets():
        hooks.append(m.register_forward_hook(layer_hook(n)))
    for m in measures_of(run):
        hooks.append(m.register_forward_hook(qm_hook))

    # the patched ops call the raw op right after their operands went through the CustomTensorOP measures
    lt = ns.lt

    def wrap(table, key, kind):
        raw = table[key]

        def fn(*a, **kw):
            y = raw(*a, **kw)
            if pending:
                events.append((kind, list(pending), y.detach().clone()))
                pending.clear()
            return y
        table[key] = fn
        return raw

    run.fix_bn_eps()
    run.model.eval()
    reset_tensor_op_index(run)
    lt.replace_op()
    saved_raw = [(lt.raw_tensor_magic_op, "__add__", wrap(lt.raw_tensor_magic_op, "__add__", "add")),
                 (lt.raw_torch_op, "cat", wrap(lt.raw_torch_op, "cat", "cat")),
                 (lt.raw_torch_op, "mean", wrap(lt.raw_torch_op, "mean", "mean")),
                 (lt.raw_func_op, "interpolate", wrap(lt.raw_func_op, "interpolate", "interp"))]
    try:
        with torch.no_grad():
            logits = run.model(x)
    finally:
        for table, key, raw in saved_raw:
            table[key] = raw
        lt.restore_op()
        for h in hooks:
            h.remove()
    for i, (kind, ops, y) in enumerate(events):
        for j, (xin, xq, lo, hi) in enumerate(ops):
            rec[f"T/{i}/range{j}"] = np.array([lo, hi], np.float32)
            if full:
                rec[f"T/{i}/x{j}"] = xin.numpy()
            else:
                rec[f"T/{i}/xq{j}"] = codes_of(xq, lo, hi)
        rec[f"T/{i}/y"] = y.numpy().copy() if full else y.numpy().ravel()[::STRIDE].copy()
        meta["tensor_ops"].append({"kind": kind, "operands": len(ops), "in_shapes": [list(o[0].shape) for o in ops],
                                   "out_shape": list(y.shape)})
    return rec, meta, logits


def forward(run, x):
    reset_tensor_op_index(run)
    return run.forward_quantized(x)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="mobilenetv2", choices=["mobilenetv2", "deeplab"])
    ap.add_argument("--what", default="tf,top1,static")
    ap.add_argument("--cle-cache", default=None)
    ap.add_argument("--images", type=int, default=None)
    ap.add_argument("--threads", type=int, default=0)
    args = ap.parse_args()
    if args.threads:
        torch.set_num_threads(args.threads)
    what = set(args.what.split(","))
    cache = args.cle_cache or f"/tmp/ref_cle_{args.model}.pt"
    ns = RH.import_reference()
    run = build_run(ns, args.model, cache)
    fresh = save_ranges(run)
    gdir = os.path.join(ROOT, "tests", "golden")
    dl = args.model == "deeplab"

    if "tf" in what:
        g = torch.Generator().manual_seed(2)
        # the first image of the seeded batch the other goldens use
        x = (torch.randn(2, 3, 129, 129, generator=g) if dl else torch.randn(4, 3, 224, 224, generator=g))[:1].clone()
        load_ranges(run, fresh)
        rec, meta, logits = teacher_forced(run, x)
        meta["input"] = "torch.randn(seed 2)[:1]"
        meta["logits_absmax"] = float(logits.abs().max())
        rec["logits"] = logits.numpy()[:, :, ::4, ::4] if dl else logits.numpy()
        np.savez_compressed(os.path.join(gdir, f"eval_tf_{args.model}.npz"), **rec)
        json.dump(meta, open(os.path.join(gdir, f"eval_tf_{args.model}.json"), "w"), indent=1)
        print("teacher-forced records:", len(meta["layers"]), "layers,", len(meta["tensor_ops"]), "tensor ops", flush=True)

    if "tf_static" in what:
        # the same records with every measure in static mode (utils/quantize.py:125-126): the codes a layer's
        # QuantMeasure produces ARE what the producer of its input must emit in the static-range u8 pipeline
        g = torch.Generator().manual_seed(2)
        x = (torch.randn(2, 3, 129, 129, generator=g) if dl else torch.randn(4, 3, 224, 224, generator=g))[:1].clone()
        load_ranges(run, fresh)
        for m in measures_of(run):
            m.set_update_stat(False)
        rec, meta, logits = teacher_forced(run, x)
        meta["input"] = "torch.randn(seed 2)[:1]"
        meta["mode"] = "set_update_stat(False) on every QuantMeasure"
        rec["logits"] = logits.numpy()[:, :, ::4, ::4] if dl else logits.numpy()
        np.savez_compressed(os.path.join(gdir, f"eval_tf_static_{args.model}.npz"), **rec)
        json.dump(meta, open(os.path.join(gdir, f"eval_tf_static_{args.model}.json"), "w"), indent=1)
        load_ranges(run, fresh)
        print("static teacher-forced records:", len(meta["layers"]), "layers,", len(meta["tensor_ops"]), "tensor ops",
              flush=True)

    if "top1" in what:
        n_img = args.images or (4 if dl else 1024)
        batch = 2 if dl else 256                           # main_dfq.py:92 batch size 256 (seg: per-rank share)
        shape = (3, 513, 513) if dl else (3, 224, 224)
        out = {}
        info = {"images": n_img, "batch": batch, "input": "torch.randn(seed 5), batches in order",
                "perturbation": "x * (1 + 1e-7)"}
        res = []
        for pert in (False, True):
            load_ranges(run, fresh)
            g = torch.Generator().manual_seed(5)
            am, mg, first, t5 = [], [], None, []
            t = time.perf_counter()
            for b0 in range(0, n_img, batch):
                x = torch.randn(batch, *shape, generator=g)
                if pert:
                    x = x * (1.0 + 1e-7)
                lg = forward(run, x)
                top2 = lg.topk(2, dim=1).values
                am.append(lg.argmax(1).numpy())
                if not dl:
                    t5.append(lg.topk(5, dim=1).indices.numpy())
                mg.append((top2[:, 0] - top2[:, 1]).numpy() / float(lg.abs().max()))
                if first is None:
                    first = lg[:4].numpy().copy()
                print(f"  {'perturbed' if pert else 'plain'} {b0 + batch}/{n_img} {time.perf_counter() - t:.0f} s",
                      flush=True)
            res.append((np.concatenate(am), np.concatenate(mg), first, np.concatenate(t5) if t5 else None))
        (am0, mg0, lg0, t50), (am1, _, lg1, t51) = res
        if t50 is not None:
            # every image of this random-init network lands on the same class, so top-1 alone says little: keep
            # the ordered top-5 as well (and the reference's own top-5 stability under the perturbation)
            out["top5"] = t50.astype(np.int16)
            info["distinct_top1_classes"] = int(len(np.unique(am0)))
            info["reference_top5_set_self_agreement"] = float(np.mean([len(set(a) & set(b)) / 5.0 for a, b in zip(t50, t51)]))
            info["reference_top5_ordered_self_agreement"] = float((t50 == t51).all(1).mean())
        agree = float((am0 == am1).mean())
        info["reference_self_agreement_under_1e-7_input_perturbation"] = agree
        info["reference_logit_shift_under_perturbation"] = float(np.abs(lg0 - lg1).max() / np.abs(lg0).max())
        info["margin_median"] = float(np.median(mg0))
        info["margin_below_1e-3"] = float((mg0 < 1e-3).mean())
        out["argmax"] = am0.astype(np.uint8 if dl else np.int16)
        out["argmax_perturbed"] = am1.astype(np.uint8 if dl else np.int16)
        out["margin"] = mg0.astype(np.float16) if not dl else mg0[:, ::8, ::8].astype(np.float16)
        out["logits_first4"] = lg0 if not dl else lg0[:, :, ::8, ::8]
        np.savez_compressed(os.path.join(gdir, f"eval_top1_{args.model}.npz"), **out)
        json.dump(info, open(os.path.join(gdir, f"eval_top1_{args.model}.json"), "w"), indent=1)
        print("top-1 golden:", info, flush=True)

    if "static" in what:
        # static-range mode: utils/quantize.py:125-126 set_update_stat(False) on every measure, ranges as
        # set_quant_minmax left them (utils/layer_transform.py:356-618)
        load_ranges(run, fresh)
        for m in measures_of(run):
            m.set_update_stat(False)
        n_img = 2 if dl else 64
        shape = (3, 129, 129) if dl else (3, 224, 224)
        x = torch.randn(n_img, *shape, generator=torch.Generator().manual_seed(7))
        lg = forward(run, x)
        xp = forward(run, x * (1.0 + 1e-7))
        ranges = {}
        for n, m in run.targets():
            ranges[n] = [float(m.quant.running_min), float(m.quant.running_max)]
        top = [[float(m.running_min), float(m.running_max)] for m in run.model.custom_tensor_op.children()]
        info = {"images": n_img, "input": "torch.randn(seed 7)", "layer_ranges": ranges, "tensor_op_ranges": top,
                "reference_self_agreement_under_1e-7_input_perturbation": float((lg.argmax(1) == xp.argmax(1)).float().mean()),
                "reference_logit_shift_under_perturbation": float((lg - xp).abs().max() / lg.abs().max())}
        np.savez_compressed(os.path.join(gdir, f"eval_static_{args.model}.npz"),
                            logits=lg.numpy() if not dl else lg.numpy()[:, :, ::4, ::4],
                            argmax=lg.argmax(1).numpy().astype(np.uint8 if dl else np.int16))
        json.dump(info, open(os.path.join(gdir, f"eval_static_{args.model}.json"), "w"), indent=1)
        load_ranges(run, fresh)
        print("static golden:", {k: v for k, v in info.items() if not k.endswith("ranges")}, flush=True)


if __name__ == "__main__":
    main()
