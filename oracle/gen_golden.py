"""Generate tests/golden/* from the UNMODIFIED reference and pin the C oracle to it.

Build container only (needs /root/reference).  Usage:
    python -m oracle.gen_golden [--skip-eval]
For every DFQ step it runs the reference (oracle/ref_harness.py) and the restatement
(oracle/pipeline.py) on the same seeded state, asserts parity, and writes
  tests/golden/mobilenetv2_dfq.json   digests, iteration counts, parity verdicts
  tests/golden/mobilenetv2_dfq.npz    small tensors (S, biases, BN stats, samples, logits)
"""
import argparse
import hashlib
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import pipeline as P                                   # noqa: E402
from oracle import ref_harness as RH                               # noqa: E402
from data_free_quantization_b200.fixtures import synthetic_state_dict   # noqa: E402

SAMPLE_STRIDE = 997


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def to_np_state(sd):
    return {k: v.detach().cpu().numpy().copy() for k, v in sd.items() if v.dtype == torch.float32}


def compare(tag, ref_snap, st, verdict, exact=True, rtol=1e-5):
    worst = 0.0
    for k, v in ref_snap.items():
        a, b = v.numpy(), st[k]
        if exact:
            same = a.tobytes() == b.tobytes()
            if not same:
                nbad = int((a != b).sum())
                raise AssertionError(f"{tag}: {k} differs in {nbad}/{a.size} elements")
        else:
            # "relative" for reduction outputs: against max(|value|, 1 % of the tensor's largest
            # magnitude) -- fp32 summation error scales with sum|terms|, not with a cancelled result
            denom = np.maximum(np.abs(a), 1e-2 * max(float(np.max(np.abs(a))), 1e-30))
            worst = max(worst, float(np.max(np.abs(a - b) / denom)))
            if worst > rtol:
                raise AssertionError(f"{tag}: {k} rel err {worst}")
    verdict[tag] = "bit-exact" if exact else f"max rel {worst:.3e}"
    print(f"[parity] {tag}: {verdict[tag]}", flush=True)


def digest(snap):
    return {k: sha(v.numpy() if hasattr(v, "numpy") else v) for k, v in snap.items()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--skip-eval", action="store_true")
    ap.add_argument("--bits", type=int, default=8)
    ap.add_argument("--cle-cache", default=None, help="cache of the reference's post-CLE state (dev aid)")
    ap.add_argument("--model", default="mobilenetv2", choices=["mobilenetv2", "deeplab"])
    ap.add_argument("--tag", default="", help="suffix of the output files (bit-width sweep: --bits 4 --tag _w4)")
    args = ap.parse_args()
    ns = RH.import_reference()
    torch.manual_seed(0)
    if args.model == "deeplab":
        # BASELINE.json configs[3].  The reference model registers the backbone twice more under aliasing names
        # (backbone/mobilenet.py:113-114), so the seeded state comes from our alias-free twin (same names, same
        # shapes, bit-identical forward) and is loaded non-strictly; only alias keys may be missing.
        from modeling.segmentation.deeplab import DeepLab as RefDeepLab
        from data_free_quantization_b200.modeling.deeplab import DeepLab
        sd = synthetic_state_dict(DeepLab(), seed=0, bn_seed=1)
        model = RefDeepLab(sync_bn=False)
        missing, unexpected = model.load_state_dict(sd, strict=False)
        assert not unexpected and all("level_features" in k for k in missing), (missing[:3], unexpected[:3])
        run = RH.RefRun(ns, model, torch.ones(1, 3, 129, 129))    # the graph does not depend on the trace size
    else:
        model = ns.mobilenet_v2(None)
        sd = synthetic_state_dict(model, seed=0, bn_seed=1)
        model.load_state_dict(sd)
        run = RH.RefRun(ns, model, torch.ones(4, 3, 224, 224))
    out_stem = {"mobilenetv2": "mobilenetv2_dfq", "deeplab": "deeplab_dfq"}[args.model] + args.tag
    if args.tag:
        args.skip_eval = True

    gold = {"model": args.model, "seed": 0, "bn_seed": 1, "torch": torch.__version__,
            "parity": {}, "steps": {}, "timing_s": {}}
    small = {}
    st = to_np_state(sd)
    gold["init_digest"] = sha(np.concatenate([v.ravel() for v in st.values()]))
    gold["graph"] = run.graph_names()
    gold["bottoms"] = run.bottoms_names()

    with RH.ieee_sqrt():
        run.merge_batchnorm()
    rels = run.create_relation()
    bn_of = {}
    bot = gold["bottoms"]
    gmods = dict(zip(gold["graph"], run.graph.values()))
    for name, m in gmods.items():
        if isinstance(m, torch.nn.BatchNorm2d) and bot[name]:
            if type(gmods[bot[name][0]]) in run.targ_layer:
                bn_of[bot[name][0]] = name

    def relu_between(first, second):
        idx = second
        while idx != first:
            if isinstance(gmods[bot[idx][0]], torch.nn.ReLU):
                return True
            idx = bot[idx][0]
        return False

    topo = {"targets": [n for n, _ in run.targets()], "conv_bn": bn_of,
            "relations": [(a, b, c, relu_between(a, b)) for a, b, c in rels]}
    gold["topo"] = topo
    P.merge_batchnorm(st, topo)
    compare("merge_batchnorm", run.snapshot(), st, gold["parity"])
    gold["steps"]["merge_batchnorm"] = digest(run.snapshot())

    t = time.perf_counter()
    if args.cle_cache and os.path.exists(args.cle_cache):
        cache = torch.load(args.cle_cache)
        assert cache["init"] == gold["init_digest"]
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
        gold["timing_s"]["cross_layer_equalization"] = cache["time"]
    else:
        with RH.ieee_sqrt():
            run.equalize()
        gold["timing_s"]["cross_layer_equalization"] = time.perf_counter() - t
        if args.cle_cache:
            torch.save({"init": gold["init_digest"], "snap": run.snapshot(),
                        "S": [r.get_scale_vec() for r in run.relations],
                        "time": gold["timing_s"]["cross_layer_equalization"]}, args.cle_cache)
    t = time.perf_counter()
    iters, diffs, S_acc = P.cross_layer_equalization(st, topo)
    gold["timing_s"]["oracle_cross_layer_equalization"] = time.perf_counter() - t
    gold["cle_iters"], gold["cle_diffs"] = iters, diffs
    compare("cross_layer_equalization", run.snapshot(), st, gold["parity"])
    for r, (a, b, _c) in zip(run.relations, rels):
        s_ref = r.get_scale_vec().numpy()
        assert s_ref.tobytes() == S_acc[(a, b)].tobytes(), f"S mismatch {a}->{b}"
        small[f"S/{a}->{b}"] = s_ref
    gold["parity"]["relation_S"] = "bit-exact"
    snap = run.snapshot()
    gold["steps"]["cross_layer_equalization"] = digest(snap)

    t = time.perf_counter()
    run.absorb(3)
    gold["timing_s"]["bias_absorption"] = time.perf_counter() - t
    P.bias_absorption(st, topo, 3)
    snap = run.snapshot()
    compare("bias_absorption", snap, st, gold["parity"], exact=False, rtol=1e-5)
    # carry the reference's biases forward so later bit-exact checks are not polluted by
    # the (allowed) summation-order difference of the absorption GEMV
    for k, v in snap.items():
        st[k] = v.numpy().copy()
        if k.endswith(("bias", "fake_bias", "fake_weight")):
            small[f"absorb/{k}"] = v.numpy()
    gold["steps"]["bias_absorption"] = digest(snap)

    run.set_layer_bits(args.bits, 8, 8)
    run.merge_batchnorm()                      # main_dfq.py:211 -- resets fake_weight/fake_bias
    t = time.perf_counter()
    run.quantize(args.bits, 8)
    gold["timing_s"]["quantize_targ_layer"] = time.perf_counter() - t
    for bn in set(bn_of.values()):
        st[bn + ".fake_weight"].fill(1.0)
        st[bn + ".fake_bias"].fill(0.0)
    codes = P.quantize_targ_layer(st, topo, args.bits, 8, want_codes=True)
    snap = run.snapshot()
    compare("quantize_targ_layer", snap, st, gold["parity"])
    gold["steps"]["quantize_targ_layer"] = digest(snap)
    gold["codes"] = {k: sha(v.astype(np.uint8)) for k, v in codes.items()}
    for k, v in snap.items():
        small[f"quant/{k}"] = v.numpy() if k.endswith("bias") else v.numpy().ravel()[::SAMPLE_STRIDE]

    run.set_quant_minmax()
    qm = {}
    for n, m in run.targets():
        qm[n] = [float(m.quant.running_min), float(m.quant.running_max)]
    gold["quant_minmax"] = qm
    t = time.perf_counter()
    run.clip(-15, 15)
    gold["timing_s"]["clip_weight"] = time.perf_counter() - t
    P.clip_weight(st, topo)
    compare("clip_weight", run.snapshot(), st, gold["parity"])
    before, after = run.bias_correction_as_wired(args.bits)
    gold["bias_correction_as_wired"] = [len(before), len(after)]
    compare("bias_correction(as wired: no-op)", run.snapshot(), st, gold["parity"])
    gold["steps"]["final"] = digest(run.snapshot())

    if not args.skip_eval:
        g = torch.Generator().manual_seed(2)
        # DeepLab: 2 x 129 x 129 keeps the fixture small; the layer shapes that matter (dilated depthwise, atrous
        # 3x3, F.pad, bilinear interpolate, cat) do not depend on the image size
        x = torch.randn(2, 3, 129, 129, generator=g) if args.model == "deeplab" else \
            torch.randn(4, 3, 224, 224, generator=g)
        # strided samples of the first target layers' outputs: the quantised random-init network amplifies
        # rounding noise chaotically, so tight parity is checked near the input
        taps, hooks = {}, []
        for n, m in run.targets()[:6]:
            hooks.append(m.register_forward_hook(
                lambda mod, inp, out, n=n: taps.__setitem__(n, out.detach().clone())))
        t = time.perf_counter()
        logits = run.forward_quantized(x)
        gold["timing_s"]["eval_b4"] = time.perf_counter() - t
        for h in hooks:
            h.remove()
        for n, v in taps.items():
            small[f"eval/tap/{n}"] = v.numpy().ravel()[::SAMPLE_STRIDE]
        if args.model == "deeplab":
            small["eval/logits_b2"] = logits.numpy()[:, :, ::4, ::4]
            small["eval/argmax_b2"] = logits.argmax(1).numpy().astype(np.uint8)
        else:
            small["eval/logits_b4"] = logits.numpy()
        acts = {}
        for n, m in run.targets():
            acts[n] = [float(m.quant.running_min), float(m.quant.running_max)]
        gold["quant_minmax_after_eval"] = acts

    os.makedirs(os.path.join(ROOT, "tests", "golden"), exist_ok=True)
    with open(os.path.join(ROOT, "tests", "golden", out_stem + ".json"), "w") as f:
        json.dump(gold, f, indent=1)
    if args.tag:            # sweep files keep only what differs from the W8 golden: the quantised state
        gold = {k: gold[k] for k in ("model", "seed", "bn_seed", "torch", "parity", "codes", "quant_minmax")} | {
            "bits": args.bits, "steps": {k: gold["steps"][k] for k in ("quantize_targ_layer", "final")}}
        small = {k: v for k, v in small.items() if k.startswith("quant/") and k.endswith("bias")}
        with open(os.path.join(ROOT, "tests", "golden", out_stem + ".json"), "w") as f:
            json.dump(gold, f, indent=1)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", out_stem + ".npz"), **small)
    print("golden written", {k: round(v, 3) for k, v in gold["timing_s"].items()})


if __name__ == "__main__":
    main()
