"""DeepLabV3+ (MobileNetV2) W8A8 eval forward at 513 x 513: fused kernels (dense k x k and pointwise on tcgen05,
depthwise on CUDA cores) against the two-step form (QuantMeasure kernel + cuDNN fp32).  python tools/deeplab_eval_time.py [batch]"""
import contextlib, io, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.nn as nn


def main():
    from data_free_quantization_b200 import graph as G
    from data_free_quantization_b200.Cross_layer_equal import cross_layer_equalization
    from data_free_quantization_b200.bias_absorption import bias_absorption
    from data_free_quantization_b200.clip_weight import clip_weight
    from data_free_quantization_b200.fixtures import synthetic_state_dict
    from data_free_quantization_b200.modeling.deeplab import DeepLab
    from data_free_quantization_b200.utils import quantize as Q
    from data_free_quantization_b200.utils.layer_transform import (merge_batchnorm, quantize_targ_layer, replace_op,
                                                                   restore_op, set_quant_minmax)
    from data_free_quantization_b200.utils.relation import create_relation
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    dev = "cuda:0"
    torch.manual_seed(0)
    model = DeepLab()
    model.load_state_dict(synthetic_state_dict(model, seed=0, bn_seed=1))
    model = model.eval().to(dev)
    targ = [Q.QuantNConv2d, Q.QuantNLinear]
    md = {0: [(nn.ReLU6, nn.ReLU)], 1: [(nn.Conv2d, targ[0]), (nn.Linear, targ[1])]}
    with contextlib.redirect_stdout(io.StringIO()):
        model, tr = G.switch_layers(model, G.TorchTransformer(), torch.ones(1, 3, 129, 129, device=dev), md,
                                    ignore_layer=[Q.QuantMeasure], quant_op=True)
        graph, bottoms = tr.log.getGraph(), tr.log.getBottoms()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        merge_batchnorm(model, graph, bottoms, targ)
        rels = create_relation(graph, bottoms, targ)
        cross_layer_equalization(graph, rels, targ, Save_state=False, Treshhold=2e-7)
        bias_absorption(graph, rels, bottoms, N=3)
        Q.set_layer_bits(graph, 8, 8, 8, targ)
        merge_batchnorm(model, graph, bottoms, targ)
        quantize_targ_layer(graph, 8, 8, targ)
        set_quant_minmax(graph, bottoms, verbose=False)
        clip_weight(graph, [-15, 15], targ)
        torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"DeepLab DFQ pipeline (merge..clip, incl. create_relation): {(t1 - t0) * 1e3:.1f} ms")
    model.eval()
    x = torch.randn(batch, 3, 513, 513, generator=torch.Generator().manual_seed(2)).to(dev)
    replace_op()
    try:
        for fused in (True, False):
            Q.FUSED_EVAL = fused
            with torch.no_grad():
                for _ in range(2):
                    y = model(x)
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(5):
                    y = model(x)
                b.record(); b.synchronize()
            ms = a.elapsed_time(b) / 5
            print(f"fused={fused}: {ms:.2f} ms per batch of {batch} -> {batch / ms * 1e3:.1f} images/s, "
                  f"logits {tuple(y.shape)}")
    finally:
        restore_op()
        Q.FUSED_EVAL = True
    from data_free_quantization_b200.eval_plan import EvalPlan
    outs = {}
    for move in ("torch", "native"):
        plan = EvalPlan(model, tr, batch=batch, in_shape=(3, 513, 513), move_impl=move)
        outs[move] = plan(x).clone()
        replay = plan.capture()
        for _ in range(2):
            replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            replay()
        b.record(); b.synchronize()
        ms = a.elapsed_time(b) / 5
        print(f"static plan (pad/pool/interp/cat: {move}), one CUDA graph ({len(plan.ops)} launches-or-ops): "
              f"{ms:.2f} ms per batch of {batch} -> {batch / ms * 1e3:.1f} images/s")
        by, byop = {}, {}
        for name, fn in plan.ops:
            kind = plan.op_info.get(name, {}).get("kind", name.split("[")[0])
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); e1.synchronize()
            t = e0.elapsed_time(e1)
            by[kind] = by.get(kind, 0.0) + t
            if name in plan.op_info:
                byop[name] = (kind, round(t, 4), round(plan.op_info[name]["bytes"] / t / 1e6, 1),
                              plan.op_info[name]["shape"])
        print("by kind (ms, serialised):", {k: round(v, 3) for k, v in sorted(by.items(), key=lambda kv: -kv[1])})
        if move == "native":
            print("per op (kind, ms, algorithmic GB/s, shape), slowest first:")
            for name, v in sorted(byop.items(), key=lambda kv: -kv[1][1])[:int(os.environ.get("DFQ_TOP", "40"))]:
                print(f"  {name:40s} {v[0]:7s} {v[1]:8.4f} {v[2]:8.1f}  {v[3]}")
        del plan, replay
    d = (outs["native"] - outs["torch"]).abs().max().item() / outs["torch"].abs().max().item()
    same = (outs["native"].view(batch, -1, 513 * 513).argmax(1) == outs["torch"].view(batch, -1, 513 * 513).argmax(1))
    print(f"native vs torch movement ops: max |dlogit| / max |logit| = {d:.3g}, same class on "
          f"{same.float().mean().item() * 100:.4f} % of pixels")

if __name__ == "__main__":
    main()
