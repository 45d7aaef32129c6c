"""CLI of the B200 DFQ pipeline -- same flags as the reference's main_dfq.py (:36-62), same step order
(:161-258).  Everything from merge_batchnorm on runs on the GPU through libdfq_b200.

Differences from the reference script, all forced by what is (not) available:
  * weights: the reference loads pretrained checkpoints that are not distributed
    (.MISSING_LARGE_BLOBS); without --weights this script uses seeded synthetic weights (fixtures.py);
  * data: `inference_all` (main_dfq.py:66-113) reads ./val (ImageNet) / ./VOCdevkit; here --val DIR points at an
    ImageFolder (same transforms, batch 256), otherwise --batches synthetic batches are run and the accuracy line says
    so.  The forward of a --quantize run goes through the static eval plan (eval_plan.EvalPlan, or eval_q8.EvalPlanQ8
    with --static: every QuantMeasure in set_update_stat(False) mode); under torchrun every batch is split across
    the ranks (dist.inference_all_sharded);
  * --task seg builds DeepLabV3+ (MobileNetV2 backbone, main_dfq.py:131-137) on 513 x 513 inputs;
  * `--correction` does not raise: the reference passes an unknown kwarg (main_dfq.py:231 -> TypeError).
    --correction_mode ref (default) reproduces the reference function as wired (a no-op on id-keyed graphs),
    --correction_mode paper applies b -= eps.E[x].
"""
import argparse
import time

import torch
import torch.nn as nn

from .bias_absorption import bias_absorption
from .bias_correction import bias_correction
from .clip_weight import clip_weight
from .Cross_layer_equal import cross_layer_equalization
from .fixtures import synthetic_state_dict
from .graph import TorchTransformer
from .modeling.mobilenetv2 import mobilenet_v2
from .utils.layer_transform import (merge_batchnorm, quantize_targ_layer, replace_op, restore_op,
                                    set_quant_minmax, switch_layers)
from .utils.quantize import QuantConv2d, QuantLinear, QuantMeasure, set_layer_bits
from .utils.relation import create_relation


def get_argument(argv=None):
    parser = argparse.ArgumentParser()
    parser.add_argument("--gpu", action='store_true')
    parser.add_argument("--sizedisp", action='store_true')
    parser.add_argument("--visualize", action='store_true')
    parser.add_argument("--quantize", action='store_true')
    parser.add_argument("--equalize", action='store_true')
    parser.add_argument("--correction", action='store_true')
    parser.add_argument("--absorption", action='store_true')
    parser.add_argument("--relu", action='store_true')
    parser.add_argument("--clip_weight", action='store_true')
    parser.add_argument("--task", default='cls', type=str, choices=['cls', 'seg'])
    parser.add_argument("--resnet", action='store_true')
    parser.add_argument("--log", action='store_true')
    parser.add_argument("--bits_weight", type=int, default=8)
    parser.add_argument("--bits_activation", type=int, default=8)
    parser.add_argument("--bits_bias", type=int, default=8)
    # additions (not in the reference)
    parser.add_argument("--weights", default=None, help="checkpoint path; default: seeded synthetic weights")
    parser.add_argument("--batch", type=int, default=None, help="eval batch (default 256 cls / 16 seg, main_dfq.py:79,86)")
    parser.add_argument("--batches", type=int, default=2, help="synthetic eval batches when --val is not given")
    parser.add_argument("--val", default=None, help="ImageFolder root for --task cls (the reference reads ./val)")
    parser.add_argument("--static", action="store_true",
                        help="static activation ranges (QuantMeasure.set_update_stat(False)) on the u8 pipeline")
    parser.add_argument("--correction_mode", default="ref", choices=["ref", "paper"])
    return parser.parse_args(argv)


def run_dfq_steps(model, graph, bottoms, args, targ_pair=(QuantConv2d, QuantLinear), timings=None):
    """main_dfq.py:188-231 on an already traced model (graph / bottoms from switch_layers).
    Returns the relations.

    Without `timings`, the second pass over the same traced model in the same fresh state is recorded into CUDA
    graphs and later passes replay them (pipeline_plan.py); results and side effects are those of the eager
    functions below."""
    from . import pipeline_plan as PP

    def timed(name, fn):
        if timings is None:
            return fn()
        torch.cuda.synchronize()
        t = time.perf_counter()
        out = fn()
        torch.cuda.synchronize()
        timings[name] = timings.get(name, 0.0) + (time.perf_counter() - t) * 1e3
        return out

    targ_layer = tuple(targ_pair) if args.quantize else (nn.Conv2d, nn.Linear)

    def weights_phase():
        """merge_batchnorm .. quantize_targ_layer (main_dfq.py:188-214)."""
        timed("merge_batchnorm", lambda: merge_batchnorm(model, graph, bottoms, targ_layer))
        res = []
        if args.equalize:
            res = create_relation(graph, bottoms, targ_layer, delete_single=False)
            info = timed("cross_layer_equalization",
                         lambda: cross_layer_equalization(graph, res, targ_layer, Save_state=False, Treshhold=2e-7,
                                                          return_info=timings is not None))
            if timings is not None and info:
                timings["cle_kernel"] = info["kernel_ms"]
                timings["cle_iters"] = info["iters"]
                timings["cle_impl"] = info["impl"]
        if args.absorption:
            timed("bias_absorption", lambda: bias_absorption(graph, res, bottoms, N=3, visualize=args.visualize))
        if args.quantize:
            timed("set_layer_bits", lambda: set_layer_bits(graph, args.bits_weight, args.bits_activation,
                                                           args.bits_bias, targ_layer))
            timed("merge_batchnorm", lambda: merge_batchnorm(model, graph, bottoms, targ_layer))
            timed("quantize_targ_layer",
                  lambda: quantize_targ_layer(graph, args.bits_weight, args.bits_bias, targ_layer))
        return res

    def clip_phase():
        if args.clip_weight:
            timed("clip_weight", lambda: clip_weight(graph, range_clip=[-15, 15], targ_type=targ_layer))

    def correction():
        if args.correction:
            timed("bias_correction", lambda: bias_correction(graph, bottoms, targ_layer, bits_weight=args.bits_weight,
                                                             visualize=args.visualize,
                                                             mode=getattr(args, "correction_mode", "ref")))

    mode, plan = ("eager", None) if timings is not None else PP.lookup(graph, bottoms, args, targ_layer)
    if mode == "capture":
        snap = PP._snapshot(graph)
        try:
            plan = PP.PipelinePlan(model, graph, bottoms, args, targ_layer,
                                   lambda phase: weights_phase() if phase == 1 else clip_phase())
            PP.store(graph, plan)
            PP.stats["captured"] += 1
            mode = "replay"
        except Exception as exc:                 # capture refused (e.g. a driver without cooperative graph nodes)
            PP._restore(snap)
            PP.store(graph, ("failed", repr(exc)))
            mode = "eager"
    if mode == "replay":
        # bias_correction as wired in the reference touches nothing on the device: its host work (and
        # its log lines) run while the graphs execute, and the wait for set_quant_minmax's statistics comes last.
        # The other modes launch kernels on the final weights: the reference's order.
        host_only = args.correction and getattr(args, "correction_mode", "ref") == "ref"
        res = plan.replay(finish=not host_only)
        PP.stats["replayed"] += 1
        correction()
        if host_only:
            plan.finish_ranges()
        return res
    PP.stats["eager"] += 1
    res = weights_phase()
    if args.quantize:
        timed("set_quant_minmax", lambda: set_quant_minmax(graph, bottoms))
    clip_phase()
    correction()
    return res


def run_dfq(model, data, args, targ_pair=(QuantConv2d, QuantLinear), timings=None):
    """main_dfq.py:148-231 on an already constructed CUDA model.  Returns (model, graph, bottoms, relations)."""
    transformer = TorchTransformer()
    module_dict = {}
    if args.quantize:
        module_dict[1] = [(nn.Conv2d, targ_pair[0]), (nn.Linear, targ_pair[1])]
    if args.relu:
        module_dict[0] = [(torch.nn.ReLU6, torch.nn.ReLU)]
    model, transformer = switch_layers(model, transformer, data, module_dict, ignore_layer=[QuantMeasure],
                                       quant_op=args.quantize)
    graph = transformer.log.getGraph()
    bottoms = transformer.log.getBottoms()
    res = run_dfq_steps(model, graph, bottoms, args, targ_pair, timings)
    run_dfq.last_transformer = transformer
    return model, graph, bottoms, res


run_dfq.last_transformer = None


def synthetic_batches(n_batches, batch, shape, seed=2):
    g = torch.Generator().manual_seed(seed)
    for _ in range(n_batches):
        yield torch.randn(batch, *shape, generator=g), None


def imagefolder_batches(root, batch):
    """main_dfq.py:70-79: ImageNet validation folder, Resize(256) / CenterCrop(224) / Normalize, batch 256."""
    from torch.utils.data import DataLoader
    from torchvision import datasets, transforms
    ds = datasets.ImageFolder(root, transforms.Compose([
        transforms.Resize(256), transforms.CenterCrop(224), transforms.ToTensor(),
        transforms.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])]))
    for image, label in DataLoader(ds, batch_size=batch, shuffle=False, num_workers=4, pin_memory=True):
        yield image, label


def inference_all(model, transformer, task, batches, batch, in_shape, static=False, quantized=True):
    """main_dfq.py:66-113.  Every batch is split across the ranks of the default process group (one process per GPU);
    the forward of a quantised model is the static eval plan -- one launch list of libdfq_b200 kernels per shard
    size, no cuDNN, no per-layer host synchronisation.  Returns (accuracy or None, images, seconds)."""
    from .dist import rank_world, shard_range
    import torch.distributed as dist
    print("Start inference")
    rank, world = rank_world()
    dev = next(model.parameters()).device
    plans = {}

    def forward(x):
        if not quantized:
            with torch.no_grad():
                return model(x.to(dev))
        n = x.shape[0]
        if n not in plans:
            if static:
                from .eval_q8 import EvalPlanQ8
                plans[n] = EvalPlanQ8(model, transformer, batch=n, in_shape=in_shape)
            else:
                from .eval_plan import EvalPlan
                plans[n] = EvalPlan(model, transformer, batch=n, in_shape=in_shape)
        return plans[n](x.to(dev, non_blocking=True)).view(n, -1) if task == 'cls' else plans[n](x.to(dev)).view(
            n, -1, in_shape[1], in_shape[2])

    correct = total = labelled = 0
    t0 = time.time()
    for ii, (x, label) in enumerate(batches):
        a, b = shard_range(x.shape[0], rank, world)
        if b > a:
            logits = forward(x[a:b])
            pred = logits.argmax(1)
            total += b - a
            if label is not None:
                correct += int((pred.cpu() == label[a:b]).sum())
                labelled += b - a
                print(f"Batch {ii + 1}: Accuracy: {correct / max(1, labelled) * 100:.2f}%")
    torch.cuda.synchronize()
    t = torch.tensor([correct, labelled, total], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(t)
    correct, labelled, total = (int(v) for v in t.tolist())
    acc = correct / labelled if labelled else None
    if acc is not None:
        print(f"Final Accuracy: {acc * 100:.2f}%")
    return acc, total, time.time() - t0


def main(argv=None):
    args = get_argument(argv)
    assert args.relu or args.relu == args.equalize, 'must replace relu6 to relu while equalization'
    assert args.equalize or args.absorption == args.equalize, 'must use absorption with equalize'
    if args.resnet:
        raise SystemExit("--resnet: the reference loads pytorchcv's resnet18, which is not available offline")
    if not torch.cuda.is_available():
        raise SystemExit("no CUDA device: the DFQ pipeline has no CPU path")
    import os
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 and not dist.is_initialized():
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group("nccl")
    dev = torch.device("cuda", torch.cuda.current_device())
    if args.task == 'seg':                                   # main_dfq.py:131-137
        from .modeling.deeplab import DeepLab
        torch.manual_seed(0)
        model = DeepLab()
        if args.weights:
            model.load_state_dict(torch.load(args.weights, map_location="cpu"))
        in_shape, batch = (3, 513, 513), args.batch or 16
    else:
        model = mobilenet_v2(args.weights)
        in_shape, batch = (3, 224, 224), args.batch or 256
    if args.weights is None:
        model.load_state_dict(synthetic_state_dict(model, seed=0, bn_seed=1))
    model = model.to(dev).eval()
    data = torch.ones((4,) + in_shape, device=dev) if args.task == 'cls' else torch.ones((1, 3, 129, 129), device=dev)
    timings = {}
    model, graph, bottoms, _ = run_dfq(model, data, args, timings=timings)
    for k, v in timings.items():
        if isinstance(v, float):
            print(f"{k}: {v:.3f} ms")
    print(f"DFQ pipeline total: {sum(v for v in timings.values() if isinstance(v, float)):.3f} ms")
    model.eval()
    if args.static:
        if not args.quantize:
            raise SystemExit("--static needs --quantize (the u8 pipeline runs the quantised eval forward)")
        for m in model.modules():
            if hasattr(m, "set_update_stat"):
                m.set_update_stat(False)
    transformer = run_dfq.last_transformer
    if args.val and args.task == 'cls':
        batches = imagefolder_batches(args.val, batch)
    else:
        batches = synthetic_batches(args.batches, batch, in_shape)
    accuracy, images, seconds = inference_all(model, transformer, args.task, batches, batch, in_shape,
                                              static=args.static, quantized=args.quantize)
    print(f"Inference time is {seconds} seconds ({images} images, {images / max(seconds, 1e-9):.1f} images/s"
          f"{', synthetic data: no accuracy' if accuracy is None else ''})")
    if args.log and (not dist.is_initialized() or dist.get_rank() == 0):
        with open("dfq_result.txt", 'a+') as ww:     # main_dfq.py:262-267, format unchanged
            ww.write("task: {}, resnet: {}, relu: {}, equalize: {}, absorption: {}, quantize: {}, correction: {}, "
                     "clip: {}, bits_weight: {}, bits_activation: {}, bits_bias: {}\n".format(
                         args.task, args.resnet, args.relu, args.equalize, args.absorption, args.quantize,
                         args.correction, args.clip_weight, args.bits_weight, args.bits_activation, args.bits_bias))
            ww.write("Accuracy: {} %\n\n".format(float('nan') if accuracy is None else accuracy * 100))
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
