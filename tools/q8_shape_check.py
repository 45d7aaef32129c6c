"""Shape generality of the static u8 plan: odd image sizes / batch sizes against the reference-shaped module forward of the
same model in static mode (QuantMeasure.set_update_stat(False)): python tools/q8_shape_check.py"""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from helpers import dfq_quantised
from data_free_quantization_b200.eval_q8 import EvalPlanQ8
from data_free_quantization_b200.utils import quantize as Q
dev = "cuda:0"
for which, shapes in (("deeplab", [(3, 200, 264), (1, 97, 131), (2, 513, 513)]), ("mobilenetv2", [(7, 224, 224), (3, 160, 192)])):
    with contextlib.redirect_stdout(io.StringIO()):
        model, graph, bottoms, targ, tr = dfq_quantised(which, dev, np.load(os.path.join(ROOT, f"tests/golden/{which}_dfq.npz")))
    for m in model.modules():
        if hasattr(m, "set_update_stat"):
            m.set_update_stat(False)
    for n, h, w in shapes:
        x = torch.randn(n, 3, h, w, generator=torch.Generator().manual_seed(n + h)).to(dev)
        plan = EvalPlanQ8(model, tr, batch=n, in_shape=(3, h, w))
        got = plan(x).clone()
        Q.FUSED_EVAL = False
        with torch.no_grad():
            ref = model(x)
        Q.FUSED_EVAL = True
        shift = float((got - ref).abs().max() / ref.abs().max())
        agree = float((got.argmax(1) == ref.argmax(1)).float().mean())
        print(f"{which} {n}x3x{h}x{w}: logits shift {shift:.2e} of the largest, classes agree {agree:.4f}, launches {len(plan.ops)}")
