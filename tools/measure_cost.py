"""What the per-tensor measure launches cost inside the dynamic-range plans: the same fp32 plan captured with dynamic
ranges (a one-block measure kernel between producer and consumer) and with frozen ranges (no measure launches).
  python tools/measure_cost.py [batch of MobileNetV2 = 256] [batch of DeepLab = 16]"""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from helpers import dfq_quantised
from data_free_quantization_b200.eval_plan import EvalPlan
dev = "cuda:0"
def t(replay, n=10):
    for _ in range(3): replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): replay()
    b.record(); b.synchronize()
    return a.elapsed_time(b) / n
B_M = int(sys.argv[1]) if len(sys.argv) > 1 else 256
B_D = int(sys.argv[2]) if len(sys.argv) > 2 else 16
for which, B, shape in (("mobilenetv2", B_M, (3, 224, 224)), ("deeplab", B_D, (3, 513, 513))):
    with contextlib.redirect_stdout(io.StringIO()):
        model, graph, bottoms, targ, tr = dfq_quantised(which, dev, np.load(os.path.join(ROOT, f"tests/golden/{which}_dfq.npz")))
    x = torch.randn(B, *shape, device=dev)
    plan = EvalPlan(model, tr, batch=B, in_shape=shape); plan(x)
    d = t(plan.capture()); n_dyn = len(plan.ops)
    for m in model.modules():
        if hasattr(m, "set_update_stat"): m.set_update_stat(False)
    plan2 = EvalPlan(model, tr, batch=B, in_shape=shape); plan2(x)
    s = t(plan2.capture())
    print(f"{which} batch {B}: dynamic {d:.3f} ms ({n_dyn} ops), the same fp32 plan with frozen ranges {s:.3f} ms ({len(plan2.ops)} ops): measures cost {d - s:.3f} ms")
