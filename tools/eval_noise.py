"""Diagnostic: how far apart are (a) the cuDNN module path, (b) the static plan, (c) the reference golden logits,
and how sensitive is the quantised network to a 1-ulp perturbation of its input?"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import contextlib, io
from test_gpu_eval import _quantised_fresh
from data_free_quantization_b200.eval_plan import EvalPlan
from data_free_quantization_b200.utils.layer_transform import replace_op, restore_op
npz = np.load(os.path.join(ROOT, "tests/golden/mobilenetv2_dfq.npz"))
want = npz["eval/logits_b4"]
x = torch.randn(4, 3, 224, 224, generator=torch.Generator().manual_seed(2)).cuda()
def rel(a, b): return float(np.abs(a - b).max() / np.abs(b).max())
def module_run(xx):
    with contextlib.redirect_stdout(io.StringIO()):
        model, graph, targ, tr = _quantised_fresh(npz)
    replace_op()
    with torch.no_grad(): out = model(xx).cpu().numpy()
    restore_op()
    return out
def plan_run(xx):
    with contextlib.redirect_stdout(io.StringIO()):
        model, graph, targ, tr = _quantised_fresh(npz)
    return EvalPlan(model, tr, batch=4)(xx).cpu().numpy()
a = module_run(x); b = plan_run(x)
xp = x * (1 + 1e-7)
c = module_run(xp); d = plan_run(xp)
print("module vs golden", rel(a, want), "plan vs golden", rel(b, want), "module vs plan", rel(a, b))
print("module: x vs x*(1+1e-7)", rel(c, a), " plan: same", rel(d, b))
print("argmax golden", want.argmax(1), "module", a.argmax(1), "plan", b.argmax(1))
print("mean |err| / max|logit|: module", float(np.abs(a - want).mean() / np.abs(want).max()), "plan", float(np.abs(b - want).mean() / np.abs(want).max()))
print("logit range", float(want.min()), float(want.max()), "top1-top2 gaps", np.sort(want, 1)[:, -1] - np.sort(want, 1)[:, -2])
