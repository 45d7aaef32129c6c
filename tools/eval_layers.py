"""Diagnostic: per-layer difference between the cuDNN module path and the static plan (same model state)."""
import os, sys, contextlib, io
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_eval import _quantised_fresh
from data_free_quantization_b200.eval_plan import EvalPlan
from data_free_quantization_b200.utils.layer_transform import replace_op, restore_op
npz = np.load(os.path.join(ROOT, "tests/golden/mobilenetv2_dfq.npz"))
x = torch.randn(4, 3, 224, 224, generator=torch.Generator().manual_seed(2)).cuda()
with contextlib.redirect_stdout(io.StringIO()):
    model, graph, targ, tr = _quantised_fresh(npz)
    model2, graph2, targ2, tr2 = _quantised_fresh(npz)
ref, relu_after = {}, {}
mods = dict(model.named_modules())
for name, m in mods.items():
    if type(m) in targ:
        m.register_forward_hook(lambda mod, inp, out, name=name: ref.__setitem__(name, out.detach().clone()))
replace_op()
with torch.no_grad(): out_ref = model(x)
restore_op()
plan = EvalPlan(model2, tr2, batch=4)
got = plan.debug_run(x)
relu_fused = {st: True for st in got}
for name in ref:
    a = ref[name]
    b = got[name]
    a2 = torch.relu(a)
    e_raw = float((a - b).abs().max() / a.abs().max())
    e_relu = float((a2 - b).abs().max() / a.abs().max())
    e = min(e_raw, e_relu)
    d = (a2 if e_relu <= e_raw else a) - b
    nbad = int((d.abs() > 1e-4 * a.abs().max()).sum())
    print(f"{name:24s} shape {tuple(a.shape)} maxrel {e:.2e}  elems>1e-4: {nbad} / {a.numel()}")
