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
replace_op()
with torch.no_grad(): model(x)
restore_op()
plan = EvalPlan(model2, tr2, batch=4)
got = plan.debug_run(x)
m1, m2 = dict(model.named_modules()), dict(model2.named_modules())
k = 0
for name in m1:
    if type(m1[name]) in targ:
        a, b = m1[name].quant, m2[name].quant
        print(name, "module", float(a.running_min), float(a.running_max), "plan", float(b.running_min), float(b.running_max))
        k += 1
        if k > 5: break
print(plan.slots[:8].cpu().numpy())
print([n for n, _ in plan.ops[:12]])
print("stats0", plan.stats[:3].cpu().numpy().view(np.uint32))
