"""Two forwards of the static u8 plan at batch 256 (profiling target).  python tools/q8_once.py"""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from helpers import dfq_quantised
from data_free_quantization_b200.eval_q8 import EvalPlanQ8
dev = "cuda:0"
with contextlib.redirect_stdout(io.StringIO()):
    model, graph, bottoms, targ, tr = dfq_quantised("mobilenetv2", dev, np.load(os.path.join(ROOT, "tests/golden/mobilenetv2_dfq.npz")))
for m in model.modules():
    if hasattr(m, "set_update_stat"):
        m.set_update_stat(False)
plan = EvalPlanQ8(model, tr, batch=256)
x = torch.randn(256, 3, 224, 224, device=dev)
plan(x); torch.cuda.synchronize()
plan(x); torch.cuda.synchronize()
print("ok")
