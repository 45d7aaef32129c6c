"""Two forwards of the dynamic-range plan (profiling target): python tools/dyn_once.py [mobilenetv2|deeplab] [batch]
  ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,\
dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/dyn.csv python tools/dyn_once.py"""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from helpers import dfq_quantised
from data_free_quantization_b200.eval_plan import EvalPlan
dev = "cuda:0"
which = sys.argv[1] if len(sys.argv) > 1 else "mobilenetv2"
B = int(sys.argv[2]) if len(sys.argv) > 2 else (256 if which == "mobilenetv2" else 16)
shape = (3, 224, 224) if which == "mobilenetv2" else (3, 513, 513)
with contextlib.redirect_stdout(io.StringIO()):
    model, graph, bottoms, targ, tr = dfq_quantised(which, dev, np.load(os.path.join(ROOT, f"tests/golden/{which}_dfq.npz")))
plan = EvalPlan(model, tr, batch=B, in_shape=shape)
x = torch.randn(B, *shape, device=dev)
plan(x); torch.cuda.synchronize()
torch.cuda.nvtx.range_push("second")
plan(x); torch.cuda.synchronize()
torch.cuda.nvtx.range_pop()
print("ops", len(plan.ops))
