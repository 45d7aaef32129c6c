"""Two forwards of the static u8 plan of DeepLabV3+ at 16 x 513 x 513 (profiling target).
  ncu --set full --import-source on --clock-control none --kernel-name regex:pw_i8_kernel --launch-skip 5 --launch-count 5 \
      -o gpurun_out/r2_conv_q8 python tools/q8_deeplab_once.py"""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from helpers import dfq_quantised
from data_free_quantization_b200.eval_q8 import EvalPlanQ8
dev = "cuda:0"
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
with contextlib.redirect_stdout(io.StringIO()):
    model, graph, bottoms, targ, tr = dfq_quantised("deeplab", dev, np.load(os.path.join(ROOT, "tests/golden/deeplab_dfq.npz")))
for m in model.modules():
    if hasattr(m, "set_update_stat"):
        m.set_update_stat(False)
plan = EvalPlanQ8(model, tr, batch=B, in_shape=(3, 513, 513))
x = torch.randn(B, 3, 513, 513, device=dev)
plan(x); torch.cuda.synchronize()
plan(x); torch.cuda.synchronize()
