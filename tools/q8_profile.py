"""Per-op CUDA-event times of the static u8 plan (serialised) + graph replay time.
  python tools/q8_profile.py [batch] [mobilenetv2|deeplab]   (defaults: 256 mobilenetv2; deeplab runs 513 x 513)"""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from helpers import dfq_quantised
from data_free_quantization_b200.eval_q8 import EvalPlanQ8
dev = "cuda:0"
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
WHICH = sys.argv[2] if len(sys.argv) > 2 else "mobilenetv2"
SHAPE = (3, 513, 513) if WHICH == "deeplab" else (3, 224, 224)
with contextlib.redirect_stdout(io.StringIO()):
    model, graph, bottoms, targ, tr = dfq_quantised(WHICH, dev, np.load(os.path.join(ROOT, f"tests/golden/{WHICH}_dfq.npz")))
for m in model.modules():
    if hasattr(m, "set_update_stat"):
        m.set_update_stat(False)
plan = EvalPlanQ8(model, tr, batch=B, in_shape=SHAPE)
x = torch.randn(B, *SHAPE, device=dev)
plan(x); torch.cuda.synchronize()
tot = {}
rows = []
for name, fn in plan.ops:
    ts = []
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    ms = sorted(ts)[2]
    io_ = plan.step_io[name]
    kind = io_["kind"]
    tot[kind] = tot.get(kind, 0.0) + ms
    inb = sum(plan.codes[i].numel() for i in io_["in_slots"]) if kind != "stem" else plan.input.buf.numel() * 4
    if kind == "pad":
        inb = sum(plan.pre[i].numel() for i in io_["pre_slots"])
    outb = sum(int(np.prod(io_["out"].shape)) for _ in (io_["out_slots"] or [0]))
    rows.append((name, kind, ms, (inb + outb) / ms / 1e6, tuple(io_["out"].shape[1:]), len(io_["out_slots"])))
for r in sorted(rows, key=lambda r: -r[2])[:45]:
    print("%-28s %-6s %7.3f ms %8.1f GB/s out %s x%d" % r)
print({k: round(v, 3) for k, v in tot.items()}, "sum", round(sum(tot.values()), 3))
replay = plan.capture()
for _ in range(3): replay()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10): replay()
b.record(); b.synchronize()
ms = a.elapsed_time(b) / 10
print("graph replay %.3f ms per batch of %d = %.0f images/s; activation code bytes written %.1f MB/img" % (ms, B, B / ms * 1e3, plan.act_bytes / B / 1e6))
