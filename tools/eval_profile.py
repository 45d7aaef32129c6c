"""Per-op timing of the static eval plan at a given batch (CUDA events around each launch, mean of reps).
Usage: python tools/eval_profile.py [batch] [pw_impl]   -> table on stdout (ms, algorithmic GB/s per op)."""
import contextlib
import io
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_eval import _quantised_fresh  # noqa: E402
from data_free_quantization_b200.eval_plan import EvalPlan  # noqa: E402

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 256
kw = {"pw_impl": sys.argv[2]} if len(sys.argv) > 2 else {}
npz = np.load(os.path.join(ROOT, "tests/golden/mobilenetv2_dfq.npz"))
with contextlib.redirect_stdout(io.StringIO()):
    model, graph, targ, tr = _quantised_fresh(npz)
plan = EvalPlan(model, tr, batch=batch, **kw)
x = torch.randn(batch, 3, 224, 224, generator=torch.Generator().manual_seed(2)).cuda()
plan(x)
plan(x)
torch.cuda.synchronize()
reps = 5
acc = {}
for _ in range(reps):
    plan.stats.zero_()
    for i, (name, fn) in enumerate(plan.ops):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        acc[i] = acc.get(i, 0.0) + a.elapsed_time(b)
total = sum(acc.values()) / reps
kinds = {}
print(f"batch {batch}: {len(plan.ops)} launches, {total:.3f} ms per forward (event-timed per op, serialised)")
for i, (name, fn) in enumerate(plan.ops):
    ms = acc[i] / reps
    info = plan.op_info.get(name, {}) if hasattr(plan, "op_info") else {}
    kind = info.get("kind", name.split("[")[0])
    kinds[kind] = kinds.get(kind, 0.0) + ms
    by = info.get("bytes")
    gbs = f"{by / ms / 1e6:8.0f} GB/s" if by else ""
    print(f"{name:34s} {kind:10s} {ms:8.4f} ms {gbs} {info.get('shape', '')}")
print("by kind:", {k: round(v, 3) for k, v in sorted(kinds.items(), key=lambda kv: -kv[1])})
