"""Where one replayed DFQ pipeline pass spends its time: CUDA events between the pieces of PipelinePlan.replay
(first graph, statistics copy, second graph, host side effects, set_quant_minmax) + host wall clock of each piece.
  python tools/pipeline_timeline.py"""
import contextlib, io, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from data_free_quantization_b200 import pipeline_plan, _lib as L

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
case = bench.PipelineCase(dev)
with contextlib.redirect_stdout(io.StringIO()):
    for _ in range(4):
        case.rearm(False); case.step()
torch.cuda.synchronize()
plan = next(v for v in pipeline_plan._plans.values() if isinstance(v, pipeline_plan.PipelinePlan))
ev = lambda: torch.cuda.Event(enable_timing=True)
rows = []
for rep in range(5):
    case.rearm(False)
    torch.cuda.synchronize()
    e = [ev() for _ in range(6)]
    h = [time.perf_counter()]
    e[0].record(); plan.g1.replay(); e[1].record(); h.append(time.perf_counter())
    plan.mm.start(); e[2].record(); h.append(time.perf_counter())
    if plan.g2 is not None:
        plan.g2.replay()
    e[3].record(); h.append(time.perf_counter())
    for d, k, v in plan.fix:
        d[k] = v
    for q, us, nb, tr in plan.measures:
        q.update_stat, q.num_bits, q.training = us, nb, tr
    L.bump_weights_epoch()
    h.append(time.perf_counter())
    plan.mm.finish(); e[4].record(); h.append(time.perf_counter())
    torch.cuda.synchronize()
    rows.append(([e[i].elapsed_time(e[i + 1]) for i in range(4)], [1e3 * (h[i + 1] - h[i]) for i in range(5)]))
names = ["graph 1 (merge..quantise)", "stats D2H", "graph 2 (clip)", "fix-ups + set_quant_minmax finish"]
for r in rows[1:]:
    print("device ms:", {n: round(v, 3) for n, v in zip(names, r[0])}, "total", round(sum(r[0]), 3))
    print("  host ms:", {n: round(v, 3) for n, v in zip(["g1.replay", "mm.start", "g2.replay", "fix-ups", "mm.finish"], r[1])})
print("launches in the graphs:", plan.launches, " fix-ups:", len(plan.fix))
