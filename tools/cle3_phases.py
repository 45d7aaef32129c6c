"""Per-phase timer stamps of the resident CLE kernel (iteration 2).  python tools/cle3_phases.py"""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from helpers import fresh_model, switch
from data_free_quantization_b200 import Cross_layer_equal as CLE
from data_free_quantization_b200.utils.layer_transform import merge_batchnorm
from data_free_quantization_b200.utils.relation import create_relation
dev = "cuda:0"
for rep in range(2):
    model, _ = fresh_model()
    model, tr, targ = switch(model, dev)
    graph, bottoms = tr.log.getGraph(), tr.log.getBottoms()
    with contextlib.redirect_stdout(io.StringIO()):
        merge_batchnorm(model, graph, bottoms, targ)
        rels = create_relation(graph, bottoms, targ)
    CLE.DEBUG_STAMPS = torch.zeros(148 * 33 * 8, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    with contextlib.redirect_stdout(io.StringIO()):
        info = CLE.cross_layer_equalization(graph, rels, targ, Save_state=False, Treshhold=2e-7, return_info=True)
print("impl", info["impl"], "iters", info["iters"], "kernel_ms", info["kernel_ms"], "per iter us", 1e3 * info["kernel_ms"] / info["iters"])
print(CLE.LAST_PLAN)
st = CLE.DEBUG_STAMPS.cpu().numpy().reshape(148, 33, 8).astype(np.float64)
cta0 = 0
for ci, (ln, n) in enumerate(CLE.LAST_PLAN["chains"]):
    blk = st[cta0:cta0 + n]
    t0 = blk[:, 0, 0].min()
    rows = []
    for i in range(ln):
        s = blk[:, i, :7]
        rows.append("ph%d start %.1f wait %.1f stage %.1f passA %.1f arr %.1f passB %.1f end %.1f" % (
            i, (s[:, 0].mean() - t0) / 1e3, (s[:, 1] - s[:, 0]).mean() / 1e3, (s[:, 2] - s[:, 1]).mean() / 1e3,
            (np.where(s[:, 3] > 0, s[:, 3], s[:, 2]) - s[:, 2]).mean() / 1e3,
            (s[:, 4] - np.where(s[:, 3] > 0, s[:, 3], s[:, 2])).mean() / 1e3, (s[:, 5] - s[:, 4]).mean() / 1e3,
            (s[:, 6] - s[:, 5]).mean() / 1e3))
    e = blk[:, 32, :3]
    print(f"chain {ci} len {ln} ctas {n}: iteration {(e[:, 2].max() - t0) / 1e3:.1f} us; phases done at "
          f"{(e[:, 0].mean() - t0) / 1e3:.1f}, grid barrier {(e[:, 1] - e[:, 0]).mean() / 1e3:.1f}, decision "
          f"{(e[:, 2] - e[:, 1]).mean() / 1e3:.1f}")
    if n > 8 or ci == 0:
        for r in rows:
            print("   ", r)
    cta0 += n
