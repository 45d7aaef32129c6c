"""Per-phase timing of the streaming CLE kernel (iteration 2): work time per block vs grid-barrier wait."""
import contextlib, io, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import fresh_model, switch
from data_free_quantization_b200 import Cross_layer_equal as C
from data_free_quantization_b200.utils.layer_transform import merge_batchnorm
from data_free_quantization_b200.utils.relation import create_relation
bps = int(os.environ.get("DFQ_CLE_BLOCKS_PER_SM", "1"))
with contextlib.redirect_stdout(io.StringIO()):
    model, _ = fresh_model()
    model, tr, targ = switch(model, "cuda:0")
    graph, bottoms = tr.log.getGraph(), tr.log.getBottoms()
    merge_batchnorm(model, graph, bottoms, targ)
    rels = create_relation(graph, bottoms, targ)
    blocks = 148 * bps
    C.DEBUG_STAMPS = torch.zeros(blocks * 64 * 5, dtype=torch.int64, device="cuda:0")
    info = C.cross_layer_equalization(graph, rels, targ, Save_state=False, return_info=True)
torch.cuda.synchronize()
nph = C.DEBUG_STAMPS.n_phases
st = C.DEBUG_STAMPS.cpu().numpy()[:blocks * nph * 5].reshape(blocks, nph, 5).astype(np.int64)
print("impl", info["impl"], "iters", info["iters"], "phases", nph, "blocks", blocks)
for ph in range(nph):
    t0, t1, t2 = st[:, ph, 0], st[:, ph, 1], st[:, ph, 2]
    start = t0.min()
    work = (t1 - t0)
    print(f"phase {ph}: work per block min/median/max {work.min()/1e3:.1f}/{np.median(work)/1e3:.1f}/{work.max()/1e3:.1f} us; "
          f"last block done at {(t1.max()-start)/1e3:.1f} us; barrier released at {(t2.min()-start)/1e3:.1f}..{(t2.max()-start)/1e3:.1f} us")
    b = int(st[:, ph, 3].argmax())
    print(f"   longest task: {st[b, ph, 3]/1e3:.1f} us, task {C.DEBUG_STAMPS.tasks[int(st[b, ph, 4])]} (tensor,row0,nrows,kind,strip,O,L,up,down)")
