"""One MobileNetV2 merge + CLE pass (profiling target).  python tools/cle_once.py [reps]"""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from helpers import fresh_model, switch
from data_free_quantization_b200 import Cross_layer_equal as CLE
from data_free_quantization_b200.utils.layer_transform import merge_batchnorm
from data_free_quantization_b200.utils.relation import create_relation
dev = "cuda:0"
for rep in range(int(sys.argv[1]) if len(sys.argv) > 1 else 1):
    model, _ = fresh_model()
    model, tr, targ = switch(model, dev)
    graph, bottoms = tr.log.getGraph(), tr.log.getBottoms()
    with contextlib.redirect_stdout(io.StringIO()):
        merge_batchnorm(model, graph, bottoms, targ)
        rels = create_relation(graph, bottoms, targ)
        info = CLE.cross_layer_equalization(graph, rels, targ, Save_state=False, Treshhold=2e-7, return_info=True)
    print(info["impl"], info["iters"], info["kernel_ms"])
