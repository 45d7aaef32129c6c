"""Generate tests/golden/biascorr_mobilenetv2.npz from the UNMODIFIED reference's bias_correction.

Build container only (needs /root/reference).  Two fixtures on the seeded MobileNetV2 after merge_batchnorm
(weights un-quantised, so eps = Q(W) - W is non-zero):
  ref-body : the reference function body, driven with a position-keyed graph / bottoms
             (with the id()-keyed graph every script builds it is a no-op);
  paper    : b - eps . E[x], assembled from the reference's own helpers (_quantize_error,
             find_prev_bn, _calculate_bias_correction_for_branches) on the id()-keyed graph.
Usage: python -m oracle.gen_biascorr_golden [--bits 4 --tag _w4]     (bit-width sweep, BASELINE.json configs[2])
"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness as RH                                       # noqa: E402
from data_free_quantization_b200.fixtures import synthetic_state_dict     # noqa: E402


def fresh(ns):
    torch.manual_seed(0)
    model = ns.mobilenet_v2(None)
    model.load_state_dict(synthetic_state_dict(model, seed=0, bn_seed=1))
    run = RH.RefRun(ns, model, torch.ones(4, 3, 224, 224))
    with RH.ieee_sqrt():
        run.merge_batchnorm()
    return run


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bits", type=int, default=8)
    ap.add_argument("--tag", default="")
    args = ap.parse_args()
    ns = RH.import_reference()
    out = {}
    # ---- ref-body ---------------------------------------------------------------------------------
    run = fresh(ns)
    keys = list(run.graph.keys())
    pos = {k: i for i, k in enumerate(keys)}
    re = lambda k: k if isinstance(run.graph[k], str) else pos[k]
    graph_pos = type(run.graph)((re(k), v) for k, v in run.graph.items())
    bottoms_pos = {re(k): (None if v is None else [re(b) for b in v]) for k, v in run.bottoms.items()}
    before, after = ns.bc.bias_correction(graph_pos, bottoms_pos, tuple(run.targ_layer), bits_weight=args.bits)
    print("ref-body: corrected", len(after), "layers")
    for name, t in run.snapshot().items():
        if name.endswith("bias"):
            out["refbody/" + name] = t.numpy()
    # ---- paper ------------------------------------------------------------------------------------
    run = fresh(ns)
    bc, lt = ns.bc, ns.lt
    from scipy.stats import norm
    pdf = lambda x: torch.from_numpy(norm(0, 1).pdf(x)).float()
    cdf = lambda x: torch.from_numpy(norm.cdf(x)).float()
    calc_mean = lambda w, b: w * pdf(-b / w) + b * (1 - cdf(-b / w))
    bn_module, relu_attached = {}, {}
    g, bt = run.graph, run.bottoms
    with torch.no_grad():
        for key, node in g.items():
            bot = bt[key]
            if bot is None or bot[0] == 'Data':
                continue
            if isinstance(node, torch.nn.BatchNorm2d):
                bn_module[key] = node
                relu_attached[key] = False
                continue
            if isinstance(node, torch.nn.ReLU) and bot[0] in bn_module:
                relu_attached[bot[0]] = True
            if isinstance(node, tuple(run.targ_layer)):
                bn_list, relu_list, kind_list, _ = lt.find_prev_bn(bn_module, relu_attached, g, bt, bot[:])
                branch = {"0": [(bn, relu_list[i], kind_list[i]) for i, (bn, _b) in enumerate(bn_list)]}
                (_kind, expect), = bc._calculate_bias_correction_for_branches(branch, calc_mean).values()
                W = node.weight.detach()
                eps = bc._quantize_error(W, args.bits, reduction=None, signed=False)
                eps = eps.view(W.size(0), W.size(1), -1).sum(-1).double()          # [O, I/g]
                groups = getattr(node, "groups", 1)
                e = expect.double().view(groups, -1)                                # [g, I/g]
                og = W.size(0) // groups
                corr = torch.stack([eps[o] @ e[o // og] for o in range(W.size(0))])
                node.bias.data.sub_(corr.float())
    for name, t in run.snapshot().items():
        if name.endswith(".bias"):
            out["paper/" + name] = t.numpy()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "biascorr_mobilenetv2" + args.tag + ".npz"), **out)
    print("written", len(out), "tensors")


if __name__ == "__main__":
    main()
