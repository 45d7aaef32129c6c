"""Worst relative error of bias_correction (ref-body / paper) against the reference fixtures per weight width."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import fresh_model, rel_err, snapshot, switch  # noqa: E402


def main():
    from data_free_quantization_b200.bias_correction import bias_correction
    from data_free_quantization_b200.utils.layer_transform import merge_batchnorm
    for bits in (8, 6, 4):
        fx = np.load(os.path.join(ROOT, "tests", "golden", "biascorr_mobilenetv2%s.npz" % ("" if bits == 8 else f"_w{bits}")))
        for mode in ("paper", "refbody"):
            model, _ = fresh_model()
            model, tr, targ = switch(model, "cuda:0")
            graph, bottoms = tr.log.getGraph(), tr.log.getBottoms()
            merge_batchnorm(model, graph, bottoms, targ)
            if mode == "paper":
                bias_correction(graph, bottoms, tuple(targ), bits_weight=bits, mode="paper")
            else:
                pos = {k: i for i, k in enumerate(graph.keys())}
                re = lambda k: k if isinstance(graph[k], str) else pos[k]
                gp = type(graph)((re(k), v) for k, v in graph.items())
                bp = {re(k): (None if v is None else [re(b) for b in v]) for k, v in bottoms.items()}
                bias_correction(gp, bp, tuple(targ), bits_weight=bits)
            snap = snapshot(model, graph, targ)
            errs = sorted(((rel_err(fx[k], snap[k[len(mode) + 1:]]), k) for k in fx.files if k.startswith(mode + "/")),
                          reverse=True)
            print(bits, mode, ["%.2e %s" % e for e in errs[:4]], flush=True)


if __name__ == "__main__":
    main()
