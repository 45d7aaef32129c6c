"""How much of "bit-exact with the reference" rests on the IEEE-sqrt shim (oracle/ref_harness.py ieee_sqrt)?

TEST INFRASTRUCTURE, build container only.  Runs the UNMODIFIED reference's merge_batchnorm + cross-layer
equalisation on the seeded MobileNetV2 twice -- with the shim every golden was taken under, and WITHOUT it
(torch.sqrt through MKL VML, 1 ulp low on ~0.6 % of inputs) -- and records: iteration counts, how many weights /
Relation.S entries / 8-bit weight codes differ and by how much.  Writes tests/golden/sqrt_unpatched_report.json.
    python -m oracle.gen_sqrt_report
"""
import copy as _copy
import json
import os
import sys
import time
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_harness as RH                               # noqa: E402
from data_free_quantization_b200.fixtures import synthetic_state_dict   # noqa: E402


def one_run(ns, patched):
    torch.manual_seed(0)
    model = ns.mobilenet_v2(None)
    model.load_state_dict(synthetic_state_dict(model, seed=0, bn_seed=1))
    run = RH.RefRun(ns, model, torch.ones(4, 3, 224, 224))
    cle_mod = sys.modules["Cross_layer_equal"]
    count = {"n": 0}

    def counting(x, *a, **k):                      # the loop deep-copies the graph once per iteration (:83)
        count["n"] += 1
        return _copy.deepcopy(x, *a, **k)
    cle_mod.copy = types.SimpleNamespace(deepcopy=counting)
    t = time.perf_counter()
    try:
        if patched:
            with RH.ieee_sqrt():
                run.merge_batchnorm()
                run.create_relation()
                run.equalize()
        else:
            run.merge_batchnorm()
            run.create_relation()
            run.equalize()
    finally:
        cle_mod.copy = _copy
    return run.snapshot(), [r.get_scale_vec().clone() for r in run.relations], count["n"], time.perf_counter() - t


def main():
    ns = RH.import_reference()
    snap_p, S_p, it_p, t_p = one_run(ns, True)
    print("patched", it_p, round(t_p, 1), flush=True)
    snap_u, S_u, it_u, t_u = one_run(ns, False)
    print("unpatched", it_u, round(t_u, 1), flush=True)
    rep = {"torch": torch.__version__, "iterations_ieee_sqrt": it_p, "iterations_unpatched": it_u,
           "seconds": [round(t_p, 1), round(t_u, 1)], "tensors": {}}
    tot = nd = 0
    worst = 0.0
    for k, v in snap_u.items():
        a, b = v.numpy(), snap_p[k].numpy()
        d = int((a != b).sum())
        tot += a.size
        nd += d
        if d:
            rel = float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-30)))
            worst = max(worst, rel)
            rep["tensors"][k] = {"differ": d, "of": int(a.size), "max_rel": rel}
    rep["state_elements"], rep["state_elements_differing"], rep["max_rel_diff"] = tot, nd, worst
    sd = sn = 0
    sw = 0.0
    for a, b in zip(S_u, S_p):
        a, b = a.numpy(), b.numpy()
        sd += int((a != b).sum())
        sn += a.size
        if (a != b).any():
            sw = max(sw, float(np.max(np.abs(a - b) / np.abs(b))))
    rep["S_entries"], rep["S_entries_differing"], rep["S_max_rel_diff"] = sn, sd, sw
    cd = cn = 0                                     # 8-bit codes on quantize_targ_layer's grid (utils/quantize.py:51-76)
    for k, v in snap_u.items():
        if not k.endswith(".weight"):
            continue
        a, b = v.numpy().astype(np.float64), snap_p[k].numpy().astype(np.float64)
        qa = np.rint((a - a.min()) / max((a.max() - a.min()) / 255.0, 1e-8))
        qb = np.rint((b - b.min()) / max((b.max() - b.min()) / 255.0, 1e-8))
        cd += int((qa != qb).sum())
        cn += a.size
    rep["weight_codes"], rep["weight_codes_differing"] = cn, cd
    json.dump(rep, open(os.path.join(ROOT, "tests", "golden", "sqrt_unpatched_report.json"), "w"), indent=1)
    print({k: v for k, v in rep.items() if k != "tensors"})


if __name__ == "__main__":
    main()
