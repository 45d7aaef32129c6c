"""Weight clipping on B200 -- host mirror of the reference's clip_weight.py (clip_weight.py:4-33)."""
import ctypes

import torch.nn as nn

from . import _lib as L


def clip_weight(graph, range_clip=None, targ_type=[nn.Conv2d, nn.Linear], verbose=False):
    """layer.weight.data.clamp_(lo, hi) for every target layer, in place on the GPU.

    The reference prints one warning line per non-target graph node (:32-33); pass verbose=True for that."""
    if range_clip is None:
        range_clip = [-15, 15]
    assert isinstance(range_clip, (list, tuple)) and len(range_clip) == 2, \
        "range_clip should be a list or tuple of two elements"
    todo = []
    for idx, layer in graph.items():
        if isinstance(layer, tuple(targ_type)):
            if hasattr(layer, 'weight'):
                todo.append(layer.weight.data)
            else:
                print(f"Warning: Layer at index {idx} does not have 'weight' attribute")
        elif verbose:
            print(f"Warning: Layer at index {idx} is not in the target type list for clipping")
    if not todo:
        return
    descs = (L.DfqClipDesc * len(todo))()
    for d, w in zip(descs, todo):
        d.w, d.n = L.dev_f32(w, "weight"), w.numel()
    blocks = L.i64()
    L.check(L.lib().dfq_clip_plan_host(descs, len(todo), ctypes.byref(blocks)), "dfq_clip_plan_host")
    table = L.table_to_device(descs, todo[0].device)
    L.check(L.lib().dfq_clip_multi_f32(table.data_ptr(), len(todo), blocks.value, float(range_clip[0]),
                                       float(range_clip[1]), L.stream()), "dfq_clip_multi_f32")
