"""Host mirror of utils/relation.py: the (layer_first, layer_second, bn) pairs CLE and absorption run on.

Pure graph logic (microseconds); kept in Python.  Keys are whatever the graph dict uses
(``id(module)`` ints or op-name strings such as ``add_25`` / ``torch.mean_150``).
"""
from collections import OrderedDict

from torch.nn import AvgPool2d, BatchNorm2d, ReLU

from .quantize import QConv2d, QuantMeasure


class Relation():
    """utils/relation.py:6-33."""

    def __init__(self, layer_idx_1, layer_idx_2, bn_idx_1):
        self.layer_first = layer_idx_1
        self.layer_second = layer_idx_2
        self.bn_idx = bn_idx_1
        self.S = None

    def __repr__(self):
        return '({}, {})'.format(self.layer_first, self.layer_second)

    def get_idxs(self):
        return self.layer_first, self.layer_second, self.bn_idx

    def set_scale_vec(self, S):
        if self.S is None:
            self.S = S
        else:
            self.S *= S

    def get_scale_vec(self):
        return self.S


_PASS_THROUGH = (BatchNorm2d, ReLU, QuantMeasure, AvgPool2d)


def _previous_target(graph, bottoms, key, targ_type, fan_out):
    """Walk up single-input, single-consumer nodes to the preceding target layer (relation.py:47-63)."""
    bot, last_bn = bottoms[key], None
    while len(bot) == 1 and bot[0] != "Data" and fan_out[bot[0]] == 1:
        node = graph[bot[0]]
        if type(node) == BatchNorm2d:
            last_bn = bot[0]
        if type(node) in targ_type:
            return bot[0], last_bn
        transparent = type(node) in _PASS_THROUGH or (
            type(node) == str and ("F.pad" in bot[0] or "torch.mean" in bot[0]))
        if not transparent:
            return None, None
        bot = bottoms[bot[0]]
    return None, None


def create_relation(graph, bottoms, targ_type=[QConv2d], delete_single=False):
    """utils/relation.py:36-106 -- same relations, same order."""
    fan_out = {}
    for key in graph:
        if key == "Data":
            continue
        for b in bottoms[key]:
            fan_out[b] = fan_out.get(b, 0) + 1

    found = OrderedDict()
    for key in graph:
        if type(graph[key]) in targ_type:
            prev, bn = _previous_target(graph, bottoms, key, targ_type, fan_out)
            if prev in found:
                found.pop(prev)                      # relation.py:77-78 (second consumer drops the pair)
            elif prev is not None:
                found[prev] = Relation(prev, key, bn)

    rels = list(found.values())
    if not delete_single:
        return rels
    groups = []                                      # keep only chains of >= 2 relations (:85-101)
    for r in rels:
        hit = -1
        for gi, grp in enumerate(groups):
            for q in grp:
                if r.get_idxs()[0] == q.get_idxs()[1]:
                    hit = gi
                    break
        if hit != -1:
            groups[hit].append(r)
        else:
            groups.append([r])
    return [r for grp in groups if len(grp) > 1 for r in grp]
