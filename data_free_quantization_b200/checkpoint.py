"""Packed integer checkpoint of a DFQ-quantised model.

The reference only ever stores fake-quantised fp32 weights (`print_model_size` never shrinks, ourplots.py:122-132).
Here the quantised state of every target layer is kept as what it is: integer codes (u8, two per byte for grids of
<= 4 bits; u16 above 8 bits) plus the (min, max, scale) of utils/quantize.py:61-68 per tensor, and the bias the same
way.  Unpacking runs the chain's own two fp32 roundings (code * scale, + min: utils/quantize.py:75-76) on the GPU
(dfq_dequant_codes_f32), so the restored weights and biases equal quantize_targ_layer's
(utils/layer_transform.py:288-305) bit for bit.  It is the same payload the layer-sharded all-gather of dist.py ships.
"""
import torch

from . import _lib as L
from .utils.quantize import fake_quant_tensors

FORMAT = "dfq-b200-packed-v1"


def _pack_codes(codes, bits):
    """uint8 / int16 codes -> storage tensor on the CPU (nibble-packed for <= 4 bits)."""
    flat = codes.reshape(-1)
    if bits <= 4:
        if flat.numel() % 2:
            flat = torch.cat([flat, flat.new_zeros(1)])
        flat = flat[0::2] | (flat[1::2] << 4)
    return flat.cpu()


def _unpack_codes(stored, bits, n, device):
    flat = stored.to(device)
    if bits <= 4:
        flat = torch.stack([flat & 0xF, flat >> 4], dim=1).reshape(-1)[:n].contiguous()
    return flat


def _quantise(tensors, bits):
    """In-place fake-quant of a list of fp32 CUDA tensors -> [(codes, slot{min, max, scale, 0})] (one launch set)."""
    if not tensors:
        return []
    dt = torch.uint8 if bits <= 8 else torch.int16
    codes = [torch.empty(t.shape, dtype=dt, device=t.device) for t in tensors]
    slots = fake_quant_tensors(tensors, bits, 0, codes=codes)
    return [(c, slots[i].clone()) for i, c in enumerate(codes)]


def pack_quantized(model, graph, targ_type, bit_weight=8, bits_bias=16):
    """quantize_targ_layer (utils/layer_transform.py:288-305) that also returns the packed integer state:
    {"format", "bits_weight", "bits_bias", "layers": {module name: {...}}}.  The modules are quantised in place
    exactly as quantize_targ_layer does (biases only when bits_bias < 32)."""
    assert targ_type is not None, "targ_type cannot be None!"
    names = {id(m): n for n, m in model.named_modules()}
    layers = [(names[id(graph[k])], graph[k]) for k in graph if type(graph[k]) in targ_type]
    with torch.no_grad():
        wq = _quantise([m.weight.data for _, m in layers], bit_weight)
        with_bias = [(n, m) for n, m in layers if m.bias is not None and bits_bias < 32]
        bq = dict(zip((n for n, _ in with_bias), _quantise([m.bias.data for _, m in with_bias], bits_bias)))
    out = {"format": FORMAT, "bits_weight": int(bit_weight), "bits_bias": int(bits_bias), "layers": {}}
    for (name, m), (codes, slot) in zip(layers, wq):
        entry = {"shape": tuple(m.weight.shape), "codes": _pack_codes(codes, bit_weight), "slot": slot.cpu()}
        if name in bq:
            entry["bias_codes"], entry["bias_slot"] = _pack_codes(bq[name][0], bits_bias), bq[name][1].cpu()
        elif m.bias is not None:
            entry["bias_fp32"] = m.bias.data.cpu().clone()
        out["layers"][name] = entry
    return out


def packed_nbytes(packed):
    """(packed bytes, bytes of the same tensors as fp32)."""
    small = full = 0
    for e in packed["layers"].values():
        n = 1
        for s in e["shape"]:
            n *= s
        small += e["codes"].numel() * e["codes"].element_size() + 16
        full += 4 * n
        for k in ("bias_codes", "bias_fp32"):
            if k in e:
                small += e[k].numel() * e[k].element_size() + (16 if k == "bias_codes" else 0)
                full += 4 * e["shape"][0]
    return small, full


def save_packed(packed, path):
    torch.save(packed, path)


def load_packed(path):
    packed = torch.load(path, map_location="cpu", weights_only=True)
    if packed.get("format") != FORMAT:
        raise ValueError(f"{path}: not a {FORMAT} file")
    return packed


def _dequant(codes, slot, out):
    L.check(L.lib().dfq_dequant_codes_f32(codes.data_ptr(), codes.element_size(), L.dev_f32(out, "tensor"),
                                          out.numel(), 1, slot.data_ptr(), L.stream()), "dfq_dequant_codes_f32")
    L.bump_weights_epoch()          # raw-pointer write: packed-code caches of the eval kernels are stale now


def unpack_into(model, packed):
    """Write the dequantised weights / biases of `packed` into the same-named modules of `model` (CUDA)."""
    mods = dict(model.named_modules())
    bw, bb = packed["bits_weight"], packed["bits_bias"]
    with torch.no_grad():
        for name, e in packed["layers"].items():
            m = mods[name]
            w = m.weight.data
            if tuple(w.shape) != tuple(e["shape"]):
                raise ValueError(f"{name}: shape {tuple(w.shape)} != packed {tuple(e['shape'])}")
            _dequant(_unpack_codes(e["codes"], bw, w.numel(), w.device), e["slot"].to(w.device), w)
            if "bias_codes" in e or "bias_fp32" in e:
                if m.bias is None:
                    m.bias = torch.nn.Parameter(torch.empty(w.shape[0], device=w.device), requires_grad=False)
                b = m.bias.data
                if "bias_codes" in e:
                    _dequant(_unpack_codes(e["bias_codes"], bb, b.numel(), b.device), e["bias_slot"].to(b.device), b)
                else:
                    b.copy_(e["bias_fp32"])
    return model


# ---- true integer biases (the "8-bit weights with 32-bit bias" scheme of the reference's report,
# mla_report_latex/main.tex:157-166) ------------------------------------------------------------------------------
# The reference quantises a bias on its OWN (min, max) grid (utils/layer_transform.py:301-303), which an integer
# pipeline cannot add to an s32 accumulator.  With static activation ranges the accumulator of a layer counts units of
# s_x * s_w (input scale x weight scale), so the bias that an integer engine adds is round(b / (s_x s_w)) as int32.
# This is an optional export next to the reference-exact one: |b_int s_x s_w - b| <= s_x s_w / 2.
def integer_biases(model, graph, targ_type):
    """{module name: {"bias_i32": int32 [O] (CPU), "scale": s_x * s_w}} for every target layer with a bias whose weight
    sits on an integer grid and whose QuantMeasure holds a static range (after set_quant_minmax)."""
    import ctypes
    import numpy as np
    from . import pw_i8
    names = {id(m): n for n, m in model.named_modules()}
    out = {}
    for k in graph:
        m = graph[k]
        if type(m) not in targ_type or m.bias is None or not hasattr(m, "quant"):
            continue
        rec = None
        for b in (8, 7, 6, 5, 4, 3, 2):
            rec = pw_i8._recover_codes(m.weight.data, b)
            if rec is not None:
                break
        if rec is None:
            continue
        s_w = rec[0]
        p = L.DfqQuantParams()
        L.lib().dfq_quant_params_host(float(m.quant.running_min), float(m.quant.running_max), int(m.quant.num_bits), 0,
                                      ctypes.byref(p))
        scale = float(np.float32(p.scale)) * float(s_w)
        bi = torch.round(m.bias.data.double().cpu() / scale).clamp_(-2 ** 31, 2 ** 31 - 1).to(torch.int32)
        out[names[id(m)]] = {"bias_i32": bi, "scale": scale}
    return out


def apply_integer_biases(model, table):
    """Replace the layers' biases by their integer form b_int * (s_x s_w) (what an integer-only engine computes)."""
    mods = dict(model.named_modules())
    with torch.no_grad():
        for name, e in table.items():
            b = mods[name].bias.data
            b.copy_((e["bias_i32"].double() * e["scale"]).float().to(b.device))
    L.bump_weights_epoch()
