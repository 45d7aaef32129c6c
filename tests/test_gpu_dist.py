"""GPU: the device side of dist.py -- code output + dfq_dequant_codes_f32 reproduce the fake-quantised weights bit
for bit (what the all-gather ships), and the sharded pass with one rank equals quantize_targ_layer.  The 2-rank
exchange itself is covered on CPU (tests/test_dist_cpu.py, gloo) and by tools/dist_check.py under torchrun."""
import numpy as np
import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("shape,bits", [((96, 16, 1, 1), 8), ((1000, 1280), 8), ((32, 1, 3, 3), 4), ((7,), 6)])
def test_codes_roundtrip_bit_exact(shape, bits):
    from data_free_quantization_b200.dist import CudaBackend
    be = CudaBackend()
    w = torch.randn(shape, device=DEV, generator=torch.Generator(device=DEV).manual_seed(1)) * 0.3 + 0.05
    (codes, slot), = be.quantize_weights([w], bits)
    assert int(codes.max()) <= 2 ** bits - 1
    back = torch.empty_like(w)
    be.dequantize(codes, slot, back)
    assert back.cpu().numpy().tobytes() == w.cpu().numpy().tobytes()


def test_single_rank_sharded_equals_quantize_targ_layer(oracle):
    from data_free_quantization_b200.dist import quantize_targ_layer_sharded
    g = torch.Generator().manual_seed(3)
    graph, ref = {}, {}
    for i, s in enumerate([(32, 3, 3, 3), (64, 32, 1, 1), (10, 64)]):
        m = nn.Conv2d(s[1], s[0], s[2]) if len(s) == 4 else nn.Linear(s[1], s[0])
        m.weight.data = torch.randn(s, generator=g)
        m.bias.data = torch.randn(s[0], generator=g)
        ref[i] = (m.weight.data.numpy().copy(), m.bias.data.numpy().copy())
        graph[i] = m.to(DEV)
    info = quantize_targ_layer_sharded(graph, 8, 8, [nn.Conv2d, nn.Linear])
    assert info["owner"] == [0, 0, 0]
    for i, (w, b) in ref.items():
        lo, hi = oracle.minmax(w)
        assert graph[i].weight.data.cpu().numpy().tobytes() == oracle.fake_quant(w, 8, lo, hi).tobytes()
        lo, hi = oracle.minmax(b)
        assert graph[i].bias.data.cpu().numpy().tobytes() == oracle.fake_quant(b, 8, lo, hi).tobytes()
