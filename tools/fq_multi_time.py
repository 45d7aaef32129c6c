"""Times dfq_fake_quant_multi_f32 (range reduce -> decode -> apply) on one 2^28-element tensor: per call with CUDA
events, and five calls deep between one event pair (host latency hidden).  With `--once` it runs the multi path, the
single-tensor min/max and the single-tensor apply once each, for an ncu launch list:
  ncu --metrics gpu__time_duration.sum --clock-control none --csv python tools/fq_multi_time.py --once"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from data_free_quantization_b200 import _lib as L
from data_free_quantization_b200.utils.quantize import fake_quant_tensors

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(11)
lib = L.lib()


x = torch.randn(1 << 28, device=dev, generator=g) * 0.1
y = torch.empty_like(x)
if "--once" in sys.argv:
    from data_free_quantization_b200.utils.quantize import minmax, quantize
    for _ in range(2):
        fake_quant_tensors([x], 8, 0, outs=[y])
    lo, hi = (float(v) for v in minmax(x).tolist())
    quantize(x, 8, lo, hi)
    torch.cuda.synchronize()
    sys.exit(0)
for _rep in range(2):
    for _ in range(3):
        fake_quant_tensors([x], 8, 0, outs=[y])
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fake_quant_tensors([x], 8, 0, outs=[y]); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    # queue-depth form: 5 calls between one event pair (host latency hidden)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        fake_quant_tensors([x], 8, 0, outs=[y])
    b.record(); b.synchronize()
    ms = sum(ts) / len(ts)
    print(f"mean {ms:.4f} ms  best {min(ts):.4f}  {12 * x.numel() / ms / 1e6:.0f} GB/s   5-deep {a.elapsed_time(b) / 5:.4f} ms")
