"""Bias-correction reduction (range reduce + eps.E[x]) on a [1024, 2^18] weight: CUDA-event time (--time) or one pass
for an ncu launch list."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn as nn
from data_free_quantization_b200.bias_correction import eps_dot

dev = torch.device("cuda:0")
x = torch.randn(1 << 28, device=dev) * 0.1
lin = nn.Linear(2, 2)
lin.weight = nn.Parameter(x.view(1024, -1), requires_grad=False)
v = torch.randn((1 << 28) // 1024, device=dev)
for _ in range(3):
    eps_dot([lin], [v], 8)
torch.cuda.synchronize()
if "--time" in sys.argv:
    ts = []
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); eps_dot([lin], [v], 8); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    ms = sum(ts) / len(ts)
    print(f"bias_correct mean {ms:.4f} ms best {min(ts):.4f}  {8 * x.numel() / ms / 1e6:.0f} GB/s")
