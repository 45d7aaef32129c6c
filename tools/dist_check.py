"""torchrun check of the NCCL paths of dist.py on real GPUs (one process per GPU):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dist_check.py
Every rank quantises the seeded MobileNetV2 (a) alone and (b) layer-sharded with the all-gather, and the two state
dicts must be bit-identical; then the batch-sharded eval must agree with the un-sharded count."""
import contextlib
import io
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import fresh_model  # noqa: E402
from data_free_quantization_b200.dist import (fake_quant_sliced, quantize_targ_layer_sharded, shard_range)  # noqa: E402
from data_free_quantization_b200.utils.layer_transform import quantize_targ_layer  # noqa: E402
from data_free_quantization_b200.utils.quantize import quantize  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
targ = [nn.Conv2d, nn.Linear]


def graph_of(model):
    return {n: m for n, m in model.named_modules() if type(m) in targ}


with contextlib.redirect_stdout(io.StringIO()):
    a, _ = fresh_model()
    b, _ = fresh_model()
    a, b = a.to(dev), b.to(dev)
    quantize_targ_layer(graph_of(a), 8, 8, targ)
    info = quantize_targ_layer_sharded(graph_of(b), 8, 8, targ)
torch.cuda.synchronize()
bad = [n for (n, p), (_, q) in zip(a.state_dict().items(), b.state_dict().items())
       if p.dtype == torch.float32 and not torch.equal(p, q)]
assert not bad, (rank, bad[:5])
assert len(set(info["owner"])) == world
# one tensor in slices
full = torch.randn(1 << 22, device=dev, generator=torch.Generator(device=dev).manual_seed(5))
lo_, hi_ = float(full.min()), float(full.max())
want = quantize(full, 8, lo_, hi_)
s0, s1 = shard_range(full.numel(), rank, world)
mine = full[s0:s1].clone()
lo, hi = fake_quant_sliced(mine, 8)
assert (lo, hi) == (lo_, hi_) and torch.equal(mine, want[s0:s1])
# batch-sharded eval in sync_stats mode: the per-sample extrema of all shards are gathered before every
# running-range update, so each rank's shard of the logits and every QuantMeasure range equal the un-sharded run
import numpy as np  # noqa: E402
from helpers import dfq_quantised  # noqa: E402
from data_free_quantization_b200.eval_plan import EvalPlan  # noqa: E402
from data_free_quantization_b200.dist import inference_all_sharded  # noqa: E402
gold = np.load(os.path.join(ROOT, "tests", "golden", "mobilenetv2_dfq.npz"))
B = 4 * world
x = torch.randn(B, 3, 224, 224, generator=torch.Generator().manual_seed(2))
with contextlib.redirect_stdout(io.StringIO()):
    m1, g1, _b, targ1, tr1 = dfq_quantised("mobilenetv2", dev, gold)
    full = EvalPlan(m1, tr1, batch=B)(x.to(dev)).clone()
    m2, g2, _b, targ2, tr2 = dfq_quantised("mobilenetv2", dev, gold)
    s0, s1 = shard_range(B, rank, world)
    plan = EvalPlan(m2, tr2, batch=s1 - s0, sync_group=dist.group.WORLD)
    part = plan(x[s0:s1].to(dev)).clone()
torch.cuda.synchronize()
assert torch.equal(part.view(s1 - s0, -1), full.view(B, -1)[s0:s1]), (rank, float((part.view(s1 - s0, -1) - full.view(B, -1)[s0:s1]).abs().max()))
for (k1, a1), (k2, a2) in zip(g1.items(), g2.items()):
    if type(a1) in targ1:
        assert float(a1.quant.running_min) == float(a2.quant.running_min) and float(a1.quant.running_max) == float(a2.quant.running_max)
# the eval loop itself (main_dfq.py:66-113) with the counts all-reduced over NCCL (device defaults to the CUDA device)
labels = full.view(B, -1).argmax(1).cpu()
correct, total = inference_all_sharded(lambda xs: plan(xs.to(dev)).view(xs.shape[0], -1), [(x, labels)])
assert (correct, total) == (B, B), (correct, total)
dist.barrier()
if rank == 0:
    print(f"dist_check ok: world {world}, sharded state dict bit-identical ({len(info['owner'])} layers, "
          f"{info['bytes_gathered']} B received per rank, path {info['path']}"
          f"{'' if info.get('p2p_unavailable') is None else ' [' + info['p2p_unavailable'] + ']'}), sliced tensor "
          f"bit-identical, batch-sharded eval with sync_stats equals the un-sharded forward bit for bit")
dist.destroy_process_group()
