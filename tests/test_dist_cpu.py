"""CPU (gloo, world_size 2): the multi-GPU host logic of data_free_quantization_b200/dist.py -- layer balancing,
the packed all-gather of codes + ranges + biases and the local dequantisation, the sliced single-tensor pass and
the batch-sharded eval loop -- with the oracle standing in for the CUDA kernels (the tests' checker role: the
product's default backend is the CUDA library).  Results must equal the un-sharded oracle bit for bit."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

from data_free_quantization_b200 import dist as D


class OracleBackend:
    """numpy/C oracle in place of the kernels (CPU tensors)."""

    def __init__(self):
        from oracle import oracle as O
        O.build()
        self.O = O

    def quantize_weights(self, tensors, bits):
        out = []
        for t in tensors:
            lo, hi = self.O.minmax(t.numpy())
            y, codes = self.O.fake_quant(t.numpy(), bits, lo, hi, want_codes=True)
            t.copy_(torch.from_numpy(y))
            scale = self.O.quant_params(lo, hi, bits)[1]
            out.append((torch.from_numpy(codes.astype(np.uint8)), torch.tensor([lo, hi, scale, 0.0])))
        return out

    def quantize_bias(self, tensors, bits):
        for t in tensors:                       # utils/layer_transform.py:301-303: float(min), float(max) given
            lo, hi = self.O.minmax(t.numpy())
            t.copy_(torch.from_numpy(self.O.fake_quant(t.numpy(), bits, lo, hi)))

    def dequantize(self, codes, slot, out):
        s, lo = np.float32(slot[2].item()), np.float32(slot[0].item())
        q = codes.numpy().astype(np.float32)
        out.copy_(torch.from_numpy(((q * s).astype(np.float32) + lo).astype(np.float32)).view(out.shape))

    def minmax(self, x):
        return torch.tensor(self.O.minmax(x.numpy()))

    def apply(self, x, bits, lo, hi):
        x.copy_(torch.from_numpy(self.O.fake_quant(x.numpy(), bits, lo, hi)))
        return x


def _layers(seed=0):
    g = torch.Generator().manual_seed(seed)
    shapes = [(32, 3, 3, 3), (32, 1, 3, 3), (16, 32, 1, 1), (96, 16, 1, 1), (96, 1, 3, 3), (24, 96, 1, 1), (100, 64)]
    graph = {}
    for i, s in enumerate(shapes):
        if len(s) == 4:
            m = nn.Conv2d(s[1], s[0], s[2], bias=(i % 3 != 1))
            m.weight.data = torch.randn(s, generator=g) * (0.05 + 0.1 * i)
        else:
            m = nn.Linear(s[1], s[0])
            m.weight.data = torch.randn(s, generator=g) * 0.3
        if m.bias is not None:
            m.bias.data = torch.randn(s[0], generator=g)
        graph[f"layer{i}"] = m
    graph["relu"] = nn.ReLU()
    return graph


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        be = OracleBackend()
        # --- layer-sharded quantisation
        graph = _layers()
        info = D.quantize_targ_layer_sharded(graph, 8, 8, [nn.Conv2d, nn.Linear], backend=be)
        ref = _layers()
        for k, m in ref.items():
            if isinstance(m, (nn.Conv2d, nn.Linear)):
                w = m.weight.data.numpy()
                lo, hi = be.O.minmax(w)
                want = be.O.fake_quant(w, 8, lo, hi)
                assert graph[k].weight.data.numpy().tobytes() == want.tobytes(), (rank, k, "weight")
                if m.bias is not None:
                    blo, bhi = be.O.minmax(m.bias.data.numpy())
                    wantb = be.O.fake_quant(m.bias.data.numpy(), 8, blo, bhi)
                    assert graph[k].bias.data.numpy().tobytes() == wantb.tobytes(), (rank, k, "bias")
        assert sorted(set(info["owner"])) == list(range(world))
        # --- one tensor in contiguous slices
        full = torch.randn(10007, generator=torch.Generator().manual_seed(7))
        a, b = D.shard_range(full.numel(), rank, world)
        mine = full[a:b].clone()
        lo, hi = D.fake_quant_sliced(mine, 6, backend=be)
        glo, ghi = be.O.minmax(full.numpy())
        assert (lo, hi) == (glo, ghi)
        want = be.O.fake_quant(full.numpy(), 6, glo, ghi)[a:b]
        assert mine.numpy().tobytes() == want.tobytes()
        # --- batch-sharded eval loop
        xs = [(torch.arange(10 * 4, dtype=torch.float32).view(10, 4) + i, torch.tensor([3] * 5 + [0] * 5))
              for i in range(3)]
        correct, total = D.inference_all_sharded(lambda x: x, xs)      # "logits" = x: argmax is always column 3
        assert (correct, total) == (15, 30)
        q.put((rank, "ok"))
    except Exception as e:                                  # noqa: BLE001
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_sharded_paths_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(results) == [(0, "ok"), (1, "ok")], results


def test_balance_and_ranges():
    sizes = [1280000, 409600, 307200, 153600, 153600, 864, 288]
    owner = D.balance(sizes, 4)
    loads = [sum(s for s, o in zip(sizes, owner) if o == r) for r in range(4)]
    assert owner[0] == 0 and loads[0] == 1280000                     # the classifier alone on one rank
    assert max(loads[1:]) <= 1280000 and sum(loads) == sum(sizes)
    assert D.balance(sizes, 1) == [0] * len(sizes)
    cover = []
    for r in range(3):
        a, b = D.shard_range(10, r, 3)
        cover += list(range(a, b))
    assert cover == list(range(10))
    assert D.shard_range(2, 3, 4) == (2, 2)                           # more ranks than items: empty shard
