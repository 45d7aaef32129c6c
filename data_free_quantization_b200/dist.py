"""Multi-GPU partitioning of the DFQ hot path: one process per GPU under torchrun, NCCL for the two
places where the path has a real exchange.

* Weight-side passes (quantize_targ_layer, utils/layer_transform.py:288-305) shard BY LAYER: every rank fake-
  quantises its share of the layers on its own GPU and the quantised state dict is assembled on every rank with ONE
  all-gather of {8-bit codes, (min, max, scale), quantised biases} -- 4x fewer bytes over NVLink than fp32 weights --
  followed by a local dequantisation that reproduces the owner's fp32 values bit for bit (dfq_dequant_codes_f32).
* One huge tensor (the 4 G-element microbench) shards by contiguous slice: all-reduce of (min, max), local apply.
* Eval (main_dfq.py:66-113) shards BY BATCH with replicated weights; the top-1 counts are all-reduced.
* CLE, absorption, set_quant_minmax run replicated (deterministic, bit-identical on every rank): at MobileNetV2 /
  DeepLab size one all-reduce per CLE iteration costs more than the iteration.

The compute callables are injected (`backend`) so the partition / pack / all-gather / unpack logic is testable on
CPU with gloo; the default backend is the CUDA library and raises without it (no CPU fallback in the product).
"""
import torch
import torch.distributed as dist


def rank_world(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def balance(sizes, parts):
    """Greedy longest-processing-time assignment: owner[i] in [0, parts) for every item, largest items first onto
    the least loaded part (ties -> lowest rank), deterministic on every rank."""
    owner = [0] * len(sizes)
    load = [0] * parts
    for i in sorted(range(len(sizes)), key=lambda i: (-sizes[i], i)):
        r = min(range(parts), key=lambda r: (load[r], r))
        owner[i] = r
        load[r] += sizes[i]
    return owner


def shard_range(n, rank, world):
    """Contiguous [start, stop) of n items for `rank` (first n % world ranks get one more)."""
    base, extra = divmod(n, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


class CudaBackend:
    """The product path: libdfq_b200 kernels on the current CUDA device."""

    def quantize_weights(self, tensors, bits):
        """In-place fake-quant of fp32 tensors; returns [(uint8 codes, fp32 slot[4] = min, max, scale, 0)]."""
        from .utils.quantize import fake_quant_tensors
        if not tensors:
            return []
        codes = [torch.empty(t.shape, dtype=torch.uint8, device=t.device) for t in tensors]
        slots = fake_quant_tensors(tensors, bits, 0, codes=codes)
        return [(c, slots[i]) for i, c in enumerate(codes)]

    def quantize_bias(self, tensors, bits):
        from .utils.quantize import fake_quant_tensors
        if tensors:
            fake_quant_tensors(tensors, bits, 0)

    def dequantize(self, codes, slot, out):
        from . import _lib as L
        slot = slot.contiguous()
        L.check(L.lib().dfq_dequant_codes_f32(codes.data_ptr(), 1, L.dev_f32(out, "weight"), out.numel(), 1,
                                              slot.data_ptr(), L.stream()), "dfq_dequant_codes_f32")
        L.bump_weights_epoch()      # raw-pointer write: packed-code caches of the eval kernels are stale now

    def minmax(self, x):
        from .utils.quantize import minmax
        return minmax(x)

    def apply(self, x, bits, lo, hi):
        from .utils.quantize import quantize
        return quantize(x, bits, lo, hi, inplace=True)


def _pad16(n):
    return (n + 15) // 16 * 16


def _layout(layers, owner, world, with_bias):
    """Byte layout of every rank's send buffer: {layer index: (codes offset, slot offset, bias offset or None)}."""
    offs, totals = {}, [0] * world
    for i, layer in enumerate(layers):
        r = owner[i]
        o = totals[r]
        co = o
        o += _pad16(layer.weight.numel())
        so = o
        o += 16
        bo = None
        if with_bias and layer.bias is not None:
            bo = o
            o += _pad16(4 * layer.bias.numel())
        offs[i] = (co, so, bo)
        totals[r] = o
    return offs, totals


def quantize_targ_layer_sharded(graph, bit_weight=8, bits_bias=16, targ_type=None, group=None, backend=None):
    """quantize_targ_layer (utils/layer_transform.py:288-305) with the layers dealt to the ranks of `group`; on
    return every rank holds the complete quantised weights and biases, identical to the single-GPU pass.
    Returns {"owner": [...], "bytes_gathered": per-rank payload}."""
    assert targ_type is not None, "targ_type cannot be None!"
    assert 1 <= bit_weight <= 8, "the sharded pass ships 8-bit codes"
    backend = backend or CudaBackend()
    rank, world = rank_world(group)
    layers = [graph[k] for k in graph if type(graph[k]) in targ_type]
    owner = balance([layer.weight.numel() for layer in layers], world)
    with_bias = bits_bias < 32
    mine = [i for i in range(len(layers)) if owner[i] == rank]
    with torch.no_grad():
        got = backend.quantize_weights([layers[i].weight.data for i in mine], bit_weight)
        if with_bias:
            backend.quantize_bias([layers[i].bias.data for i in mine if layers[i].bias is not None], bits_bias)
        if world == 1:
            return {"owner": owner, "bytes_gathered": 0}
        offs, totals = _layout(layers, owner, world, with_bias)
        width = max(totals)
        dev = layers[0].weight.device
        send = torch.zeros(width, dtype=torch.uint8, device=dev)
        for i, (codes, slot) in zip(mine, got):
            co, so, bo = offs[i]
            n = codes.numel()
            send[co:co + n].copy_(codes.reshape(-1))
            send[so:so + 16].view(torch.float32).copy_(slot.reshape(-1)[:4])
            if bo is not None:
                b = layers[i].bias.data
                send[bo:bo + 4 * b.numel()].view(torch.float32).copy_(b)
        recv = torch.empty(world * width, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(recv, send, group=group)
        recv = recv.view(world, width)
        for i, layer in enumerate(layers):
            r = owner[i]
            if r == rank:
                continue
            co, so, bo = offs[i]
            n = layer.weight.numel()
            backend.dequantize(recv[r, co:co + n], recv[r, so:so + 16].view(torch.float32), layer.weight.data)
            if bo is not None:
                layer.bias.data.copy_(recv[r, bo:bo + 4 * layer.bias.numel()].view(torch.float32))
    return {"owner": owner, "bytes_gathered": width}


def fake_quant_sliced(x_slice, bits=8, group=None, backend=None):
    """utils/quantize.py:24-78 on ONE tensor split into contiguous slices across the ranks: global (min, max) by an
    all-reduce of two floats, then the local elementwise pass in place.  Returns (min, max) as Python floats."""
    backend = backend or CudaBackend()
    mm = backend.minmax(x_slice)
    red = torch.stack([-mm[0], mm[1]])
    if rank_world(group)[1] > 1:
        dist.all_reduce(red, op=dist.ReduceOp.MAX, group=group)
    lo, hi = -float(red[0]), float(red[1])
    backend.apply(x_slice, bits, lo, hi)
    return lo, hi


def inference_all_sharded(run_batch, batches, group=None, device=None):
    """main_dfq.py:66-113 with every batch split across the ranks.  run_batch(x_shard) -> logits of this rank's
    shard; `batches` yields (x [B,...], labels [B]) host tensors identical on every rank.  Returns (correct, total)
    summed over ranks."""
    rank, world = rank_world(group)
    correct = total = 0
    for x, labels in batches:
        a, b = shard_range(x.shape[0], rank, world)
        if b > a:
            logits = run_batch(x[a:b])
            pred = logits.argmax(1).cpu()
            correct += int((pred == labels[a:b]).sum())
            total += b - a
    t = torch.tensor([correct, total], dtype=torch.int64, device=device or "cpu")
    if world > 1:
        dist.all_reduce(t, group=group)
    return int(t[0]), int(t[1])
