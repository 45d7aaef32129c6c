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
    """Byte layout of every rank's exchange buffer: [codes of its layers, each padded to 16 B][one 16-byte slot
    (min, max, scale, 0) per layer][fp32 biases].  Returns ({layer: (codes offset, slot offset, bias offset | None)},
    per-rank sizes)."""
    offs, totals = {}, [0] * world
    by_rank = [[i for i in range(len(layers)) if owner[i] == r] for r in range(world)]
    for r, idxs in enumerate(by_rank):
        o = 0
        co = {}
        for i in idxs:
            co[i] = o
            o += _pad16(layers[i].weight.numel())
        so = {}
        for i in idxs:
            so[i] = o
            o += 16
        for i in idxs:
            bo = None
            if with_bias and layers[i].bias is not None:
                bo = o
                o += _pad16(4 * layers[i].bias.numel())
            offs[i] = (co[i], so[i], bo)
        totals[r] = o
    return offs, totals


class _Exchange:
    """The buffer the quantised state crosses GPUs through.  Preferred: a SYMMETRIC-MEMORY buffer
    (torch.distributed._symmetric_memory: every rank maps every peer's buffer over NVLink), so that the dequantising
    kernel reads the codes straight out of their owner's HBM -- one pass, 1 byte per element over the link, no
    all-gather, no staging copy.  Fallback (no P2P / DFQ_DIST_P2P=0): one all_gather_into_tensor of the buffers."""

    _cache = {}

    def __init__(self, width, dev, group):
        import os
        self.width, self.dev, self.group = width, dev, group
        self.handle, self.peers = None, None
        self.world = dist.get_world_size(group)
        if os.environ.get("DFQ_DIST_P2P", "1") != "0" and dev.type == "cuda":
            try:
                import torch.distributed._symmetric_memory as symm
                self.buf = symm.empty(width, dtype=torch.uint8, device=dev)
                self.handle = symm.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
                self.peers = [self.handle.get_buffer(r, (width,), torch.uint8) for r in range(self.world)]
            except Exception as exc:                          # noqa: BLE001 -- reported, then the collective path
                self.handle, self.peers = None, None
                self.why = repr(exc)[:200]
        if self.handle is None:
            self.buf = torch.zeros(width, dtype=torch.uint8, device=dev)
            self.recv = torch.empty(self.world * width, dtype=torch.uint8, device=dev)

    @classmethod
    def get(cls, width, dev, group):
        key = (width, str(dev), id(group))
        if key not in cls._cache:
            if len(cls._cache) > 4:
                cls._cache.clear()
            cls._cache[key] = cls(width, dev, group)
        return cls._cache[key]

    def exchange(self):
        """After every rank has filled its buffer: returns [rank -> uint8 view of that rank's buffer]."""
        if self.handle is not None:
            self.handle.barrier()                             # all buffers complete and visible to their peers
            return self.peers
        dist.all_gather_into_tensor(self.recv, self.buf, group=self.group)
        return list(self.recv.view(self.world, self.width))

    def release(self):
        """Peers are done reading this rank's buffer (before it is overwritten by the next pass)."""
        if self.handle is not None:
            self.handle.barrier()


_PIECE = 1 << 20          # elements per piece of a foreign layer (a multiple of the kernel's 16 K-element block)
_foreign_cache = {}


def _foreign_table(layers, owner, offs, theirs, bufs, rank, world, dev):
    """Descriptor table of the one launch that dequantises every foreign layer, cached per (buffers, weights).

    Order matters over NVLink: blocks run roughly in table order, and with the layers listed owner by owner all ranks
    read from the SAME peer at the same time -- its egress (900 GB/s) shared by N - 1 readers while the other peers sit
    idle (measured at N = 8: 5.1 ms for 0.94 GB per rank, slower than the all-gather).  So every layer is cut into
    1 Mi-element pieces and the pieces are dealt round-robin over the owners, starting with the next rank: at any time a
    rank has loads in flight to all N - 1 peers, and no two ranks start on the same one."""
    from . import _lib as L
    import ctypes
    key = (rank, world, tuple(b.data_ptr() for b in bufs), tuple(layers[i].weight.data_ptr() for i in theirs),
           tuple(layers[i].weight.numel() for i in theirs))
    hit = _foreign_cache.get(key)
    if hit is not None:
        return hit
    per_owner = {}
    for i in theirs:
        base = bufs[owner[i]].data_ptr()
        y, n = L.dev_f32(layers[i].weight.data, "weight"), layers[i].weight.numel()
        for o in range(0, n, _PIECE):
            per_owner.setdefault(owner[i], []).append((base + offs[i][0] + o, base + offs[i][1], y + 4 * o,
                                                       min(_PIECE, n - o)))
    order = [(rank + j) % world for j in range(1, world) if (rank + j) % world in per_owner]
    pieces = []
    for k in range(max(len(v) for v in per_owner.values())):
        pieces += [per_owner[r][k] for r in order if k < len(per_owner[r])]
    descs = (L.DfqDequantDesc * len(pieces))()
    for d, (codes, slot, y, n) in zip(descs, pieces):
        d.codes, d.slot, d.y, d.n = codes, slot, y, n
    blocks = L.i64()
    L.check(L.lib().dfq_dequant_plan_host(descs, len(pieces), ctypes.byref(blocks)), "dfq_dequant_plan_host")
    if len(_foreign_cache) > 8:
        _foreign_cache.clear()
    hit = (L.table_to_device(descs, dev), len(pieces), blocks.value)
    _foreign_cache[key] = hit
    return hit


def quantize_targ_layer_sharded(graph, bit_weight=8, bits_bias=16, targ_type=None, group=None, backend=None):
    """quantize_targ_layer (utils/layer_transform.py:288-305) with the layers dealt to the ranks of `group`; on
    return every rank holds the complete quantised weights and biases, identical to the single-GPU pass.

    Every rank fake-quantises its layers with the codes written IN PLACE into its exchange buffer (one multi-tensor
    launch), then ONE multi-tensor launch dequantises all foreign layers -- from the peers' buffers over NVLink when
    symmetric memory is available, else from an all-gathered copy.
    Returns {"owner": [...], "bytes_gathered": bytes this rank received, "path": "p2p" | "all_gather" | ...}."""
    assert targ_type is not None, "targ_type cannot be None!"
    assert 1 <= bit_weight <= 8, "the sharded pass ships 8-bit codes"
    rank, world = rank_world(group)
    layers = [graph[k] for k in graph if type(graph[k]) in targ_type]
    owner = balance([layer.weight.numel() for layer in layers], world)
    with_bias = bits_bias < 32
    mine = [i for i in range(len(layers)) if owner[i] == rank]
    if backend is not None:
        return _sharded_generic(layers, owner, mine, rank, world, bit_weight, bits_bias, with_bias, group, backend)
    from . import _lib as L
    from .utils.quantize import fake_quant_tensors
    import ctypes
    with torch.no_grad():
        if world == 1:
            CudaBackend().quantize_weights([layers[i].weight.data for i in mine], bit_weight)
            if with_bias:
                CudaBackend().quantize_bias([layers[i].bias.data for i in mine if layers[i].bias is not None], bits_bias)
            return {"owner": owner, "bytes_gathered": 0, "path": "single"}
        offs, totals = _layout(layers, owner, world, with_bias)
        width = _pad16(max(totals))
        dev = layers[0].weight.device
        ex = _Exchange.get(width, dev, group)
        send = ex.buf
        # own layers: codes land in the exchange buffer, ranges in one copy, biases in one concatenation
        if mine:
            ws = [layers[i].weight.data for i in mine]
            codes = [send[offs[i][0]:offs[i][0] + layers[i].weight.numel()].view(layers[i].weight.shape) for i in mine]
            slots = fake_quant_tensors(ws, bit_weight, 0, codes=codes)            # [n_mine, 4] {min, max, scale, 0}
            s0 = offs[mine[0]][1]
            send[s0:s0 + 16 * len(mine)].view(torch.float32).view(len(mine), 4).copy_(slots)
            if with_bias:
                bs = [layers[i].bias.data for i in mine if layers[i].bias is not None]
                if bs:
                    fake_quant_tensors(bs, bits_bias, 0)
                    for i in mine:
                        bo = offs[i][2]
                        if bo is not None:
                            b = layers[i].bias.data
                            send[bo:bo + 4 * b.numel()].view(torch.float32).copy_(b)
        bufs = ex.exchange()
        theirs = [i for i in range(len(layers)) if owner[i] != rank]
        if theirs:
            table, n_desc, blocks = _foreign_table(layers, owner, offs, theirs, bufs, rank, world, dev)
            L.check(L.lib().dfq_dequant_multi_f32(table.data_ptr(), n_desc, blocks, L.stream()),
                    "dfq_dequant_multi_f32")
            if with_bias:
                for i in theirs:
                    bo = offs[i][2]
                    if bo is not None:
                        b = layers[i].bias.data
                        b.copy_(bufs[owner[i]][bo:bo + 4 * b.numel()].view(torch.float32))
        ex.release()
        got = sum(totals[r] for r in range(world) if r != rank)
    return {"owner": owner, "bytes_gathered": got, "path": "p2p" if ex.handle is not None else "all_gather",
            "p2p_unavailable": getattr(ex, "why", None)}


def _sharded_generic(layers, owner, mine, rank, world, bit_weight, bits_bias, with_bias, group, backend):
    """The same exchange with injected compute callables (CPU tests with gloo: tests/test_dist_cpu.py)."""
    with torch.no_grad():
        got = backend.quantize_weights([layers[i].weight.data for i in mine], bit_weight)
        if with_bias:
            backend.quantize_bias([layers[i].bias.data for i in mine if layers[i].bias is not None], bits_bias)
        if world == 1:
            return {"owner": owner, "bytes_gathered": 0, "path": "single"}
        offs, totals = _layout(layers, owner, world, with_bias)
        width = _pad16(max(totals))
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
    return {"owner": owner, "bytes_gathered": width, "path": "all_gather"}


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
    if device is None:          # NCCL reduces CUDA tensors only
        device = torch.device("cuda", torch.cuda.current_device()) if (
            world > 1 and dist.get_backend(group) == "nccl") else "cpu"
    t = torch.tensor([correct, total], dtype=torch.int64, device=device)
    if world > 1:
        dist.all_reduce(t, group=group)
    return int(t[0]), int(t[1])
