"""dfq_measure_update_noalloc (include/dfq_b200.h): the allocation-free variant of dfq_measure_update, scratch in a
file-scope device array of the library.  Checked against the existing entry point bit for bit (the tolerance of
tests/test_gpu_parity.py for QuantMeasure: exact), on two calls in a row with different inputs (the shared array keeps
the first call's contents), then with the static-range mode and the size limit.  Collected last on purpose."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _call(fn, x, rmin, rmax, slot, update_stat, scratch=None):
    from data_free_quantization_b200 import _lib as L
    n = x.size(0)
    args = [x.data_ptr(), n, x.numel() // n, rmin.data_ptr(), rmax.data_ptr(), slot.data_ptr(), 8, 0, update_stat]
    if scratch is not None:
        args.append(scratch.data_ptr())
    args.append(L.stream())
    return fn(*args)


def _bits(t):
    return t.detach().cpu().view(torch.int32).tolist()


def test_noalloc_matches_measure_update_twice_in_a_row():
    from data_free_quantization_b200 import _lib as L
    lib = L.lib()
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(5)
    # everything the calls touch is allocated up front; the second input is smaller than the first, so stale
    # per-sample extrema of call 1 sit in the shared array behind the ones call 2 uses
    xs = [torch.randn(64, 24, 17, 17, device=dev, generator=g) * 3.0,
          torch.randn(7, 5, 9, 11, device=dev, generator=g) * 0.25 + 1.0]
    state = {k: (torch.full((1,), -0.5, device=dev), torch.full((1,), 0.75, device=dev),
                 torch.zeros(4, device=dev)) for k in ("ref", "new")}
    scratch = torch.empty(2 * 64, dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    for x in xs:
        assert _call(lib.dfq_measure_update, x, *state["ref"], 1, scratch) == 0
        before = torch.cuda.memory_allocated(dev)
        assert _call(lib.dfq_measure_update_noalloc, x, *state["new"], 1) == 0
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == before
        for a, b in zip(state["ref"], state["new"]):
            assert _bits(a) == _bits(b)
        # the slot is {running_min, running_max, scale, 0} and the range did move
        assert _bits(state["new"][2])[0] == _bits(state["new"][0])[0]
    assert float(state["new"][1]) > 0.75 and float(state["new"][0]) < -0.5


def test_noalloc_static_mode_and_limit(capfd):
    from data_free_quantization_b200 import _lib as L
    lib = L.lib()
    dev = torch.device("cuda:0")
    x = torch.randn(4, 8, device=dev)
    ref = (torch.full((1,), -1.25, device=dev), torch.full((1,), 2.5, device=dev), torch.zeros(4, device=dev))
    new = (torch.full((1,), -1.25, device=dev), torch.full((1,), 2.5, device=dev), torch.zeros(4, device=dev))
    # update_stat == 0 (utils/quantize.py:119-123): the range is the running range, untouched
    assert _call(lib.dfq_measure_update, x, *ref, 0, torch.empty(8, device=dev)) == 0
    assert _call(lib.dfq_measure_update_noalloc, x, *new, 0) == 0
    torch.cuda.synchronize()
    for a, b in zip(ref, new):
        assert _bits(a) == _bits(b)
    assert float(new[0]) == -1.25 and float(new[1]) == 2.5
    # one sample more than the built-in scratch holds: refused before anything is launched
    big = torch.zeros(4097, 1, device=dev)
    rc = _call(lib.dfq_measure_update_noalloc, big, *new, 1)
    assert rc != 0
    assert "invalid argument" in lib.dfq_error_string(rc).decode()
    assert "exceeds the built-in scratch" in capfd.readouterr().err
    # the limit itself is accepted and agrees with the existing entry point
    edge = torch.randn(4096, 3, device=dev)
    assert _call(lib.dfq_measure_update, edge, *ref, 1, torch.empty(2 * 4096, device=dev)) == 0
    assert _call(lib.dfq_measure_update_noalloc, edge, *new, 1) == 0
    torch.cuda.synchronize()
    for a, b in zip(ref, new):
        assert _bits(a) == _bits(b)
