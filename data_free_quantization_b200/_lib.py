"""ctypes binding of libdfq_b200.so -- the C ABI declared in include/dfq_b200.h.

There is no CPU fallback: if the shared object is missing or a tensor is not a contiguous fp32 CUDA
tensor, the call raises.  PyTorch is used for device memory and streams only.
"""
import ctypes
import os

import torch

from . import _build

c_f32p = ctypes.POINTER(ctypes.c_float)
c_void = ctypes.c_void_p
i32, i64, f32, f64 = ctypes.c_int32, ctypes.c_int64, ctypes.c_float, ctypes.c_double


class DfqQuantParams(ctypes.Structure):
    _fields_ = [("neg_min", f32), ("scale", f32), ("qmin", f32), ("qmax", f32), ("add_back", f32)]


class DfqQuantDesc(ctypes.Structure):
    _fields_ = [("x", c_void), ("y", c_void), ("codes", c_void), ("n", i64), ("rows", i64), ("bits", i32),
                ("flags", i32), ("clip_lo", f32), ("clip_hi", f32), ("first_block", i64), ("range_slot", i64)]


class DfqClePair(ctypes.Structure):
    _fields_ = [("W1", c_void), ("W2", c_void), ("b1", c_void), ("bn_w", c_void), ("bn_b", c_void), ("S", c_void),
                ("C1", i32), ("K1", i32), ("O2", i32), ("I2", i32), ("KK", i32), ("stage", i32),
                ("first_unit", i32), ("n_units", i32)]


class DfqCleLayer(ctypes.Structure):
    _fields_ = [("W", c_void), ("W_old", c_void), ("n", i64), ("first_chunk", i32), ("n_chunks", i32)]


class DfqCleTensor(ctypes.Structure):
    _fields_ = [("W", c_void), ("W_old", c_void), ("O", i32), ("row_len", i32), ("KK", i32), ("I2", i32), ("Og", i32),
                ("up", i32), ("down", i32), ("layer", i32)]


class DfqCleTask(ctypes.Structure):
    _fields_ = [("tensor", i32), ("row0", i32), ("nrows", i32), ("kind", i32)]


class DfqCle3Tensor(ctypes.Structure):
    _fields_ = [("W", c_void), ("S", c_void), ("b1", c_void), ("bn_w", c_void), ("bn_b", c_void)] + \
               [(k, i32) for k in ("O", "L", "KK", "C", "dw", "up", "down", "s_off", "up_off", "cp_off", "dp_off",
                                   "G", "layer")]


class DfqCle3Chain(ctypes.Structure):
    _fields_ = [(k, i32) for k in ("t0", "len", "cta0", "ncta")]


class DfqCle3Plan(ctypes.Structure):
    _fields_ = [(k, i32) for k in ("n_chains", "n_tensors", "grid", "smem_bytes", "total_C", "cp_words",
                                   "dp_doubles", "fvec_floats", "cp_sm_words", "max_rows")]


class DfqAbsorbPair(ctypes.Structure):
    _fields_ = [("W2", c_void), ("b1", c_void), ("b2", c_void), ("bn_w", c_void), ("bn_b", c_void), ("c", c_void),
                ("wc", c_void), ("C1", i32), ("O2", i32), ("I2", i32), ("KK", i32), ("first_row", i32)]


class DfqCorrDesc(ctypes.Structure):
    _fields_ = [("W", c_void), ("v", c_void), ("out", c_void), ("bias", c_void), ("fold", c_void), ("O", i32),
                ("I2", i32), ("KK", i32), ("bits", i32), ("flags", i32), ("alpha", f32), ("fold_mod", i32),
                ("groups", i32), ("first_unit", i64),
                ("first_block", i64), ("range_slot", i64), ("first_stream", i64)]


class DfqMergeDesc(ctypes.Structure):
    _fields_ = [("W", c_void), ("b", c_void), ("gamma", c_void), ("beta", c_void), ("mean", c_void), ("var", c_void),
                ("fake_weight", c_void), ("fake_bias", c_void), ("O", i64), ("K", i64), ("eps", f32),
                ("first_row", i32)]


class DfqDequantDesc(ctypes.Structure):
    _fields_ = [("codes", c_void), ("y", c_void), ("n", i64), ("slot", c_void), ("first_block", i64)]


class DfqClipDesc(ctypes.Structure):
    _fields_ = [("w", c_void), ("n", i64), ("first_block", i64)]


class DfqConvDesc(ctypes.Structure):
    _fields_ = [("x", c_void), ("y", c_void), ("w", c_void), ("bias", c_void), ("in_slot", c_void),
                ("out_stats", c_void)] + [(n, i32) for n in (
                    "N", "C_in", "H", "W", "C_out", "OH", "OW", "kh", "kw", "stride", "pad", "dil", "groups", "relu",
                    "bits", "reserved")]


class DfqPwI8Desc(ctypes.Structure):
    _fields_ = [("x", c_void), ("y", c_void), ("wpk", c_void), ("wr", c_void), ("bias", c_void), ("in_slot", c_void),
                ("out_stats", c_void), ("G", i64), ("xs_n", i64), ("xs_k", i64), ("xs_p", i64), ("ys_n", i64),
                ("ys_o", i64), ("ys_p", i64), ("P", i32), ("K", i32), ("O", i32), ("bits", i32), ("relu", i32),
                ("s_w", f32), ("c_w", f32), ("o_cap", i32)]


class DfqConvI8Desc(ctypes.Structure):
    _fields_ = [("x", c_void), ("y", c_void), ("wpk", c_void), ("wr", c_void), ("wrt", c_void), ("bias", c_void),
                ("in_slot", c_void), ("out_stats", c_void)] + \
               [(k, i32) for k in ("N", "C", "H", "W", "O", "OH", "OW", "kh", "kw", "stride", "pad", "dil", "bits",
                                   "relu")] + [("s_w", f32), ("c_w", f32), ("scratch", c_void), ("o_cap", i32), ("halo", i32)]


class DfqQ8OutMap(ctypes.Structure):
    _fields_ = [(k, i32) for k in ("OH", "OW", "pl", "pr", "pt", "pb")]


class DfqPwQ8Desc(ctypes.Structure):
    _fields_ = [("x", c_void), ("wpk", c_void), ("wr", c_void), ("bias", c_void), ("in_slot", c_void),
                ("y0", c_void), ("out_slot0", c_void), ("y1", c_void), ("out_slot1", c_void), ("yf", c_void),
                ("yf_g", i64), ("yf_o", i64), ("G", i64), ("K", i32), ("O", i32), ("relu", i32), ("o_cap", i32),
                ("s_w", f32), ("c_w", f32), ("map0", DfqQ8OutMap), ("map1", DfqQ8OutMap)]


class DfqDwQ8Desc(ctypes.Structure):
    _fields_ = [("x", c_void), ("wrow", c_void), ("bias", c_void), ("in_slot", c_void), ("y0", c_void),
                ("out_slot0", c_void), ("y1", c_void), ("out_slot1", c_void)] + \
               [(k, i32) for k in ("N", "H", "W", "C", "stride", "relu")] + [("s_w", f32), ("c_w", f32)] + \
               [(k, i32) for k in ("pad", "dil")]


class DfqConvQ8Desc(ctypes.Structure):
    _fields_ = [("x", c_void), ("wpk", c_void), ("wr", c_void), ("wrt", c_void), ("bias", c_void), ("in_slot", c_void),
                ("y0", c_void), ("out_slot0", c_void), ("y1", c_void), ("out_slot1", c_void)] + \
               [(k, i32) for k in ("N", "C", "H", "W", "O", "OH", "OW", "kh", "kw", "stride", "pad", "dil", "relu")] + \
               [("s_w", f32), ("c_w", f32), ("o_cap", i32), ("tma", i32)]


class DfqStemQ8Desc(ctypes.Structure):
    _fields_ = [("x", c_void), ("w", c_void), ("bias", c_void), ("in_slot", c_void), ("y0", c_void),
                ("out_slot0", c_void), ("y1", c_void), ("out_slot1", c_void)] + \
               [(k, i32) for k in ("N", "H", "W", "O", "stride", "pad", "relu", "reserved")] + \
               [("map0", DfqQ8OutMap), ("map1", DfqQ8OutMap), ("wq", c_void), ("s_w", f32), ("c_w", f32), ("yf", c_void),
                ("out_stats", c_void)]


Q_SYMMETRIC, Q_RANGE_F32, Q_PER_CHANNEL, Q_CLAMP_OUT = 1, 2, 4, 8
ABI_VERSION = 23

# every symbol include/dfq_b200.h declares: (restype, argtypes)
P = ctypes.POINTER
PROTOTYPES = {
    "dfq_abi_version": (ctypes.c_int, []),
    "dfq_error_string": (ctypes.c_char_p, [ctypes.c_int]),
    "dfq_launch_count": (i64, []),
    "dfq_quant_params_host": (None, [f64, f64, ctypes.c_int, ctypes.c_int, P(DfqQuantParams)]),
    "dfq_minmax_f32": (ctypes.c_int, [c_void, i64, c_void, c_void]),
    "dfq_minmax_rows_f32": (ctypes.c_int, [c_void, i64, i64, c_void, c_void]),
    "dfq_fake_quant_apply_f32": (ctypes.c_int, [c_void, c_void, c_void, i64, ctypes.c_int, P(DfqQuantParams), c_void]),
    "dfq_measure_update": (ctypes.c_int, [c_void, i64, i64, c_void, c_void, c_void, ctypes.c_int, ctypes.c_int,
                                          ctypes.c_int, c_void, c_void]),
    "dfq_measure_update_noalloc": (ctypes.c_int, [c_void, i64, i64, c_void, c_void, c_void, ctypes.c_int, ctypes.c_int,
                                                  ctypes.c_int, c_void]),
    "dfq_fake_quant_apply_slot_f32": (ctypes.c_int, [c_void, c_void, c_void, i64, ctypes.c_int, ctypes.c_int, c_void,
                                                     c_void]),
    "dfq_measure_from_stats": (ctypes.c_int, [c_void, ctypes.c_int, i64, c_void, c_void, c_void, ctypes.c_int,
                                              ctypes.c_int, ctypes.c_int, c_void]),
    "dfq_qconv2d_f32": (ctypes.c_int, [P(DfqConvDesc), c_void]),
    "dfq_qlinear_f32": (ctypes.c_int, [c_void, c_void, c_void, c_void, i64, ctypes.c_int, ctypes.c_int, c_void,
                                       ctypes.c_int, ctypes.c_int, c_void, c_void]),
    "dfq_qadd_f32": (ctypes.c_int, [c_void, c_void, c_void, i64, i64, c_void, c_void, ctypes.c_int, ctypes.c_int,
                                    c_void, c_void]),
    "dfq_qmean_f32": (ctypes.c_int, [c_void, c_void, i64, ctypes.c_int, ctypes.c_int, c_void, ctypes.c_int, c_void,
                                     c_void]),
    "dfq_qinterp_bilinear_f32": (ctypes.c_int, [c_void, c_void, i64] + [ctypes.c_int] * 6 +
                                 [c_void, ctypes.c_int, c_void, c_void]),
    "dfq_qcopy_f32": (ctypes.c_int, [c_void, c_void, i64, i64, i64, i64, c_void, ctypes.c_int, c_void, c_void]),
    "dfq_pad2d_f32": (ctypes.c_int, [c_void, c_void, i64] + [ctypes.c_int] * 7 + [c_void, c_void]),
    "dfq_pw_i8_plan_host": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, P(ctypes.c_int), P(ctypes.c_int),
                                           P(ctypes.c_int), P(ctypes.c_int), P(i64)]),
    "dfq_pw_i8_pack_host": (ctypes.c_int, [c_void, ctypes.c_int, ctypes.c_int, c_void]),
    "dfq_pw_i8_plan_cap_host": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, P(ctypes.c_int),
                                               P(ctypes.c_int), P(ctypes.c_int), P(ctypes.c_int), P(i64)]),
    "dfq_pw_i8_pack_cap_host": (ctypes.c_int, [c_void, ctypes.c_int, ctypes.c_int, ctypes.c_int, c_void]),
    "dfq_pw_i8": (ctypes.c_int, [P(DfqPwI8Desc), c_void]),
    "dfq_conv_i8_k_host": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int]),
    "dfq_conv_i8_scratch_bytes": (i64, [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int]),
    "dfq_conv_i8": (ctypes.c_int, [P(DfqConvI8Desc), c_void]),
    "dfq_pw_q8_plan_host": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, P(ctypes.c_int), P(ctypes.c_int),
                                           P(ctypes.c_int), P(ctypes.c_int), P(i64)]),
    "dfq_pw_q8_pack_host": (ctypes.c_int, [c_void, ctypes.c_int, ctypes.c_int, ctypes.c_int, c_void]),
    "dfq_pw_q8": (ctypes.c_int, [P(DfqPwQ8Desc), c_void]),
    "dfq_dw_q8": (ctypes.c_int, [P(DfqDwQ8Desc), c_void]),
    "dfq_stem_q8": (ctypes.c_int, [P(DfqStemQ8Desc), c_void]),
    "dfq_add_q8": (ctypes.c_int, [c_void, c_void, c_void, c_void, c_void, c_void, c_void, c_void, i64, ctypes.c_int,
                                  P(DfqQ8OutMap), P(DfqQ8OutMap), c_void]),
    "dfq_mean_q8": (ctypes.c_int, [c_void, c_void, c_void, c_void, c_void, i64, ctypes.c_int, ctypes.c_int, c_void]),
    "dfq_conv_q8": (ctypes.c_int, [P(DfqConvQ8Desc), c_void]),
    "dfq_conv_q8_halo_k_host": (ctypes.c_int, [ctypes.c_int]),
    "dfq_conv_q8_plan_host": (ctypes.c_int, [ctypes.c_int] * 4 + [P(ctypes.c_int), P(i64)]),
    "dfq_conv_q8_pack_host": (ctypes.c_int, [c_void] + [ctypes.c_int] * 4 + [c_void]),
    "dfq_interp_q8": (ctypes.c_int, [c_void, c_void] + [ctypes.c_int] * 7 + [c_void] * 6),
    "dfq_regrid_q8": (ctypes.c_int, [c_void, c_void, i64, ctypes.c_int, c_void, c_void, c_void, c_void, ctypes.c_int,
                                     ctypes.c_int, ctypes.c_int, c_void]),
    "dfq_pad_q8": (ctypes.c_int, [c_void, c_void] + [ctypes.c_int] * 8 + [c_void, c_void]),
    "dfq_pool_q8": (ctypes.c_int, [c_void, i64, ctypes.c_int, ctypes.c_int, ctypes.c_int, c_void, c_void, c_void, c_void,
                                   c_void]),
    "dfq_quant_plan_host": (ctypes.c_int, [P(DfqQuantDesc), ctypes.c_int, P(i64), P(i64)]),
    "dfq_fake_quant_multi_f32": (ctypes.c_int, [c_void, ctypes.c_int, i64, i64, c_void, c_void]),
    "dfq_dequant_codes_f32": (ctypes.c_int, [c_void, ctypes.c_int, c_void, i64, i64, c_void, c_void]),
    "dfq_dequant_plan_host": (ctypes.c_int, [P(DfqDequantDesc), ctypes.c_int, P(i64)]),
    "dfq_dequant_multi_f32": (ctypes.c_int, [c_void, ctypes.c_int, i64, c_void]),
    "dfq_cle_plan_host": (ctypes.c_int, [P(DfqClePair), ctypes.c_int, P(ctypes.c_int), P(ctypes.c_int)]),
    "dfq_cle_layers_plan_host": (ctypes.c_int, [P(DfqCleLayer), ctypes.c_int, P(ctypes.c_int)]),
    "dfq_cle_result_bytes": (i64, [ctypes.c_int]),
    "dfq_cle_run": (ctypes.c_int, [c_void, ctypes.c_int, ctypes.c_int, c_void, ctypes.c_int, ctypes.c_int, f64, f64,
                                   ctypes.c_int, f32, f64, ctypes.c_int, ctypes.c_int, c_void, c_void, c_void]),
    "dfq_cle2_plan_host": (ctypes.c_int, [P(DfqClePair), ctypes.c_int, P(DfqCleLayer), ctypes.c_int, P(DfqCleTensor),
                                          P(DfqCleTask), ctypes.c_int, P(ctypes.c_int), P(ctypes.c_int), P(i32),
                                          P(ctypes.c_int), P(i32)]),
    "dfq_cle2_run": (ctypes.c_int, [c_void, ctypes.c_int, c_void, ctypes.c_int, c_void, ctypes.c_int, ctypes.c_int,
                                    P(i32), ctypes.c_int, c_void, ctypes.c_int, c_void, c_void, f64, f64, ctypes.c_int,
                                    f32, f64, ctypes.c_int, ctypes.c_int, c_void, c_void, ctypes.c_int, c_void, c_void, c_void]),
    "dfq_cle2_fat_bytes": (i64, [ctypes.c_int]),
    "dfq_cle3_plan_host": (ctypes.c_int, [P(DfqClePair), ctypes.c_int, P(DfqCleLayer), ctypes.c_int, ctypes.c_int,
                                          ctypes.c_int, P(DfqCle3Tensor), P(DfqCle3Chain), P(DfqCle3Plan)]),
    "dfq_cle3_scratch_bytes": (i64, [P(DfqCle3Plan)]),
    "dfq_cle3_run": (ctypes.c_int, [c_void, c_void, P(DfqCle3Plan), c_void, f64, f64, ctypes.c_int, f32, f64,
                                    ctypes.c_int, ctypes.c_int, c_void, c_void, c_void]),
    "dfq_absorb_plan_host": (ctypes.c_int, [P(DfqAbsorbPair), ctypes.c_int, P(ctypes.c_int)]),
    "dfq_absorb_clip": (ctypes.c_int, [c_void, ctypes.c_int, ctypes.c_int, ctypes.c_int, f32, ctypes.c_int, f32, f32,
                                       c_void]),
    "dfq_clip_f32": (ctypes.c_int, [c_void, i64, f32, f32, c_void]),
    "dfq_clip_plan_host": (ctypes.c_int, [P(DfqClipDesc), ctypes.c_int, P(i64)]),
    "dfq_clip_multi_f32": (ctypes.c_int, [c_void, ctypes.c_int, i64, f32, f32, c_void]),
    "dfq_merge_plan_host": (ctypes.c_int, [P(DfqMergeDesc), ctypes.c_int, P(ctypes.c_int)]),
    "dfq_merge_bn_multi": (ctypes.c_int, [c_void, ctypes.c_int, ctypes.c_int, ctypes.c_int, c_void]),
    "dfq_corr_plan_host": (ctypes.c_int, [P(DfqCorrDesc), ctypes.c_int, P(i64), P(i64), P(i64)]),
    "dfq_bias_correct": (ctypes.c_int, [c_void, ctypes.c_int, i64, i64, i64, c_void, ctypes.c_int, c_void]),
    "dfq_merge_bn": (ctypes.c_int, [c_void, c_void, i64, i64, c_void, c_void, c_void, c_void, f32, c_void, c_void,
                                    c_void]),
}

_lib = None


def so_path():
    return _build.SO


def lib():
    """Load libdfq_b200.so (once).  Raises if it has not been built: there is no fallback path."""
    global _lib
    if _lib is None:
        path = so_path()
        if not os.path.exists(path):
            raise RuntimeError(
                f"{path} is missing: build it with `python -m data_free_quantization_b200._build` "
                "(the DFQ hot path has no CPU/PyTorch fallback)")
        h = ctypes.CDLL(path)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(h, name)               # AttributeError if the .so is stale
            fn.restype = res
            fn.argtypes = args
        _lib = h
    return _lib


def check(err, what=""):
    if err != 0:
        msg = lib().dfq_error_string(int(err)).decode()
        raise RuntimeError(f"libdfq_b200 {what}: CUDA error {err}: {msg}")


def stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def dev_f32(t, name="tensor", allow_none=False):
    """data_ptr of a contiguous fp32 CUDA tensor; anything else is an error (no CPU path)."""
    if t is None:
        if allow_none:
            return None
        raise ValueError(f"{name} is None")
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name}: expected a torch.Tensor, got {type(t)}")
    if not t.is_cuda:
        raise RuntimeError(f"{name} lives on {t.device}: the DFQ kernels run on CUDA only (no CPU fallback)")
    if t.dtype != torch.float32:
        raise TypeError(f"{name}: expected float32, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    return t.data_ptr()


_weights_epoch = 0


def weights_epoch():
    """Counter that moves whenever a multi-tensor (weight-side) pass is launched: caches derived from weight values
    (packed integer codes of the eval kernels) are keyed on it."""
    return _weights_epoch


_capture = None          # capture mode (pipeline_plan.py): {"arena": pinned uint8 tensor, "off": int, "keep": list}


def bump_weights_epoch():
    """Call after any write to layer weights that does not go through table_to_device (single-tensor kernels,
    checkpoint unpacking, CUDA-graph replays of weight-side passes): invalidates the packed-code caches of the eval
    kernels (utils/quantize.py _packed)."""
    global _weights_epoch
    _weights_epoch += 1


def begin_capture(arena_bytes=1 << 20):
    """Descriptor tables uploaded from now on are staged in ONE pinned arena that stays alive (and unchanged) with
    the captured CUDA graph: a replay re-reads the host staging memory of every captured copy."""
    global _capture
    _capture = {"arena": torch.empty(arena_bytes, dtype=torch.uint8, pin_memory=True), "off": 0, "keep": []}
    return _capture


def end_capture():
    global _capture
    cap, _capture = _capture, None
    return cap


def table_to_device(arr, device):
    """ctypes array of descriptor structs -> uint8 CUDA tensor holding the same bytes.  Staged through pinned
    memory and copied asynchronously on the current stream: a pageable copy would block the host until every
    kernel already queued (e.g. the whole CLE loop) has finished.  torch's caching host allocator keeps the
    pinned block alive until the copy has run."""
    global _weights_epoch
    _weights_epoch += 1                       # every weight-side pass uploads a descriptor table
    n = ctypes.sizeof(arr)
    cap = _capture
    if cap is not None and cap["off"] + n <= cap["arena"].numel():
        host = cap["arena"][cap["off"]:cap["off"] + n]
        cap["off"] += (n + 63) // 64 * 64
    else:
        host = torch.empty(n, dtype=torch.uint8, pin_memory=True)
        if cap is not None:
            cap["keep"].append(host)
    ctypes.memmove(host.data_ptr(), ctypes.addressof(arr), n)
    out = host.to(device, non_blocking=True)
    if cap is not None:
        cap["keep"].append(out)
    return out


_replayed_launches = 0


def add_launches(n):
    """Kernel launches executed by replaying a captured CUDA graph (the C-side counter only sees direct calls)."""
    global _replayed_launches
    _replayed_launches += int(n)


def launch_count():
    return int(lib().dfq_launch_count()) + _replayed_launches
