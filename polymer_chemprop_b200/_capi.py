"""ctypes binding of include/wdmpnn.h (the C-ABI of libwdmpnn.so).

There is no CPU fallback: if the library is missing or the device is not CUDA the
callers raise.  torch is used only for device memory and the current stream.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("WDMPNN_LIB") or os.path.join(_HERE, "libwdmpnn.so")

WD_OK, WD_ERR_INVALID_ARG, WD_ERR_CUDA, WD_ERR_UNSUPPORTED, WD_ERR_GRAPH = 0, -1, -2, -3, -4
ACTIVATIONS = {"ReLU": 0, "LeakyReLU": 1, "tanh": 2, "SELU": 3, "ELU": 4}
AGGREGATIONS = {"mean": 0, "sum": 1, "norm": 2}
MATH_MODES = {"fp32": 0, "tf32": 1, "tf32x3": 2, "bf16x3": 3}
FLAG_NO_SIGN_BITS, FLAG_NO_GRAPH, FLAG_NO_SIDE_BRANCH, FLAG_NO_DEFER_SUM = 1, 2, 4, 8

_f32p, _i32p, _i64p = C.POINTER(C.c_float), C.POINTER(C.c_int32), C.POINTER(C.c_int64)


class WdConfig(C.Structure):
    _fields_ = [("hidden_size", C.c_int32), ("atom_fdim", C.c_int32), ("bond_fdim", C.c_int32),
                ("depth", C.c_int32), ("activation", C.c_int32), ("aggregation", C.c_int32),
                ("undirected", C.c_int32), ("bias", C.c_int32), ("math_mode", C.c_int32),
                ("aggregation_norm", C.c_float), ("dropout", C.c_float), ("dropout_seed_lo", C.c_uint32),
                ("dropout_seed_hi", C.c_uint32), ("flags", C.c_uint32)]


class WdGraph(C.Structure):
    _fields_ = [("n_atoms", C.c_int64), ("n_bonds", C.c_int64), ("n_mols", C.c_int64),
                ("f_atoms", C.c_void_p), ("f_bonds", C.c_void_p), ("w_atoms", C.c_void_p),
                ("w_bonds", C.c_void_p), ("in_ptr", C.c_void_p), ("in_idx", C.c_void_p),
                ("out_idx", C.c_void_p), ("src", C.c_void_p), ("rev", C.c_void_p), ("dst", C.c_void_p),
                ("mol_of", C.c_void_p), ("a_start", C.c_void_p), ("a_size", C.c_void_p),
                ("dop", C.c_void_p), ("ld_atoms", C.c_int32), ("ld_bonds", C.c_int32), ("rev_adjacent", C.c_int32)]


class WdAdam(C.Structure):
    _fields_ = [("lr", C.c_double), ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double),
                ("weight_decay", C.c_double), ("max_grad_norm", C.c_double), ("step", C.c_int64), ("decoupled", C.c_int32)]


class WdStore(C.Structure):
    _fields_ = [("n_graphs", C.c_int64), ("atom_fdim", C.c_int32), ("tail_fdim", C.c_int32)] + \
               [(n, C.c_void_p) for n in ("f_atoms", "bond_tail", "w_atoms", "w_bonds", "b2a", "b2revb", "in_idx", "out_idx",
                                          "in_ptr_local", "atom_off", "bond_off", "dop")]


HEAD_MAX_LAYERS = 8


class WdHead(C.Structure):
    _fields_ = [("n_layers", C.c_int32), ("activation", C.c_int32), ("in_dim", C.c_int32), ("hidden_dim", C.c_int32),
                ("out_dim", C.c_int32), ("W", C.c_void_p * HEAD_MAX_LAYERS), ("b", C.c_void_p * HEAD_MAX_LAYERS)]


class WdHeadGrads(C.Structure):
    _fields_ = [("dW", C.c_void_p * HEAD_MAX_LAYERS), ("db", C.c_void_p * HEAD_MAX_LAYERS)]


class WdWeights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("W_i", "b_i", "W_h", "b_h", "W_o", "b_o")]


class WdGrads(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("dW_i", "db_i", "dW_h", "db_h", "dW_o", "db_o")]


EXPORTS = ("wd_csr_build", "wd_csr_build_check", "wd_expand_features", "wd_store_assemble", "wd_workspace_bytes", "wd_encoder_forward",
           "wd_infer_workspace_bytes", "wd_encoder_infer",
           "wd_encoder_backward", "wd_encoder_host", "wd_last_error", "wd_launch_count",
           "wd_abi_version", "wd_profile_enable", "wd_profile_reset", "wd_profile_read", "wd_debug_gemm",
           "wd_debug_dw", "wd_debug_timing", "wd_adam_scratch_floats", "wd_adam_step", "wd_adam_static_max_params",
           "wd_adam_step_static", "wd_graph_stats", "wd_head_workspace_bytes", "wd_head_forward", "wd_head_backward")
PROF_CATEGORIES = ("input_proj", "aggregate", "depth_step", "atom_update", "readout", "msg_grad", "dw",
                   "dx", "scatter", "misc", "reduce")

ABI_VERSION = 7      # include/wdmpnn.h: 2 = WdGraph gained ld_atoms / ld_bonds, 3 = wd_expand_features, 4 = wd_adam_step,
                     # 5 = WdConfig.flags, WD_MATH_TF32X3, 6 = wd_adam_step_static, 7 = wd_head_*, wd_graph_stats
_lib = None


def lib():
    """Loads libwdmpnn.so (building it in-tree if the sources are newer)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH) or os.environ.get("WDMPNN_REBUILD"):
        from . import build as _build
        _build.build()
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -m polymer_chemprop_b200.build` "
                           "(there is no CPU fallback for the encoder hot path)")
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.wd_csr_build.argtypes = [vp, C.c_int64, C.c_int64, vp, vp, C.c_int64, vp, vp, C.c_int64] + [vp] * 8 + [vp]
    L.wd_csr_build.restype = C.c_int
    L.wd_expand_features.argtypes = [vp, vp, vp, C.c_int64, C.c_int64, C.c_int32, C.c_int32, vp, C.c_int32, vp,
                                     C.c_int32, vp]
    L.wd_expand_features.restype = C.c_int
    L.wd_store_assemble.argtypes = [C.POINTER(WdStore), vp, vp, vp, vp, vp, C.c_int64, C.POINTER(WdGraph), vp, vp]
    L.wd_store_assemble.restype = C.c_int
    L.wd_adam_scratch_floats.argtypes = []
    L.wd_adam_scratch_floats.restype = C.c_int64
    L.wd_adam_step.argtypes = [vp, vp, vp, vp, C.c_int64, C.POINTER(WdAdam), vp, vp, vp]
    L.wd_adam_step.restype = C.c_int
    L.wd_adam_static_max_params.argtypes = []
    L.wd_adam_static_max_params.restype = C.c_int64
    L.wd_adam_step_static.argtypes = [vp, vp, vp, vp, C.c_int64, C.POINTER(WdAdam), vp, vp]
    L.wd_adam_step_static.restype = C.c_int
    L.wd_head_workspace_bytes.argtypes = [C.POINTER(WdHead), C.c_int64]
    L.wd_head_workspace_bytes.restype = C.c_size_t
    L.wd_head_forward.argtypes = [C.POINTER(WdHead), vp, C.c_int32, C.c_int64, vp, vp, vp, vp, C.c_size_t, vp, vp, vp]
    L.wd_head_forward.restype = C.c_int
    L.wd_head_backward.argtypes = [C.POINTER(WdHead), vp, C.c_int32, C.c_int64, vp, C.c_size_t, vp, vp, vp, C.c_int32,
                                   C.POINTER(WdHeadGrads), vp]
    L.wd_head_backward.restype = C.c_int
    L.wd_graph_stats.argtypes = [C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.wd_graph_stats.restype = None
    L.wd_csr_build_check.argtypes = [vp, vp]
    L.wd_csr_build_check.restype = C.c_int
    L.wd_workspace_bytes.argtypes = [C.POINTER(WdConfig), C.c_int64, C.c_int64, C.c_int64]
    L.wd_workspace_bytes.restype = C.c_size_t
    L.wd_encoder_forward.argtypes = [C.POINTER(WdGraph), C.POINTER(WdWeights), C.POINTER(WdConfig), vp,
                                     C.c_size_t, vp, vp]
    L.wd_encoder_forward.restype = C.c_int
    L.wd_infer_workspace_bytes.argtypes = L.wd_workspace_bytes.argtypes
    L.wd_infer_workspace_bytes.restype = C.c_size_t
    L.wd_encoder_infer.argtypes = L.wd_encoder_forward.argtypes
    L.wd_encoder_infer.restype = C.c_int
    L.wd_encoder_backward.argtypes = [C.POINTER(WdGraph), C.POINTER(WdWeights), C.POINTER(WdConfig), vp,
                                      C.c_size_t, vp, C.POINTER(WdGrads), vp]
    L.wd_encoder_backward.restype = C.c_int
    L.wd_encoder_host.argtypes = [vp] * 9 + [C.c_int64] * 4 + [C.POINTER(WdWeights), C.POINTER(WdConfig), vp, vp,
                                                               C.POINTER(WdGrads), vp]
    L.wd_encoder_host.restype = C.c_int
    L.wd_debug_gemm.argtypes = [vp, C.c_int, vp, C.c_int, C.c_int, vp, vp, C.c_int, vp, C.c_int, C.c_int64,
                                C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_size_t, vp]
    L.wd_debug_gemm.restype = C.c_int
    L.wd_debug_dw.argtypes = [vp, C.c_int, vp, C.c_int, vp, C.c_int64, C.c_int, C.c_int, C.c_int, vp,
                              C.c_size_t, vp]
    L.wd_debug_dw.restype = C.c_int
    L.wd_debug_timing.argtypes = [vp, C.c_int]
    L.wd_debug_timing.restype = C.c_int
    L.wd_last_error.restype = C.c_char_p
    L.wd_launch_count.restype = C.c_int64
    L.wd_abi_version.restype = C.c_int
    if L.wd_abi_version() != ABI_VERSION:
        raise RuntimeError(f"{LIB_PATH} has ABI version {L.wd_abi_version()}, this package needs {ABI_VERSION}: rebuild it")
    L.wd_profile_enable.argtypes = [C.c_int]
    L.wd_profile_enable.restype = None
    L.wd_profile_reset.restype = None
    L.wd_profile_read.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    L.wd_profile_read.restype = C.c_int
    _lib = L
    return L


def check(rc: int):
    """Maps the C error convention onto the reference's Python exceptions (mpn.py raises
    ValueError / NotImplementedError; everything else is a RuntimeError)."""
    if rc == WD_OK:
        return
    msg = lib().wd_last_error().decode("utf-8", "replace")
    if rc == WD_ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    if rc in (WD_ERR_INVALID_ARG, WD_ERR_GRAPH):
        raise ValueError(msg)
    raise RuntimeError(f"wdmpnn error {rc}: {msg}")


def bucket(n: int) -> int:
    """Capacity to request for a buffer of ``n`` elements: sizes of 1 Mi elements and more are rounded up to a grid of
    16-32 steps per octave (<= 6 % slack).  Batches drawn at random differ by a fraction of a per cent in their atom /
    bond counts; with exact sizes every new maximum misses torch's caching allocator and costs a device-synchronising
    cudaMalloc of up to several GB in the middle of a training loop (measured: 38 ms stalls in bench.py's train_step)."""
    if n < (1 << 20):
        return n
    g = 1 << (n.bit_length() - 5)
    return (n + g - 1) // g * g


def empty_roomy(shape, dtype, device):
    """``torch.empty(shape)`` as a view of a ``bucket()``-sized allocation (see there)."""
    import torch
    n = 1
    for d in (shape if isinstance(shape, (tuple, list)) else (shape,)):
        n *= int(d)
    cap = bucket(n)
    t = torch.empty(cap, dtype=dtype, device=device)
    return (t if cap == n else t[:n]).view(shape)


_HIGH_WATER = {}


def empty_high_water(key, shape, dtype, device):
    """``empty_roomy`` whose capacity never shrinks for ``key``: the request is the largest bucket seen so far, so after
    the first few batches of a training run every step asks the caching allocator for exactly the same block sizes
    (a new bucket in the middle of an epoch cost one rank a 19 ms cudaMalloc at 8 GPUs, profiles/r2_e_scale_n8.json)."""
    import torch
    n = 1
    for d in (shape if isinstance(shape, (tuple, list)) else (shape,)):
        n *= int(d)
    k = (key, str(device), dtype)
    cap, want = _HIGH_WATER.get(k, 0), bucket(n)
    if want < cap // 2:          # a much smaller batch than the largest so far (another workload): no padding up to it
        cap = want
    elif n > cap:                # a new maximum: grow with 3 % headroom, so that the next slightly larger batch still fits
        cap = bucket(n + n // 32) if n >= (1 << 20) else want
        _HIGH_WATER[k] = cap
    t = torch.empty(cap, dtype=dtype, device=device)
    return (t if cap == n else t[:n]).view(shape)


def ptr(t):
    return None if t is None else t.data_ptr()


def current_stream(device) -> int:
    """Raw cudaStream_t of torch's current stream on ``device`` (the private fast accessor when this torch has it:
    torch.cuda.current_stream() builds a Stream object, ~3 us per call, three calls per step)."""
    try:
        return torch._C._cuda_getCurrentRawStream(device.index if device.index is not None else torch.cuda.current_device())
    except AttributeError:
        return torch.cuda.current_stream(device).cuda_stream


class _NoGuard:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


_NO_GUARD = _NoGuard()


def device_guard(device):
    """``with device_guard(dev):`` makes ``dev`` the current CUDA device for the library call inside (the library
    launches on the current device; the reference never calls set_device).  Free when it already is."""
    idx = device.index
    if idx is None or torch.cuda.current_device() == idx:
        return _NO_GUARD
    return torch.cuda.device(device)


def graph_stats():
    """(graph launches, executable graphs built) since load: small batches go to the GPU as one CUDA graph per pass."""
    a, b = C.c_int64(0), C.c_int64(0)
    lib().wd_graph_stats(C.byref(a), C.byref(b))
    return a.value, b.value


def launch_count() -> int:
    return int(lib().wd_launch_count())


def profile_read():
    """{category: (total_ms, launches)} of the launches recorded since the last reset."""
    out = {}
    for i, name in enumerate(PROF_CATEGORIES):
        ms, n = C.c_double(0.0), C.c_int64(0)
        check(lib().wd_profile_read(i, C.byref(ms), C.byref(n)))
        out[name] = (ms.value, n.value)
    return out
