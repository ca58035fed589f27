"""ctypes binding of libp2p_b200.so (the C ABI declared in include/p2p_b200.h).

There is no CPU fallback: importing the package without the built library, or calling an op
without a CUDA device, raises.  Build with `p2p-pytorch_b200/csrc/build.sh` (or
`python -c "import __graft_entry__ as g; g.build()"`).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "csrc", "libp2p_b200.so")

F32, BF16 = 0, 1
ACT_NONE, ACT_RELU, ACT_PRELU, ACT_LRELU, ACT_TANH = 0, 1, 2, 3, 4
IMPL_AUTO, IMPL_SIMT, IMPL_TC = 0, 1, 2


class Conv(C.Structure):
    """struct p2p_conv of include/p2p_b200.h."""
    _fields_ = [(n, C.c_int32) for n in
                ("N", "Hi", "Wi", "Ci", "Ho", "Wo", "Co", "kh", "kw", "stride", "pad", "co_real", "ci_real", "act")]


MULTI_MAX = 24   # P2P_MULTI_MAX


class PackJob(C.Structure):
    """struct p2p_pack_job"""
    _fields_ = [("w", C.c_void_p), ("sigma", C.c_void_p), ("w_packed", C.c_void_p), ("wt_packed", C.c_void_p)] + \
               [(n, C.c_int32) for n in ("co_real", "ci_real", "Co", "Ci", "kh", "kw")]


class UnpackJob(C.Structure):
    """struct p2p_unpack_job"""
    _fields_ = [("dw_acc", C.c_void_p), ("grad", C.c_void_p), ("w", C.c_void_p), ("sigma", C.c_void_p),
                ("dot", C.c_void_p)] + [(n, C.c_int32) for n in ("co_real", "ci_real", "Co", "Ci", "kh", "kw")]


class LossJob(C.Structure):
    """struct p2p_loss_job"""
    _fields_ = [("p", C.c_void_p), ("r", C.c_void_p), ("g", C.c_void_p), ("target", C.c_float), ("coef", C.c_float),
                ("pixels", C.c_int64), ("C", C.c_int32), ("c_real", C.c_int32), ("mode", C.c_int32), ("blocks", C.c_int32)]


class SnGradJob(C.Structure):
    """struct p2p_sn_grad_job"""
    _fields_ = [("grad_wbar", C.c_void_p), ("dot", C.c_void_p), ("sigma", C.c_void_p), ("u", C.c_void_p),
                ("v", C.c_void_p), ("rows", C.c_int32), ("cols", C.c_int32)]


def call_multi(name: str, jobs, *extra) -> None:
    """Run a batched entry point over a list of job structures, P2P_MULTI_MAX per launch."""
    for i in range(0, len(jobs), MULTI_MAX):
        chunk = jobs[i:i + MULTI_MAX]
        arr = (type(chunk[0]) * len(chunk))(*chunk)
        call(name, len(chunk), C.cast(arr, C.c_void_p), *extra)


class P2PError(RuntimeError):
    pass


_vp, _i, _i64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float
_cp = C.POINTER(Conv)

# name -> argument ctypes (all entry points return int)
SIGNATURES = {
    "p2p_conv_fwd": [_cp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp],
    "p2p_conv_dgrad": [_cp, _vp, _vp, _vp, _i, _i, _vp],
    "p2p_conv_wgrad": [_cp, _vp, _vp, _vp, _i, _i, _vp],
    "p2p_pack_weights": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp],
    "p2p_unpack_wgrad": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp],
    "p2p_sn_weight_grad": [_vp, _vp, _vp, _vp, _vp, _i, _i, _vp],
    "p2p_spectral_norm": [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "p2p_spectral_norm_static": [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "p2p_g_input_prep": [_vp, _vp, _i, _i, _i, _i, _vp],
    "p2p_channel_stats": [_vp, _vp, _i64, _i, _i, _vp],
    "p2p_bn_finalize": [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp],
    "p2p_bn_act_store": [_vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp],
    "p2p_bn_finalize_act_store": [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp,
                                  _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp],
    "p2p_bn_act_bwd_reduce": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp,
                              _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp],
    "p2p_bn_bwd_apply": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i, _i, _i, _vp],
    "p2p_prelu_grad_finish": [_vp, _vp, _vp],
    "p2p_d_input_prep": [_vp, _i, _vp, _i, _vp, _i, _i, _i, _vp],
    "p2p_avgpool3s2_fwd": [_vp, _vp, _i, _i, _i, _i, _vp],
    "p2p_avgpool3s2_bwd_add": [_vp, _vp, _i, _i, _i, _i, _vp],
    "p2p_cast_pad_channels": [_vp, _vp, _i64, _i, _i, _i, _vp],
    "p2p_uncast_add": [_vp, _vp, _i64, _i, _i, _i, _vp],
    "p2p_d_input_grad": [_vp, _vp, _i, _i, _i, _i, _i, _vp],
    "p2p_act_bwd": [_vp, _vp, _vp, _vp, _vp, _i64, _i, _i, _i, _i, _vp],
    "p2p_tap_gather_sum": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp],
    "p2p_tap_scatter": [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp],
    "p2p_kx_gather_sum": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp],
    "p2p_kx_scatter": [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp],
    "p2p_c_output": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp],
    "p2p_maxpool2_fwd": [_vp, _vp, _i, _i, _i, _i, _i, _vp],
    "p2p_maxpool2_bwd": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp],
    "p2p_zero_fill": [_vp, _i64, _vp],
    "p2p_decode_u8": [_vp, _vp, _vp, _i64, _i, _vp],
    "p2p_nhwc_to_nchw": [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp],
    "p2p_nchw_to_nhwc": [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp],
    "p2p_mse_const_fwd": [_vp, _f, _f, _vp, _i64, _i, _i, _i, _vp],
    "p2p_mse_const_bwd": [_vp, _f, _f, _vp, _vp, _i64, _i, _i, _i, _vp],
    "p2p_mse_flat_fwd": [_vp, _f, _f, _vp, _i64, _vp],
    "p2p_mse_pair_fwd": [_vp, _vp, _f, _vp, _i64, _vp],
    "p2p_psnr_sse": [_vp, _vp, _vp, _i, _i64, _vp],
    "p2p_mse_flat_bwd": [_vp, _f, _f, _vp, _vp, _i64, _vp],
    "p2p_l1_fwd": [_vp, _vp, _f, _vp, _i64, _i, _i, _i, _vp],
    "p2p_l1_bwd": [_vp, _vp, _f, _vp, _vp, _i64, _i, _i, _i, _vp],
    "p2p_map_loss_multi_fwd": [_i, _vp, _vp, _i, _vp],
    "p2p_map_loss_multi_bwd": [_i, _vp, _vp, _i, _vp],
    "p2p_tv_fwd": [_vp, _vp, _i, _i, _i, _i, _vp],
    "p2p_tv_bwd_add": [_vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "p2p_pack_weights_multi": [_i, _vp, _i, _vp],
    "p2p_unpack_wgrad_multi": [_i, _vp, _vp],
    "p2p_sn_weight_grad_multi": [_i, _vp, _vp],
    "p2p_adam_step": [_vp, _vp, _vp, _vp, _i64, _f, _f, _f, _f, _i, _f, _vp],
    "p2p_adam_step_dev": [_vp, _vp, _vp, _vp, _i64, _f, _f, _f, _f, _vp, _f, _vp],
    "p2p_adam_step_graph": [_vp, _vp, _vp, _vp, _i64, _f, _f, _f, _vp, _vp, _vp],
}
NOARG = {"p2p_version": C.c_int, "p2p_last_error": C.c_char_p, "p2p_device_supports_tc": C.c_int,
         "p2p_launch_count": C.c_longlong}
# entry points that take no stream argument: name -> (argument ctypes, result ctype)
MISC = {"p2p_trace_begin": ([], None), "p2p_trace_end": ([C.c_char_p, C.c_int], C.c_int)}

_lib = None


def load() -> C.CDLL:
    """dlopen the in-tree library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise P2PError(f"{LIB_PATH} is missing: build it with p2p-pytorch_b200/csrc/build.sh "
                           "(this package has no CPU or PyTorch fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, args in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.argtypes, fn.restype = args, C.c_int
        for name, res in NOARG.items():
            fn = getattr(lib, name)
            fn.argtypes, fn.restype = [], res
        for name, (args, res) in MISC.items():
            fn = getattr(lib, name)
            fn.argtypes, fn.restype = args, res
        _lib = lib
    return _lib


def trace_begin() -> None:
    """start recording which kernel instantiation serves each convolution call (p2p_trace_begin)"""
    load().p2p_trace_begin()


def trace_end():
    """stop recording; returns the list of recorded lines"""
    lib = load()
    need = lib.p2p_trace_end(None, 0)
    buf = C.create_string_buffer(max(need, 1))
    # the trace is closed by the first call; the text stays until the next p2p_trace_begin
    lib.p2p_trace_end(buf, need)
    return [ln for ln in buf.value.decode().split("\n") if ln]


def launch_count() -> int:
    """kernels enqueued by the library so far in this process"""
    return int(load().p2p_launch_count())


def last_error() -> str:
    return load().p2p_last_error().decode()


def ptr(t):
    """device pointer of a tensor (None -> NULL)"""
    return None if t is None else t.data_ptr()


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def call(name: str, *args) -> None:
    """Call an entry point on the current torch stream; raise on a non-zero status."""
    rc = getattr(load(), name)(*args, stream())
    if rc != 0:
        raise P2PError(f"{name} failed with status {rc}: {last_error()}")


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.float32:
        return F32
    if dt == torch.bfloat16:
        return BF16
    raise P2PError(f"unsupported activation dtype {dt}")


def require_cuda(t: torch.Tensor, what: str = "tensor") -> None:
    if not t.is_cuda:
        raise P2PError(f"{what} must live on a CUDA device: p2p_b200 has no CPU path")
