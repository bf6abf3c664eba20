"""ctypes binding of the C-ABI in ``include/ldmb200.h`` (``libldmb200.so``, built in-tree by
``dsml_thesis_b200/csrc/Makefile``).

There is NO fallback: if the shared library is missing or a call returns an error code this module
raises.  Every wrapper takes raw device addresses (``tensor.data_ptr()``) and the CUDA stream handle
(``torch.cuda.current_stream().cuda_stream``) — no torch type crosses the boundary.
"""
from __future__ import annotations

import ctypes as C
import os

F32, BF16, F16 = 0, 1, 2
ABI_VERSION = 5

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LDMB200_LIB", os.path.join(_HERE, "libldmb200.so"))


class Ldmb200Error(RuntimeError):
    pass


class GemmDesc(C.Structure):
    """``ldmb200_gemm_desc`` (include/ldmb200.h)."""
    _fields_ = [
        ("a", C.c_void_p), ("a2", C.c_void_p),
        ("c1", C.c_int32), ("c2", C.c_int32),
        ("a_dtype", C.c_int32),
        ("batch", C.c_int32), ("h_in", C.c_int32), ("w_in", C.c_int32),
        ("ksize", C.c_int32), ("stride", C.c_int32), ("upsample", C.c_int32),
        ("w", C.c_void_p),
        ("n", C.c_int32),
        ("bias", C.c_void_p),
        ("rowvec", C.c_void_p), ("rowvec_ld", C.c_int32),
        ("residual", C.c_void_p), ("residual_ld", C.c_int32),
        ("geglu", C.c_int32),
        ("out", C.c_void_p),
        ("out_dtype", C.c_int32), ("out_ld", C.c_int32), ("out_nchw", C.c_int32), ("n_valid", C.c_int32),
        ("colstats", C.c_void_p),
        ("pad_mode", C.c_int32), ("pad_", C.c_int32),
        ("splitk_ws", C.c_void_p), ("splitk_ws_bytes", C.c_int64),
        ("prefetch", C.c_void_p), ("prefetch_bytes", C.c_int64),
    ]


SPLITK_COUNTER_BYTES = 1024   # LDMB200_SPLITK_COUNTER_BYTES


class PackJob(C.Structure):
    """``ldmb200_pack_job`` (include/ldmb200.h)."""
    _fields_ = [
        ("src", C.c_void_p), ("dst", C.c_void_p),
        ("kind", C.c_int32),
        ("n", C.c_int32), ("c", C.c_int32), ("ksize", C.c_int32),
        ("n_pad", C.c_int32), ("c_pad", C.c_int32),
        ("row_offset", C.c_int32), ("block_n", C.c_int32),
        ("dst_dtype", C.c_int32), ("pad_", C.c_int32),
    ]


class LinearJob(C.Structure):
    """``ldmb200_linear_job`` (include/ldmb200.h)."""
    _fields_ = [
        ("in_", C.c_void_p), ("w", C.c_void_p), ("bias", C.c_void_p), ("out", C.c_void_p),
        ("in_ld", C.c_int32), ("out_ld", C.c_int32),
        ("rows", C.c_int32), ("k", C.c_int32), ("n", C.c_int32),
        ("silu_in", C.c_int32), ("first_block", C.c_int32), ("pad_", C.c_int32),
    ]


class WgradDesc(C.Structure):
    """``ldmb200_wgrad_desc`` (include/ldmb200.h)."""
    _fields_ = [
        ("dy", C.c_void_p), ("x", C.c_void_p), ("dw", C.c_void_p),
        ("n_stride", C.c_int64), ("c_stride", C.c_int64), ("tap_stride", C.c_int64),
        ("dy_ld", C.c_int32), ("n", C.c_int32),
        ("c_ld", C.c_int32), ("c", C.c_int32),
        ("batch", C.c_int32), ("h", C.c_int32), ("w", C.c_int32),
        ("ksize", C.c_int32), ("stride", C.c_int32), ("pad_", C.c_int32),
    ]


class LinearBwdJob(C.Structure):
    """``ldmb200_linear_bwd_job`` (include/ldmb200.h)."""
    _fields_ = [
        ("in_", C.c_void_p), ("w", C.c_void_p), ("dout", C.c_void_p), ("dw", C.c_void_p), ("dbias", C.c_void_p),
        ("din", C.c_void_p),
        ("in_ld", C.c_int32), ("dout_ld", C.c_int32), ("din_ld", C.c_int32),
        ("rows", C.c_int32), ("k", C.c_int32), ("n", C.c_int32),
        ("silu_in", C.c_int32), ("pad_", C.c_int32),
    ]


#: every symbol include/ldmb200.h declares: name -> (restype, argtypes)
_i, _i64, _f, _p = C.c_int, C.c_int64, C.c_float, C.c_void_p
SIGNATURES = {
    "ldmb200_abi_version": (_i, []),
    "ldmb200_last_error": (C.c_char_p, []),
    "ldmb200_sm_count": (_i, []),
    "ldmb200_ddim_step": (_i, [_p, _p, _p, _p, _p, _p, _i64, _f, _f, _f, _f, _f, _f, _p]),
    "ldmb200_timestep_embedding": (_i, [_p, _p, _i, _i, _f, _p]),
    "ldmb200_small_linear": (_i, [_p, _i, _p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "ldmb200_small_linear_blocks": (_i, [_i, _i]),
    "ldmb200_small_linear_batched": (_i, [_p, _i, _i, _p]),
    "ldmb200_nchw_to_nhwc": (_i, [_p, _p, _i, _i, _i, _i, _i, _p]),
    "ldmb200_nhwc_to_nchw": (_i, [_p, _i, _p, _i, _i, _i, _p]),
    "ldmb200_groupnorm_ws_floats": (_i64, [_i, _i, _i]),
    "ldmb200_groupnorm": (_i, [_p, _i, _p, _i, _p, _p, _p, _p, _i, _i, _i, _i, _f, _i, _p, _p]),
    "ldmb200_groupnorm_static": (_i, [_p, _i, _p, _i, _p, _p, _p, _p, _i, _i, _i, _i, _f, _i, _p]),
    "ldmb200_groupnorm_fused": (_i, [_p, _i, _p, _p, _i, _p, _p, _p, _p, _p, _i, _i, _i, _i, _f, _i, _p, _p]),
    "ldmb200_layernorm": (_i, [_p, _p, _p, _p, _i, _i64, _i, _f, _p]),
    "ldmb200_space_to_depth2": (_i, [_p, _p, _i, _i, _i, _i, _i, _p]),
    "ldmb200_upsample2x": (_i, [_p, _p, _i, _i, _i, _i, _i, _p]),
    "ldmb200_cast_rows": (_i, [_p, _p, _i, _i64, _p]),
    "ldmb200_count_f16_saturated": (_i, [_p, _i64, _p, _p]),
    "ldmb200_igemm_simt": (_i, [C.POINTER(GemmDesc), _p]),
    "ldmb200_igemm_tc": (_i, [C.POINTER(GemmDesc), _p]),
    "ldmb200_pack_weights": (_i, [_p, _i, _p]),
    "ldmb200_attention_simt": (_i, [_p, _i, _i, _p, _i, _p, _i, _i, _p, _i, _i, _i, _i, _i, _f, _p]),
    "ldmb200_attention_tc": (_i, [_p, _i, _i, _i, _i, _p, _i, _i, _i, _i, _i, _f, _p]),
    # ---- training step
    "ldmb200_attention_tc_lse": (_i, [_p, _i, _i, _i, _i, _p, _i, _i, _p, _i, _i, _i, _f, _p]),
    "ldmb200_q_sample": (_i, [_p, _p, _p, _p, _p, _p, _i, _i64, _p]),
    "ldmb200_mse_loss_grad": (_i, [_p, _p, _p, _p, _i, _i64, _p]),
    "ldmb200_grad_cast": (_i, [_p, _i, _p, _i64, _i, _i, _p, _i, _p, _p, _i, _p]),
    "ldmb200_wgrad": (_i, [C.POINTER(WgradDesc), _p]),
    "ldmb200_groupnorm_bwd": (_i, [_p, _i, _p, _i, _p, _p, _p, _p, _i, _p, _i, _p, _i, _p, _p, _i, _i, _i, _f, _i, _p,
                                   _i, _p, _p]),
    "ldmb200_layernorm_bwd": (_i, [_p, _p, _p, _p, _p, _p, _p, _i64, _i, _f, _p]),
    "ldmb200_geglu_fwd": (_i, [_p, _p, _i64, _i, _p]),
    "ldmb200_geglu_bwd": (_i, [_p, _p, _p, _i64, _i, _p]),
    "ldmb200_attention_bwd_ws_floats": (_i64, [_i, _i, _i]),
    "ldmb200_attention_bwd": (_i, [_p, _i, _i, _i, _i, _p, _p, _i, _p, _p, _p, _i, _i, _i, _f, _p]),
    "ldmb200_zero_insert2x": (_i, [_p, _p, _i, _i, _i, _i, _p]),
    "ldmb200_downsum2x": (_i, [_p, _p, _i, _i, _i, _i, _p]),
    "ldmb200_small_linear_bwd": (_i, [_p, _i, _p, _p, _i, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "ldmb200_small_linear_bwd_batched": (_i, [_p, _i, _p]),
    "ldmb200_adamw_ema": (_i, [_p, _p, _p, _p, _p, _i64, _f, _f, _f, _f, _f, _f, _f, _f, _p]),
    "ldmb200_vq_lookup_postconv": (_i, [_p, _p, _i, _i, _p, _p, _i, _p, _i, _i, _p, _p, _i, _i, _p]),
    "ldmb200_softmax_rows": (_i, [_p, _i64, _p, _i, _i64, _i, _i, _f, _p]),
    "ldmb200_conv1d_temporal_attention": (_i, [_p, _p, _p, _p, _p, _i, _p, _p, _i, _i, _f, _p]),
}

_lib = None


def load():
    """Loads ``libldmb200.so`` once; raises ``Ldmb200Error`` if it is missing or has the wrong ABI."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Ldmb200Error(
            f"{LIB_PATH} not found: build it with `make -C dsml_thesis_b200/csrc` "
            "(or `python -c 'import __graft_entry__ as g; g.build()'`); there is no CPU/PyTorch fallback")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    v = lib.ldmb200_abi_version()
    if v != ABI_VERSION:
        raise Ldmb200Error(f"libldmb200 ABI version {v}, binding expects {ABI_VERSION}")
    _lib = lib
    return lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().ldmb200_last_error()
        raise Ldmb200Error(f"{what or 'ldmb200 call'} failed (code {rc}): {msg.decode() if msg else ''}")


def call(name: str, *args):
    """Calls ``name`` and raises on a non-zero return code."""
    check(getattr(load(), name)(*args), name)
