"""ctypes binding of the C-ABI CUDA library (``include/psrc_b200.h``).

This is the binding a reference maintainer would add: plain device
pointers, sizes and the current CUDA stream.  There is NO CPU fallback -- if the shared
library is missing or a call fails, a ``RuntimeError`` is raised.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import torch

PKG_DIR = Path(__file__).resolve().parent
LIB_PATH = PKG_DIR / "lib" / "libpsrc_b200.so"

F32, BF16, F16 = 0, 1, 2
ACT_NONE, ACT_QUICKGELU, ACT_QUICKGELU_GRAD = 0, 1, 2
GEMM_AUTO, GEMM_SIMT, GEMM_TCGEN05, GEMM_TCGEN05_1CTA = 0, 1, 2, 3
ATTN_AUTO, ATTN_SIMT, ATTN_MMA, ATTN_TCGEN05 = 0, 1, 2, 3

_vp, _i, _i64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float


class GemmArgs(C.Structure):
    _fields_ = [
        ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32), ("ab_dtype", C.c_int32),
        ("A", _vp), ("lda", _i64), ("W", _vp), ("ldw", _i64),
        ("bias", _vp), ("act", C.c_int32), ("aux_dtype", C.c_int32),
        ("aux", _vp), ("ldaux", _i64),
        ("residual", _vp), ("ldres", _i64),
        ("out", _vp), ("ldo", _i64), ("out_dtype", C.c_int32), ("out2_dtype", C.c_int32),
        ("out2", _vp), ("ldo2", _i64),
        ("alpha", _f), ("backend", C.c_int32),
    ]


class LossArgs(C.Structure):
    _fields_ = [
        ("B", C.c_int32), ("C", C.c_int32), ("E", C.c_int32), ("Ct", C.c_int32),
        ("B_global", C.c_int32), ("Ct_global", C.c_int32),
        ("logits", _vp), ("zs_logits", _vp), ("labels", _vp),
        ("focal_alpha", _vp), ("focal_gamma", _f),
        ("t_hat", _vp), ("t_zs", _vp), ("f_hat", _vp), ("f_zs", _vp),
        ("w_kl", _f), ("w_text", _f), ("w_image", _f),
        ("d_logits", _vp), ("d_that", _vp), ("d_fhat", _vp),
        ("losses", _vp), ("row_scratch", _vp),
    ]


# name -> (restype, argtypes); every symbol include/psrc_b200.h declares
SIGNATURES = {
    "psrc_last_error": (C.c_char_p, []),
    "psrc_version": (_i, []),
    "psrc_launch_count": (C.c_longlong, []),
    "psrc_gemm": (_i, [C.POINTER(GemmArgs), _vp]),
    "psrc_gemm_f32_strided": (_i, [_i, _i, _i, _vp, _i64, _i64, _vp, _i64, _i64, _vp, _i64, _f, _f, _vp]),
    "psrc_layernorm_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _f, _vp]),
    "psrc_layernorm_bwd": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "psrc_attn_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "psrc_attn_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "psrc_attn_row_fwd": (_i, [_vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _vp]),
    "psrc_attn_row_bwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _vp, _i, _i, _i, _i, _vp]),
    "psrc_im2col_patches": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "psrc_assemble_vision_tokens": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _vp]),
    "psrc_assemble_text_prompts": (_i, [_vp, _vp, _i, _vp, _i, _vp, _i, _vp, _vp, _i, _i, _i, _i, _vp]),
    "psrc_insert_prompt": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "psrc_prompt_grad": (_i, [_vp, _vp, _i, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "psrc_gather_ln_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _f, _vp]),
    "psrc_gather_ln_bwd": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "psrc_l2norm_fwd": (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    "psrc_l2norm_bwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _vp]),
    "psrc_loss_fwd_bwd": (_i, [C.POINTER(LossArgs), _vp]),
    "psrc_loss_fwd_bwd_noalloc": (_i, [C.POINTER(LossArgs), _vp]),
    "psrc_sgd_step": (_i, [_vp, _vp, _vp, _vp, _f, _f, _f, _i, _i, _i64, _vp]),
    "psrc_cast": (_i, [_vp, _i, _vp, _i, _i64, _vp]),
    "psrc_axpy": (_i, [_vp, _vp, _f, _i64, _vp]),
    "psrc_small_linear_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "psrc_small_linear_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _vp]),
    "psrc_group_logits_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _f, _vp]),
    "psrc_group_logits_bwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _f, _vp]),
    "psrc_ctx_bias_grad": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "psrc_permute_rows": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "psrc_crop_resize_normalize": (_i, [_vp, _i64, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i,
                                        C.POINTER(_f), C.POINTER(_f), _vp]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen the in-tree library and bind every declared symbol (raises if absent)."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python __graft_entry__.py build` "
            "(nvcc, sm_100a).  This package has no CPU or PyTorch fallback.")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the export is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().psrc_last_error().decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed (code {rc}): {last_error()}")


def launch_count() -> int:
    return int(load().psrc_launch_count())


def dt(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    if t.dtype == torch.float16:
        return F16
    raise TypeError(f"unsupported dtype {t.dtype}")


def ptr(t) -> int | None:
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("psrc_b200 kernels need CUDA tensors (there is no CPU path)")
    return t.data_ptr()


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream
