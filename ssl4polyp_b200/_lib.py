"""ctypes binding of libssl4p.so (the C-ABI declared in include/ssl4p.h).

The product path has NO fallback: if the library is missing `load()` raises, and every op raises `RuntimeError`
with the library's own message when a call returns non-zero.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libssl4p.so")
_lib = None

vp, i32, i64, f32, sz = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_size_t

# name -> (restype, argtypes); must list every symbol include/ssl4p.h declares (tests/test_host.py::test_abi_header_symbols_exported_and_bound checks this)
SIGNATURES = {
    "ssl4p_abi_version": (i32, []),
    "ssl4p_reserve_sms": (i32, [i32]),
    "ssl4p_last_error": (C.c_char_p, []),
    "ssl4p_launch_count": (C.c_ulonglong, []),
    "ssl4p_gemm_bf16": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "ssl4p_gemm_set_cta_group": (i32, [i32]),
    "ssl4p_gemm_max_cta_pairs": (i32, []),
    "ssl4p_gemm_set_scheduler_buffer": (i32, [vp, sz]),
    "ssl4p_gemm_pick_splits": (i32, [i32, i32, i32]),
    "ssl4p_gemm_workspace_bytes": (sz, [i32, i32, i32, i32, i32]),
    "ssl4p_splitk_reduce": (i32, [vp, i32, i64, vp, i64, i32, vp]),
    "ssl4p_reduce_rows_f32": (i32, [vp, i32, i32, vp, i32, vp, sz, vp]),
    "ssl4p_crop_resize_workspace_bytes": (sz, [i32, i32]),
    "ssl4p_crop_resize_normalize_u8": (i32, [vp, vp, i32, i32, i32, i32, vp, vp, vp, vp, sz, vp]),
    "ssl4p_crop_resize_u8": (i32, [vp, vp, i32, i32, i32, i32, vp, vp, sz, vp]),
    "ssl4p_color_jitter_u8": (i32, [vp, vp, vp, i32, i32, vp]),
    "ssl4p_gaussian_blur_workspace_bytes": (sz, [i32, i32]),
    "ssl4p_gaussian_blur_u8": (i32, [vp, vp, vp, i32, i32, i32, vp, sz, vp]),
    "ssl4p_flip_rotate_normalize_u8": (i32, [vp, vp, i32, i32, vp, vp, vp, vp, vp]),
    "ssl4p_mask_from_noise": (i32, [vp, i32, i32, i32, vp, vp, vp, vp]),
    "ssl4p_gather_tokens_f32": (i32, [vp, vp, vp, i32, i32, i32, i32, vp]),
    "ssl4p_patchify": (i32, [vp, vp, i32, i32, i32, i32, i32, vp]),
    "ssl4p_unpatchify": (i32, [vp, vp, i32, i32, i32, i32, i32, vp]),
    "ssl4p_im2col_patches": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]),
    "ssl4p_assemble_tokens": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, vp]),
    "ssl4p_assemble_tokens_bwd": (i32, [vp, vp, vp, i32, i32, i32, i32, vp]),
    "ssl4p_layernorm_fwd": (i32, [vp, vp, vp, f32, vp, vp, vp, i32, i32, vp]),
    "ssl4p_layernorm_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp, sz, vp]),
    "ssl4p_layernorm_bwd_workspace_bytes": (sz, [i32, i32]),
    "ssl4p_host_pack_rects": (i32, [vp, vp, vp, vp, vp, vp, i32, i32]),
    "ssl4p_layernorm_bwd_static": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp]),
    "ssl4p_layernorm_bwd_max_static_d": (i32, []),
    "ssl4p_attention_fwd": (i32, [vp, vp, vp, i32, i32, i32, i32, f32, vp]),
    "ssl4p_attention_bwd": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, vp]),
    "ssl4p_decoder_unshuffle_fwd": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, i32, vp]),
    "ssl4p_decoder_unshuffle_bwd": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, sz, vp]),
    "ssl4p_decoder_unshuffle_bwd_workspace_bytes": (sz, [i32, i32, i32]),
    "ssl4p_masked_mse_fwd": (i32, [vp, vp, i64, vp, i32, vp, i32, i32, i32, i32, i32, vp, sz, vp]),
    "ssl4p_masked_mse_bwd": (i32, [vp, vp, i64, vp, i32, vp, vp, i64, i32, i32, i32, i32, i32, vp, sz, vp]),
    "ssl4p_masked_mse_workspace_bytes": (sz, [i32, i32]),
    "ssl4p_colsum_bf16": (i32, [vp, i32, i32, i32, vp, i32, vp, sz, vp]),
    "ssl4p_colsum_workspace_bytes": (sz, [i32, i32]),
    "ssl4p_cast_f32_bf16": (i32, [vp, vp, i64, vp]),
    "ssl4p_adamw_flat": (i32, [vp, vp, vp, vp, vp, vp, i64, f32, f32, f32, f32, f32, f32, f32, f32, vp, vp]),
    "ssl4p_adamw_flat_dev": (i32, [vp, vp, vp, vp, vp, vp, i64, vp, f32, f32, f32, f32, f32, vp, vp]),
    "ssl4p_adamw_flat_dev_range": (i32, [vp, vp, vp, vp, vp, vp, i64, vp, f32, f32, f32, f32, f32, vp, vp]),
    "ssl4p_adamw_advance_step": (i32, [vp, vp]),
    "ssl4p_pool_head_fwd": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]),
    "ssl4p_pool_head_bwd": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]),
}


def load():
    """Load libssl4p.so (once).  Raises if it has not been built: there is no CPU or PyTorch fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m ssl4polyp_b200.csrc.build` "
            "(ssl4polyp_b200 has no CPU/PyTorch fallback for its CUDA kernels)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here == ABI mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    if lib.ssl4p_abi_version() != 1:
        raise RuntimeError(f"libssl4p.so ABI version {lib.ssl4p_abi_version()} != 1")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().ssl4p_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


def launch_count() -> int:
    return int(load().ssl4p_launch_count())
