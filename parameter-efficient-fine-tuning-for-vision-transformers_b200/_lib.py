"""ctypes binding of csrc/libpeftb200.so (the C ABI declared in include/peft_b200.h).

The library is built in-tree by `__graft_entry__.build()` (or `make -C csrc`).  Loading works without a GPU (cudart is
static, libcuda is resolved lazily), which lets the CPU test-suite check the exported symbols; every compute entry
point fails loudly without a device — there is deliberately no fallback path.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# PEFT_B200_LIB: A/B measurements of two builds of the SAME ABI inside one GPU session (developer knob, not a fallback)
LIB_PATH = os.environ.get("PEFT_B200_LIB") or os.path.join(_HERE, "csrc", "libpeftb200.so")

_c_ll = C.c_longlong
_vp = C.c_void_p

# name -> argtypes (restype is always int unless listed in _RESTYPES)
SIGNATURES = {
    "peft_version": [],
    "peft_last_error": [C.c_char_p, C.c_size_t],
    "peft_launch_count": [],
    "peft_kernel_count": [C.c_int],
    "peft_set_gemm_algo": [C.c_int],
    "peft_set_pdl": [C.c_int],
    "peft_set_gemm_streamk": [C.c_int],
    "peft_debug_gemm_stats": [_vp, C.c_int],
    "peft_debug_attn_stats": [_vp, C.c_int],
    "peft_debug_mma_probe": [C.c_int, C.c_int, C.c_int, C.c_int, _vp],
    "peft_set_attention_algo": [C.c_int],
    "peft_gemm_bf16": [_vp, _c_ll, _vp, _c_ll, C.c_int, C.c_int, C.c_int, _vp, _vp, _c_ll, _vp, _c_ll, _vp, _c_ll, C.c_int, _vp],
    "peft_gemm_bf16_noalloc": [_vp, _c_ll, _vp, _c_ll, C.c_int, C.c_int, C.c_int, _vp, _vp, _c_ll, _vp, _c_ll, _vp, _c_ll, C.c_int, _vp],
    "peft_gemm_bf16_rs": [_vp, _c_ll, _vp, _c_ll, C.c_int, C.c_int, C.c_int, _vp, _vp, _c_ll, _vp, _c_ll, _vp, _c_ll, C.c_int, _vp, _vp],
    "peft_gemm_f32_rs": [_vp, _c_ll, _vp, _c_ll, C.c_int, C.c_int, C.c_int, _vp, _vp, _c_ll, _vp, _c_ll, _vp, _c_ll, C.c_int, _vp, _vp],
    "peft_layernorm_fwd": [_vp, _c_ll, C.c_int, _vp, _vp, _vp, _c_ll, C.c_int, _vp, _vp, C.c_int, C.c_int, C.c_float, _vp],
    "peft_layernorm_fwd_add": [_vp, _c_ll, _vp, _vp, _vp, _c_ll, _vp, _vp, _vp, _c_ll, _vp, _c_ll, C.c_int, C.c_int, C.c_float, _vp],
    "peft_layernorm_bwd": [_vp, _c_ll, C.c_int, _vp, _c_ll, C.c_int, _vp, _vp, _vp, _vp, _c_ll, _vp, _c_ll, _vp, _vp, C.c_int, C.c_int, _vp],
    "peft_layernorm_bwd_f32g": [_vp, _c_ll, C.c_int, _vp, _c_ll, C.c_int, _vp, _vp, _vp, _vp, _c_ll, _vp, _c_ll, _vp, _c_ll, _vp, _vp, C.c_int, C.c_int, _vp],
    "peft_attention_fwd": [_vp, _c_ll, _vp, _c_ll, _vp, C.c_int, C.c_int, C.c_int, C.c_float, _vp],
    "peft_attention_bwd": [_vp, _c_ll, _vp, _c_ll, _vp, _c_ll, _vp, _vp, _vp, _c_ll, C.c_int, C.c_int, C.c_int, C.c_float, _vp],
    "peft_im2col_patches": [_vp, _vp, _c_ll, C.c_int, C.c_int, C.c_int, _vp],
    "peft_assemble_tokens": [_vp, _vp, _vp, _vp, C.c_int, C.c_int, C.c_int, _vp],
    "peft_tn_reduce": [_vp, _c_ll, C.c_int, _vp, _c_ll, C.c_int, _vp, _c_ll, C.c_int, _vp],
    "peft_clip_delta": [_vp, _c_ll, _vp, _c_ll, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, _vp],
    "peft_gemm_tn_bf16": [_vp, _c_ll, _vp, _c_ll, C.c_int, C.c_int, C.c_int, _vp, _c_ll, _vp],
    "peft_colsum_bf16": [_vp, _c_ll, C.c_int, C.c_int, _vp, _vp],
    "peft_transpose_bf16": [_vp, _c_ll, C.c_int, C.c_int, _vp, _c_ll, _vp],
    "peft_transpose_colsum_bf16": [_vp, _c_ll, C.c_int, C.c_int, _vp, _c_ll, _vp, _vp],
    "peft_phm_fold": [_vp, _vp, _vp, C.c_int, C.c_int, _vp],
    "peft_kron_factors": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _vp],
    "peft_kron_factors_drule": [_vp, _vp, _vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp],
    "peft_fastfood_fwd": [_vp, C.c_int, C.c_int, _vp, C.c_int, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "peft_fastfood_bwd": [_vp, C.c_int, C.c_int, _vp, C.c_int, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "peft_allreduce_sgd": [_vp, _vp, C.c_int, C.c_int, _vp, _vp, _vp, _c_ll, _vp, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, _vp],
    "peft_preprocess_clip": [_vp, C.c_int, C.c_int, _vp, _vp, C.c_int, _vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                             _vp, _vp, _vp, _vp, _vp],
    # fp32 verification mode
    "peft_gemm_f32": [_vp, _c_ll, _vp, _c_ll, C.c_int, C.c_int, C.c_int, _vp, _vp, _c_ll, _vp, _c_ll, _vp, _c_ll, C.c_int, _vp],
    "peft_tn_reduce_f32": [_vp, _c_ll, C.c_int, _vp, _c_ll, C.c_int, _vp, _c_ll, C.c_int, _vp],
    "peft_attention_fwd_f32": [_vp, _c_ll, _vp, _c_ll, _vp, C.c_int, C.c_int, C.c_int, C.c_float, _vp],
    "peft_attention_fwd_f32_ex": [_vp, _c_ll, _vp, _c_ll, _vp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, _vp],
    "peft_attention_bwd_f32": [_vp, _c_ll, _vp, _c_ll, _vp, _c_ll, _vp, _vp, _vp, _c_ll, C.c_int, C.c_int, C.c_int, C.c_float, _vp],
    "peft_clip_delta_f32": [_vp, _c_ll, _vp, _c_ll, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, _vp],
    "peft_im2col_patches_f32": [_vp, _vp, _c_ll, C.c_int, C.c_int, C.c_int, _vp],
    "peft_assemble_tokens_f32": [_vp, _vp, _vp, _vp, C.c_int, C.c_int, C.c_int, _vp],
    "peft_add_f32": [_vp, _vp, _vp, _c_ll, _vp],
}
_RESTYPES = {"peft_last_error": None, "peft_launch_count": C.c_ulonglong, "peft_kernel_count": C.c_ulonglong}

_lib = None


def load():
    """Load (once) and return the ctypes handle; raises ImportError with build instructions if the .so is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build the CUDA extension first "
            "(python -c 'import __graft_entry__ as g; g.build()' or make -C <package>/csrc). "
            "peft_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here == header/library mismatch
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, C.c_int)
    _lib = lib
    return lib


def last_error() -> str:
    buf = C.create_string_buffer(512)
    load().peft_last_error(buf, 512)
    return buf.value.decode("utf-8", "replace")


def check(rc: int, what: str = ""):
    if rc != 0:
        raise RuntimeError(f"libpeftb200 {what} failed (status {rc}): {last_error()}")


def launch_count() -> int:
    return int(load().peft_launch_count())


KIND_GEMM_1CTA, KIND_GEMM_2CTA, KIND_ATTN_TC_FWD, KIND_ATTN_TC_BWD, KIND_ATTN_LEGACY, KIND_GEMM_TN, KIND_GEMM_4CTA = range(7)


def kernel_count(kind: int) -> int:
    """Launches of one kernel family since load (see peft_kernel_count)."""
    return int(load().peft_kernel_count(kind))
