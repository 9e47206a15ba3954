"""ctypes binding of libb200llm.so (the C-ABI declared in include/b200llm.h).

There is deliberately no fallback: if the shared library is missing or a call returns a
non-zero status this module raises. The signatures below are the single Python-side copy of
the header; tests/test_abi.py checks that every symbol the header declares is exported.
"""
from __future__ import annotations

import ctypes as C
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200llm.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "b200llm.h")

_lib = None

P = C.c_void_p
I64 = C.c_int64
I = C.c_int
F = C.c_float
SZ = C.c_size_t

class GemmProblem(C.Structure):
    """b200_gemm_problem of include/b200llm.h."""
    _fields_ = [("A", C.c_void_p), ("lda", C.c_int64), ("a_mn", C.c_int),
                ("B", C.c_void_p), ("ldb", C.c_int64), ("b_mn", C.c_int),
                ("C", C.c_void_p), ("ldc", C.c_int64),
                ("M", C.c_int64), ("N", C.c_int64), ("K", C.c_int64),
                ("out_f32", C.c_int), ("accumulate", C.c_int), ("trans_out", C.c_int),
                ("alpha", C.c_float)]


GEMM_GROUP_MAX = 16

# name -> (restype, argtypes)
SIGNATURES = {
    "b200_version": (I, []),
    "b200_init": (I, [I]),
    "b200_last_error": (C.c_char_p, []),
    "b200_set_sm_limit": (I, [I]),
    "b200_launch_count": (C.c_uint64, []),
    "b200_gemm_nt_bf16": (I, [P, I64, P, I64, I64, P, I64, P, I64, I64, P, P, I64, P, I64, I64, I64, I, P]),
    "b200_gemm_ex": (I, [P, I64, I, P, I64, I, I64, P, I64, I, P, I64, I, I64, P, P, I64, P, I64,
                         I64, I64, I, I, I, I, F, I, P]),
    "b200_gemm_lora_packed_bf16": (I, [P, I64, P, I64, I64, P, I64, P, I64, I64, I, C.POINTER(C.c_int64), P, I64, I64, I64, P]),
    "b200_gemm_swiglu_bwd_bf16": (I, [P, I64, P, I64, I64, P, I64, P, I64, I64, P, I64, P, I64, I64, I64, P]),
    "b200_gemm_swiglu_fwd_bf16": (I, [P, I64, P, I64, I64, P, I64, P, I64, I64, P, I64, P, I64, I64, I64, P]),
    "b200_gemm_qkv_rope_bf16": (I, [P, I64, P, I64, I64, P, I64, P, I64, I64, I, C.POINTER(C.c_int64), P, P, I64, I64, I64, P,
                                    I64, I64, I64, P]),
    "b200_gemm_gelu_fwd_bf16": (I, [P, I64, P, I64, I64, P, I64, P, I64, I64, P, P, I64, P, I64, I64, I64, P]),
    "b200_gemm_gelu_bwd_bf16": (I, [P, I64, P, I64, I64, P, I64, P, I64, I64, P, I64, P, I64, I64, I64, P]),
    "b200_gemm_dropout_residual_bf16": (I, [P, I64, P, I64, I64, P, I64, P, I64, I64, P, P, I64, P, I64, I64, I64, F,
                                            C.c_uint64, P, P]),
    "b200_gemm_path_counts": (I, [C.POINTER(C.c_uint64), I]),
    "b200_gemm_set_option": (I, [I, I]),
    "b200_gemm_workspace_bytes": (SZ, []),
    "b200_gemm_set_workspace": (I, [P, SZ]),
    "b200_gemm_set_stream_workspace": (I, [P, P, SZ]),
    "b200_gemm_grouped_bf16": (I, [I, C.POINTER(GemmProblem), P]),
    "b200_gemm_grouped_workspace_bytes": (SZ, [I, C.POINTER(GemmProblem)]),
    "b200_gemm_grouped_streamk_bf16": (I, [I, C.POINTER(GemmProblem), P, SZ, P]),
    "b200_gemm_bf16": (I, [P, I64, I, P, I64, I, P, I64, I, I, F, I64, I64, I64, P]),
    "b200_rmsnorm_fwd_bf16": (I, [P, P, P, P, P, I64, I64, F, I, P]),
    "b200_rmsnorm_bwd_bf16": (I, [P, P, P, P, P, P, I64, I64, F, P]),
    "b200_rmsnorm_fwd_f32": (I, [P, P, P, I64, I64, F, P]),
    "b200_rmsnorm_bwd_f32": (I, [P, P, P, P, P, I64, I64, F, P]),
    "b200_layernorm_fwd_bf16": (I, [P, P, P, P, P, P, I64, I64, F, P]),
    "b200_layernorm_bwd_bf16": (I, [P, P, P, P, P, P, P, P, I64, I64, P]),
    "b200_layernorm_bwd_res_bf16": (I, [P, P, P, P, P, P, P, P, P, I64, I64, P]),
    "b200_swiglu_fwd_bf16": (I, [P, I64, P, I64, P, I64, I64, I, P]),
    "b200_swiglu_bwd_bf16": (I, [P, I64, P, I64, P, P, I64, P, I64, I64, I64, P]),
    "b200_gelu_fwd_bf16": (I, [P, P, I64, P]),
    "b200_gelu_bwd_bf16": (I, [P, P, P, I64, P]),
    "b200_rope_bf16": (I, [P, I64, P, I64, P, P, I64, I64, I64, I64, I, I, P]),
    "b200_embedding_fwd_bf16": (I, [P, P, P, P, I64, I64, I64, I64, P]),
    "b200_embedding_bwd_f32acc": (I, [P, P, P, P, I64, I64, I64, I64, P]),
    "b200_cross_entropy_workspace_bytes": (SZ, [I64]),
    "b200_cross_entropy_bf16": (I, [P, I64, P, P, I64, I64, I, P, SZ, P]),
    "b200_compact_rows": (I, [P, I64, I64, P, P, P]),
    "b200_gather_rows_bf16": (I, [P, I64, P, P, P, P, P, I64, I64, P]),
    "b200_scatter_rows_bf16": (I, [P, P, P, I64, I64, I64, P]),
    "b200_cross_entropy_bf16_noalloc": (I, [P, I64, P, P, I64, I64, I, P]),
    "b200_grad_norm_workspace_bytes": (SZ, [I64]),
    "b200_grad_norm_bf16": (I, [P, I64, F, P, P, SZ, P]),
    "b200_grad_norm_f32": (I, [P, I64, F, P, P, SZ, P]),
    "b200_adam_hyper": (I, [P, F, F, F, F, F, I64]),
    "b200_adamw_bf16": (I, [P, P, I, P, P, I64, F, F, F, F, F, I64, P, P, P]),
    "b200_adamw_f32": (I, [P, P, P, P, I64, F, F, F, F, F, I64, P, P, P]),
    "b200_adamw_master_f32": (I, [P, P, P, P, P, I64, F, F, F, F, F, I64, P, P, P]),
    "b200_cast_f32_to_bf16": (I, [P, P, I64, P]),
    "b200_cast_bf16_to_f32": (I, [P, P, I64, P]),
    "b200_transpose_bf16": (I, [P, I64, P, I64, I64, I64, F, P]),
    "b200_add_bf16": (I, [P, P, P, I64, P]),
    "b200_colsum_bf16": (I, [P, I64, P, I64, I64, P]),
    "b200_attention_fwd_bf16": (I, [P, I64, P, I64, P, I64, P, I64, P, I64, I64, I64, I64, I64, F, F, C.c_uint64, P, P]),
    "b200_attention_dropout_mask_u8": (I, [P, I64, I64, I64, F, C.c_uint64, P]),
    "b200_dropout_bf16": (I, [P, P, P, I64, F, C.c_uint64, P, P]),
    "b200_scale_bf16_dev": (I, [P, I64, P, P]),
    "b200_dropout_mask_u8": (I, [P, I64, F, C.c_uint64, P]),
    "b200_collate_i64": (I, [P, P, P, I64, I64, I64, I64, I64, P, P, P]),
    "b200_window_batch_i64": (I, [P, I64, P, I64, I64, P, P, P]),
    "b200_gemm_f32": (I, [P, I64, I64, I64, P, I64, I64, I64, P, P, I64, I64, I64, I64, I64, I64, I64, F, F, P]),
    "b200_softmax_causal_f32": (I, [P, I64, I64, F, P]),
    "b200_softmax_bwd_f32": (I, [P, P, I64, I64, F, P]),
    "b200_layernorm_fwd_f32": (I, [P, P, P, P, P, P, I64, I64, F, P]),
    "b200_layernorm_bwd_f32": (I, [P, P, P, P, P, P, P, P, I64, I64, P]),
    "b200_elementwise_f32": (I, [P, P, P, I64, I, P]),
    "b200_swiglu_bwd_f32": (I, [P, P, P, P, P, I64, P]),
    "b200_rope_f32": (I, [P, P, P, P, I64, I64, I64, I64, I, P]),
    "b200_cross_entropy_f32_workspace_bytes": (SZ, [I64]),
    "b200_cross_entropy_f32": (I, [P, I64, P, P, I64, I64, I, P, SZ, P]),
    "b200_embedding_fwd_f32": (I, [P, P, P, P, I64, I64, I64, P]),
    "b200_embedding_bwd_f32": (I, [P, P, P, P, I64, I64, I64, P]),
    "b200_attention_decode_bf16": (I, [P, I64, P, P, I64, I64, P, I64, I64, I64, I64, I64, I64, F, P]),
    "b200_attention_bwd_workspace_bytes": (SZ, [I64, I64, I64, I64]),
    "b200_attention_bwd_rope_bf16": (I, [P, I64, P, I64, P, I64, P, I64, P, I64, P, P, SZ, P, I64, P, I64, P,
                                         I64, I64, I64, I64, I64, I64, F, P, P, P]),
    "b200_attention_bwd_bf16": (I, [P, I64, P, I64, P, I64, P, I64, P, I64, P, P, SZ, P, I64, P, I64, P,
                                    I64, I64, I64, I64, I64, I64, F, F, C.c_uint64, P, P]),
}


def header_symbols(path: str = HEADER_PATH) -> list[str]:
    """All function names declared in include/b200llm.h."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text)))


def load() -> C.CDLL:
    """Load libb200llm.so, binding every signature. Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU or library fallback for the b200llm ops)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class B200Error(RuntimeError):
    pass


def check(status: int, what: str = "") -> None:
    if status != 0:
        msg = load().b200_last_error().decode(errors="replace")
        raise B200Error(f"b200llm {what} failed with status {status}: {msg}")


GEMM_PATHS = ("pair", "pair_tail_split", "single_cta", "grouped_tiles", "streamk", "packed_segments", "swiglu_fwd",
              "swiglu_bwd", "skinny_pair")


def gemm_path_counts() -> dict:
    """Launch counts of the GEMM scheduling paths so far (b200_gemm_path_counts)."""
    out = (C.c_uint64 * len(GEMM_PATHS))()
    load().b200_gemm_path_counts(out, len(GEMM_PATHS))
    return dict(zip(GEMM_PATHS, [int(v) for v in out]))


GEMM_OPT_TAIL_SPLIT, GEMM_OPT_STREAMK, GEMM_OPT_N_FASTEST, GEMM_OPT_SKINNY_PAIR = 0, 1, 2, 3


def gemm_set_option(option: int, value: int) -> None:
    """-1 = heuristic, 0 = off, 1 = forced (b200_gemm_set_option)."""
    check(load().b200_gemm_set_option(option, value), "gemm_set_option")


def launch_count() -> int:
    return int(load().b200_launch_count())
