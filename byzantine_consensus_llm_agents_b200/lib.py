"""ctypes binding of libbcg_b200.so (include/bcg_b200.h).

There is no CPU or PyTorch fallback: if the library is missing, or a call fails, this module
raises.  PyTorch only owns device memory and streams; all arithmetic on the product path
happens inside the library's sm_100a kernels.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "_lib" / "libbcg_b200.so"

EPI_BF16, EPI_F32, EPI_SWIGLU = 0, 1, 2
GEMM_PDL = 1
STEP_FUSED_NORM = 0x20000   # bcg_model_decode: fold add+RMSNorm into the GEMMs (opt-in, slower)
STEP_FUSED_ROPE = 0x80000     # bcg_model_decode: qk-norm/RoPE/KV-write folded into decode attention
STEP_DEFERRED_NORM = 0x40000  # bcg_model_decode: norm statistics kernel + 1/RMS in the next GEMM's epilogue

c_int_p = C.POINTER(C.c_int)
c_float_p = C.POINTER(C.c_float)
vp = C.c_void_p


class BcgError(RuntimeError):
    pass


class SamplerArgs(C.Structure):
    _fields_ = [
        ("logits", vp), ("ld_logits", C.c_longlong),
        ("B", C.c_int), ("V", C.c_int), ("mask_words", C.c_int), ("max_steps", C.c_int),
        ("eos_id0", C.c_int), ("eos_id1", C.c_int), ("page_size", C.c_int),
        ("max_blocks_per_seq", C.c_int),
        ("grammar_id", vp), ("mask_tables", vp), ("trans_tables", vp),
        ("tok_bytes", vp), ("tok_offsets", vp),
        ("temperature", vp), ("seed", vp), ("max_tokens", vp), ("forced", vp),
        ("state", vp), ("n_out", vp), ("finished", vp), ("out_tokens", vp), ("chosen", vp),
        ("next_ids", vp), ("positions", vp), ("seq_lens", vp), ("slot_map", vp),
        ("block_table", vp), ("n_active", vp),
        ("shard_v", C.c_int), ("shard_stride", C.c_longlong),
        ("tp_rank", C.c_int), ("tp_size", C.c_int), ("vocab_base", C.c_int),
    ]


class GemmFusion(C.Structure):
    _fields_ = [
        ("scale_sumsq", vp), ("scale_tiles", C.c_int), ("scale_d", C.c_int), ("scale_eps", C.c_float),
        ("red_residual", vp), ("red_w", vp), ("red_xn", vp), ("red_sumsq", vp), ("red_counters", vp),
    ]


class Model(C.Structure):
    _fields_ = [
        ("n_layers", C.c_int), ("d_model", C.c_int), ("n_heads", C.c_int), ("n_kv_heads", C.c_int),
        ("head_dim", C.c_int), ("d_ff", C.c_int), ("vocab", C.c_int),
        ("rms_eps", C.c_float), ("page_size", C.c_int),
        ("embed", vp), ("lm_head", vp), ("final_norm", vp),
        ("ln1", C.POINTER(vp)), ("ln2", C.POINTER(vp)),
        ("w_qkv", C.POINTER(vp)), ("w_o", C.POINTER(vp)),
        ("w_gate_up", C.POINTER(vp)), ("w_down", C.POINTER(vp)),
        ("q_norm", C.POINTER(vp)), ("k_norm", C.POINTER(vp)),
        ("kv_cache", C.POINTER(vp)), ("cos_sin", vp), ("tp_size", C.c_int),
    ]


class Workspace(C.Structure):
    _fields_ = [
        ("residual", vp), ("xn", vp), ("qkv", vp), ("q", vp), ("attn", vp), ("act", vp),
        ("delta", vp), ("partials", vp), ("partials_floats", C.c_longlong),
        ("attn_scratch_o", vp), ("attn_scratch_ml", vp), ("attn_counters", vp),
        ("max_splits", C.c_int),
        ("logits", vp), ("logits_local", vp), ("tp_region_floats", C.c_longlong), ("xn_last", vp),
        ("T_max", C.c_int), ("B_max", C.c_int),
        ("row_sumsq", vp), ("tile_counters", vp),
    ]


_SIGS = {
    "bcg_abi_version": (C.c_int, []),
    "bcg_last_error": (C.c_char_p, []),
    "bcg_gemm_bf16_tn": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_longlong,
                                   C.c_longlong, C.c_longlong, C.c_int, C.c_int, C.c_int, vp]),
    "bcg_gemm_bf16_tn_fused": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_longlong,
                                         C.c_longlong, C.c_longlong, C.c_int, C.c_int, C.c_int,
                                         C.POINTER(GemmFusion), vp]),
    "bcg_gemm_splits_for": (C.c_int, [C.c_int, C.c_int, C.c_int]),
    "bcg_embed_rmsnorm": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_float, vp]),
    "bcg_add_rmsnorm": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_longlong, vp, vp, vp, C.c_int,
                                  C.c_int, C.c_float, vp]),
    "bcg_add_norm_stats": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_longlong, vp, vp, vp, C.c_int, C.c_int, vp]),
    "bcg_qknorm_rope_kvwrite": (C.c_int, [vp, C.c_int, C.c_int, C.c_longlong, C.c_longlong, vp, vp,
                                          vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int,
                                          C.c_float, vp]),
    "bcg_swiglu_reduce": (C.c_int, [vp, C.c_int, C.c_longlong, vp, C.c_int, C.c_int, vp]),
    "bcg_decode_attn_paged": (C.c_int, [vp, vp, vp, C.c_int, vp, vp, vp, vp, vp, C.c_int, C.c_int,
                                        C.c_int, C.c_int, C.c_int, C.c_float, vp]),
    "bcg_decode_attn_paged_qkv": (C.c_int, [vp, C.c_int, C.c_longlong, C.c_longlong, vp, vp, vp, vp, vp, C.c_float,
                                            vp, vp, C.c_int, vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int,
                                            C.c_int, C.c_int, C.c_float, vp]),
    "bcg_decode_attn_paged_static": (C.c_int, [vp, vp, vp, C.c_int, vp, vp, C.c_int, C.c_int, C.c_int,
                                               C.c_int, C.c_int, C.c_float, vp]),
    "bcg_prefill_attn_paged": (C.c_int, [vp, vp, vp, C.c_int, vp, C.c_int, vp, C.c_int, C.c_int, C.c_int,
                                         C.c_int, C.c_float, vp]),
    "bcg_mask_sample": (C.c_int, [C.POINTER(SamplerArgs), vp]),
    "bcg_tp_unique_id": (C.c_int, [vp]),
    "bcg_tp_init": (C.c_int, [vp, C.c_int, C.c_int]),
    "bcg_tp_shutdown": (C.c_int, []),
    "bcg_tp_symm_bytes_for": (C.c_longlong, [C.c_longlong, C.c_int]),
    "bcg_tp_symm_create": (C.c_int, [C.c_longlong, vp]),
    "bcg_tp_add_norm_stats": (C.c_int, [vp, vp, C.c_int, C.c_longlong, vp, vp, vp, C.c_int, C.c_longlong,
                                        C.c_int, C.c_int, vp]),
    "bcg_tp_symm_open": (C.c_int, [vp, C.c_int, C.c_int]),
    "bcg_model_prefill": (C.c_int, [C.POINTER(Model), C.POINTER(Workspace), vp, vp, vp, vp, C.c_int,
                                    vp, C.c_int, vp, C.c_int, C.c_int, vp]),
    "bcg_model_decode": (C.c_int, [C.POINTER(Model), C.POINTER(Workspace), vp, vp, vp, vp, vp,
                                   C.c_int, C.c_int, C.c_int, vp]),
}

EXPORTED = tuple(_SIGS)

_lib = None


def load(build_if_missing: bool = False) -> C.CDLL:
    """dlopen the in-tree library; raise (never fall back) if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        if build_if_missing:
            from . import build as _build
            _build.build()
        else:
            raise BcgError(
                f"{LIB_PATH} is missing: run `python -m byzantine_consensus_llm_agents_b200.build` "
                "(there is no CPU fallback)")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().bcg_last_error()
        raise BcgError(f"{what or 'bcg call'} failed (status {rc}): "
                       f"{msg.decode() if msg else ''}")


def ptr(t) -> int:
    """Device pointer of a torch tensor (or None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream_ptr() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


# ---- thin per-kernel wrappers (tensors in, status checked) ---------------------------------

def gemm(x, w, out, epilogue=EPI_BF16, splits=1, flags=0, M=None):
    """out = x @ w.T through bcg_gemm_bf16_tn.  x [M,K] bf16, w [N,K] bf16."""
    lib = load()
    M = x.shape[0] if M is None else M
    K = x.shape[1]
    N = w.shape[0]
    ld_out = out.stride(-2)
    check(lib.bcg_gemm_bf16_tn(ptr(x), ptr(w), ptr(out), M, N, K, x.stride(0), w.stride(0), ld_out,
                               epilogue, splits, flags, stream_ptr()), "bcg_gemm_bf16_tn")
    return out


def splits_for(M, N, K) -> int:
    return load().bcg_gemm_splits_for(M, N, K)
