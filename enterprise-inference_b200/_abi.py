"""ctypes binding of include/b200_decode.h.  Fails loudly when the CUDA extension is missing:
there is no CPU fallback on the product path."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200_LIB") or os.path.join(HERE, "lib", "libb200decode.so")   # B200_LIB: A/B builds


class B200Error(RuntimeError):
    def __init__(self, code: int, text: str):
        super().__init__(f"b200 error {code}: {text}")
        self.code = code


class ModelCfg(C.Structure):
    """struct b200_model_cfg"""
    _fields_ = [(n, C.c_int32) for n in ("hidden", "n_layers", "n_q_heads", "n_kv_heads", "head_dim",
                                         "intermediate", "vocab", "qkv_bias")] + \
               [("rms_eps", C.c_float)] + \
               [(n, C.c_int32) for n in ("block_size", "max_tokens", "max_blocks_per_seq", "tp_rank", "tp_world")]


W_EMBED, W_LN1, W_QKV, W_QKV_BIAS, W_O, W_LN2, W_GATE_UP, W_DOWN, W_FINAL_NORM, W_LM_HEAD, W_GATE_UP_INTERLEAVED = range(11)

_vp, _i32, _i64, _f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float

# name -> (restype, argtypes); every symbol declared in include/b200_decode.h
SIGNATURES = {
    "b200_last_error": (C.c_char_p, []),
    "b200_abi_version": (C.c_int, []),
    "b200_device_ok": (C.c_int, []),
    "b200_ctx_create": (C.c_int, [C.POINTER(ModelCfg), C.POINTER(_vp)]),
    "b200_ctx_destroy": (C.c_int, [_vp]),
    "b200_ctx_workspace_bytes": (_i64, [_vp]),
    "b200_set_weight": (C.c_int, [_vp, _i32, _i32, _vp, _i64, _i64]),
    "b200_set_kv_cache": (C.c_int, [_vp, _vp, _vp, _i64]),
    "b200_set_rope_cache": (C.c_int, [_vp, _vp, _i64]),
    "b200_nccl_unique_id": (C.c_int, [_vp]),
    "b200_nccl_init": (C.c_int, [_vp, _vp]),
    "b200_tp_ipc_export": (C.c_int, [_vp, _vp]),
    "b200_tp_ipc_import": (C.c_int, [_vp, _vp, _i32]),
    "b200_decode_step": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _i32, _vp, _i32, _vp, _vp, _vp]),
    "b200_decode_step_host": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp, _vp]),
    "b200_set_launch_pdl": (C.c_int, [_i32]),
    "b200_ctx_set_graphs": (C.c_int, [_vp, _i32]),
    "b200_ctx_last_step_launches": (C.c_int, [_vp]),
    "b200_ctx_set_tap": (C.c_int, [_vp, _i32, _vp, _vp]),
    "b200_linear_workspace_bytes": (_i64, [_i32, _i32, _i32]),
    "b200_linear_bf16": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp, _i64, _vp]),
    "b200_linear_fused_bf16": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp,
                                         _i32, _i32, _i32, _vp, _i64, _vp]),
    "b200_embed_stream": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _vp]),
    "b200_linear_residual_bf16": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _vp, _i64, _vp]),
    "b200_linear_normed_bf16": (C.c_int, [_vp, _vp, _vp, _f32, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp, _vp, _vp,
                                          _vp, _vp, _i32, _i32, _i32, _vp, _i64, _vp]),
    "b200_rmsnorm_stream": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _f32, _vp]),
    "b200_rmsnorm": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _f32, _vp]),
    "b200_embed": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _vp]),
    "b200_rope_kv_append": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "b200_paged_attn_workspace_bytes": (_i64, [_i32, _i32, _i32, _i32]),
    "b200_paged_attn_decode": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _i64, _i32, _i32, _i32,
                                         _i32, _i32, _f32, _vp, _i64, _vp]),
    "b200_paged_attn_prefill": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _i64, _i32, _i32, _i32,
                                          _i32, _i32, _f32, _vp, _i64, _vp]),
    "b200_paged_attn_static_scratch_bytes": (_i64, []),
    "b200_paged_attn_decode_static": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _i64, _i32, _i32, _i32,
                                                _i32, _i32, _f32, _vp]),
    "b200_silu_mul": (C.c_int, [_vp, _vp, _i32, _i32, _vp]),
    "b200_argmax": (C.c_int, [_vp, _i64, _i32, _i32, _i32, _vp, _vp, _vp]),
    "b200_sample": (C.c_int, [_vp, _i64, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b200_ctx_sample_last": (C.c_int, [_vp, _i32, _vp, _vp, _vp]),
    "b200_logit_adjust": (C.c_int, [_vp, _i64, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b200_ctx_adjust_last": (C.c_int, [_vp, _i32, _vp, _i32, _vp]),
    "b200_logprobs": (C.c_int, [_vp, _i64, _i32, _i32, _vp, _i32, _vp, _vp, _vp, _vp]),
    "b200_ctx_logprobs_last": (C.c_int, [_vp, _i32, _vp, _i32, _vp, _vp]),
    "b200_advance_decode_rows": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _i32, _vp]),
    "b200_allreduce_bf16": (C.c_int, [_vp, _vp, _i64, _vp]),
    "b200_tp_allreduce": (C.c_int, [_vp, _i32, _vp, _i32, _vp, _vp, _vp, _vp]),
}

_lib = None


def lib() -> C.CDLL:
    """Load libb200decode.so (built in-tree by build.py).  No fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `python enterprise-inference_b200/build.py` "
                              "(nvcc, sm_100a). There is no CPU fallback.")
        _lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(_lib, name)
            fn.restype, fn.argtypes = res, args
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise B200Error(rc, lib().b200_last_error().decode(errors="replace"))


def ptr(t) -> int:
    """Device (or pinned host) address of a torch tensor, None -> NULL."""
    return None if t is None else t.data_ptr()
