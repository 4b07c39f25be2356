"""ctypes binding of libecho_b200.so (C ABI declared in include/echo_b200.h).

The library holds every kernel of the product path.  There is NO fallback: if the shared object is
missing or a call is made without a CUDA device, this module raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libecho_b200.so")
if os.environ.get("ECHO_B200_LIB"):                 # profiling tools only: "debug" = the -DECHO_DEBUG build (in-kernel
    LIB_PATH = os.path.join(_HERE, "libecho_b200_%s.so" % os.environ["ECHO_B200_LIB"])   # timelines); other names = A/B builds

c_p = C.c_void_p
i32, i64, f32 = C.c_int32, C.c_int64, C.c_float


class GemmSection(C.Structure):
    _fields_ = [("type", i32), ("_pad", i32), ("out", c_p), ("norm_weight", c_p)]


class GemmDesc(C.Structure):
    _fields_ = [
        ("a", c_p), ("b", c_p), ("lda", i64), ("a_batch_stride", i64), ("ldb", i64),
        ("rows", i32), ("batch", i32), ("n", i32), ("k_tap", i32), ("taps", i32),
        ("tap_shift", i32 * 8), ("epilogue", i32), ("impl", i32), ("block_n", i32), ("act", i32),
        ("bias", c_p), ("gate", c_p), ("gate_ld", i64), ("rows_per_gate", i32), ("k_splits", i32),
        ("resid", c_p), ("ld_resid", i64), ("out_f32", c_p), ("ld_out_f32", i64),
        ("out_bf16", c_p), ("ld_out_bf16", i64), ("out2_bf16", c_p), ("ld_out2_bf16", i64),
        ("snake_alpha", c_p),
        ("n_sections", i32), ("section_width", i32), ("head_dim", i32), ("rope_heads", i32),
        ("pos_offset", i32), ("pos_stride", i32), ("rows_per_seq", i32), ("a_ctx", i32),
        ("eps", f32), ("b_static", i32), ("rope_cos", c_p), ("rope_sin", c_p), ("ld_sec_out", i64),
        ("sections", GemmSection * 4), ("split_stride", i64),
        ("workspace", c_p), ("workspace_bytes", i64),
    ]


class SgemmDesc(C.Structure):
    _fields_ = [
        ("a", c_p), ("lda", i64), ("a_zstride", i64),
        ("w", c_p), ("ldw", i64), ("w_zstride", i64),
        ("bias", c_p), ("bias_zstride", i64),
        ("addend", c_p), ("ld_addend", i64), ("addend_zstride", i64),
        ("out", c_p), ("ldo", i64), ("out_zstride", i64),
        ("out_bf16", c_p), ("ldo_bf16", i64), ("out_bf16_zstride", i64),
        ("m", i32), ("n", i32), ("k", i32), ("nz", i32),
        ("pre_silu", i32), ("post", i32), ("out_scale", f32), ("_pad", f32),
    ]


class AttnSegment(C.Structure):
    _fields_ = [
        ("k", c_p), ("v", c_p), ("row_stride", i64), ("batch_stride", i64),
        ("len", i32), ("kv_batch_mod", i32), ("mask", c_p), ("mask_stride", i64),
        ("extent", c_p), ("kv_scale", c_p), ("causal", i32), ("window", i32), ("q_offset", i32), ("_pad", i32),
    ]


class AttnDesc(C.Structure):
    _fields_ = [
        ("q", c_p), ("q_row_stride", i64), ("q_batch_stride", i64), ("gate", c_p), ("out", c_p),
        ("rows", i32), ("seq", i32), ("heads", i32), ("head_dim", i32),
        ("n_seg", i32), ("impl", i32), ("softmax_scale", f32), ("kv_splits", i32),
        ("seg", AttnSegment * 4),
        ("workspace", c_p), ("workspace_bytes", i64),
    ]


class EncLayer(C.Structure):
    _fields_ = [("w_qkvg", c_p), ("w_o", c_p), ("w13", c_p), ("w2", c_p), ("q_norm", c_p), ("k_norm", c_p),
                ("attn_norm", c_p), ("mlp_norm", c_p)]


class EncodeKVDesc(C.Structure):
    _fields_ = [
        ("n_layers", i32), ("d", i32), ("heads", i32), ("ffn", i32), ("eps", f32), ("causal", i32),
        ("layers", C.POINTER(EncLayer)), ("out_norm", c_p), ("rope_cos", c_p), ("rope_sin", c_p), ("rope_len", i32),
        ("rows", i32), ("seq", i32), ("past", i32), ("x", c_p), ("mask", c_p), ("mask_stride", i64), ("extent", c_p),
        ("enc_cache", c_p), ("enc_cap", i32), ("dec_layers", i32), ("dec_d", i32), ("dec_heads", i32),
        ("w_kv", C.POINTER(c_p)), ("k_norm", C.POINTER(c_p)), ("rope_heads", i32), ("pos_stride", i32),
        ("kv_out", c_p), ("out_cap", i32), ("gemm_impl", i32), ("gemm_workspace", c_p), ("gemm_workspace_bytes", i64),
    ]


class AeTrLayer(C.Structure):
    _fields_ = [("wqkv", c_p), ("wo", c_p), ("w13", c_p), ("w2", c_p), ("attn_norm", c_p), ("ffn_norm", c_p),
                ("attn_scale", c_p), ("ffn_scale", c_p)]


class AeUpsample(C.Structure):
    _fields_ = [("factor", i32), ("hidden", i32), ("wt", c_p), ("bt", c_p), ("dw", c_p), ("dw_b", c_p), ("ln_w", c_p),
                ("ln_b", c_p), ("pw1", c_p), ("pb1", c_p), ("pw2", c_p), ("pb2", c_p), ("gamma", c_p)]


class AeUnit(C.Structure):
    _fields_ = [("dilation", i32), ("_pad", i32), ("alpha1", c_p), ("alpha2", c_p), ("w7", c_p), ("b7", c_p),
                ("w1", c_p), ("b1", c_p)]


class AeBlock(C.Structure):
    _fields_ = [("rate", i32), ("c_in", i32), ("c_out", i32), ("_pad", i32), ("alpha_in", c_p), ("alpha1_rep", c_p),
                ("wt", c_p), ("bt", c_p), ("units", AeUnit * 3)]


class AeDecoderDesc(C.Structure):
    _fields_ = [
        ("d", i32), ("tr_layers", i32), ("tr_heads", i32), ("tr_head_dim", i32), ("tr_ffn", i32), ("tr_window", i32),
        ("tr_eps", f32), ("_pad0", f32), ("tr", C.POINTER(AeTrLayer)), ("tr_norm", c_p), ("rope_cos", c_p),
        ("rope_sin", c_p), ("rope_len", i32), ("n_up", i32), ("ups", C.POINTER(AeUpsample)), ("conv0_w", c_p),
        ("conv0_b", c_p), ("c0", i32), ("n_blocks", i32), ("blocks", C.POINTER(AeBlock)), ("alpha_out", c_p),
        ("w_last", c_p), ("b_last", f32), ("gemm_impl", i32), ("gemm_workspace", c_p), ("gemm_workspace_bytes", i64),
    ]


class TailDesc(C.Structure):
    _fields_ = [
        ("h", c_p), ("ldh", i64), ("norm_w", c_p), ("w_out", c_p), ("b_out", c_p),
        ("x_t", c_p), ("v_out", c_p), ("v_comb_out", c_p), ("sched", c_p), ("step", c_p),
        ("positions", i32), ("d", i32), ("latent", i32), ("branches", i32),
        ("eps", f32), ("cfg_text", f32), ("cfg_speaker", f32), ("rescale_k", f32), ("rescale_sigma", f32),
        ("use_rescale", i32), ("update_x", i32), ("n_delta", i32), ("delta", c_p), ("delta_stride", i64),
        ("v_peer", c_p * 2), ("flag_peer", c_p * 2), ("wait_flag", c_p * 2), ("wait_value", i32), ("_pad_tail", i32),
        ("guidance", c_p),
    ]


class DitLayer(C.Structure):
    _fields_ = [("w_qkvg", c_p), ("w_o", c_p), ("w13", c_p), ("w2", c_p), ("q_norm", c_p), ("k_norm", c_p)]


class DitKV(C.Structure):
    _fields_ = [
        ("k", C.POINTER(c_p)), ("v", C.POINTER(c_p)), ("row_stride", i64), ("batch_stride", i64),
        ("len", i32), ("kv_batch_mod", i32), ("mask", c_p), ("mask_stride", i64), ("extent", c_p), ("kv_scale", c_p),
    ]


class DitStepDesc(C.Structure):
    _fields_ = [
        ("n_layers", i32), ("d", i32), ("heads", i32), ("head_dim", i32), ("ffn", i32), ("latent", i32),
        ("eps", f32), ("rope_len", i32),
        ("layers", C.POINTER(DitLayer)), ("in_w", c_p), ("in_b", c_p), ("rope_cos", c_p), ("rope_sin", c_p),
        ("x_in", c_p), ("x_rows", i32), ("rows", i32), ("seq", i32), ("pos_offset", i32),
        ("mod", c_p), ("mod_row_stride", i64), ("mod_rows", i32), ("k_splits", i32),
        ("prefix", DitKV), ("text", DitKV), ("speaker", DitKV),
        ("x", c_p), ("xn", c_p), ("qkvg", c_p), ("ao", c_p), ("hmid", c_p), ("delta", c_p),
        ("gemm_workspace", c_p), ("gemm_workspace_bytes", i64), ("gemm_impl", i32), ("attn_impl", i32),
        ("tail", C.POINTER(TailDesc)), ("pending_out", C.POINTER(i32)),
        ("attn_workspace", c_p), ("attn_workspace_bytes", i64),
    ]


EPI_GENERIC, EPI_SECTIONS, EPI_SWIGLU = 0, 1, 2
ACT_NONE, ACT_GELU, ACT_SILU = 0, 1, 2
SEC_PLAIN, SEC_NORM_ROPE, SEC_SIGMOID, SEC_ROPE = 0, 1, 2, 3
POST_NONE, POST_PLUS1, POST_TANH, POST_SILU = 0, 1, 2, 3

EXPORTS = {
    "echo_last_error": (C.c_char_p, []),
    "echo_version": (C.c_int, []),
    "echo_launch_count": (C.c_ulonglong, []),
    "echo_reset_launch_count": (None, []),
    "echo_gemm_bf16": (C.c_int, [C.POINTER(GemmDesc), c_p]),
    "echo_gemm_workspace_bytes": (C.c_longlong, []),
    "echo_gemm_bf16_splitk_noalloc": (C.c_int, [C.POINTER(GemmDesc), c_p]),
    "echo_noalloc_scratch_bytes": (C.c_longlong, []),
    "echo_debug_gemm_trace": (C.c_int, [c_p]),
    "echo_debug_attn_trace": (C.c_int, [c_p]),
    "echo_sgemm_f32": (C.c_int, [C.POINTER(SgemmDesc), c_p]),
    "echo_timestep_embedding": (C.c_int, [c_p, C.c_int, C.c_int, c_p, c_p]),
    "echo_rmsnorm_mod": (C.c_int, [c_p, i64, C.c_int, C.c_int, f32, c_p, c_p, i64, C.c_int, c_p, i64, c_p, i64, c_p]),
    "echo_add_rmsnorm_mod": (C.c_int, [c_p, i64, C.c_int, C.c_int, f32, c_p, C.c_int, i64, c_p, c_p, i64, C.c_int, c_p, i64, c_p]),
    "echo_attention": (C.c_int, [C.POINTER(AttnDesc), c_p]),
    "echo_attention_workspace_bytes": (C.c_longlong, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "echo_mask_extent": (C.c_int, [c_p, C.c_int, C.c_int, i64, c_p, c_p]),
    "echo_cfg_euler_tail": (C.c_int, [C.POINTER(TailDesc), c_p]),
    "echo_cfg_combine_euler": (C.c_int, [C.POINTER(TailDesc), c_p]),
    "echo_cfg_tail_ctas": (C.c_int, [C.c_int]),
    "echo_dit_step": (C.c_int, [C.POINTER(DitStepDesc), c_p]),
    "echo_dit_sample": (C.c_int, [C.POINTER(DitStepDesc), C.POINTER(DitStepDesc), C.c_int, c_p, i64, i64, i64, c_p]),
    "echo_step_advance": (C.c_int, [c_p, c_p]),
    "echo_embedding": (C.c_int, [c_p, C.c_int, c_p, C.c_int, c_p, c_p]),
    "echo_dwconv_layernorm": (C.c_int, [c_p, C.c_int, C.c_int, C.c_int, c_p, c_p, c_p, c_p, f32, c_p, c_p]),
    "echo_conv_out_tanh": (C.c_int, [c_p, C.c_int, C.c_int, C.c_int, c_p, f32, c_p, c_p]),
    "echo_dwconv_layernorm_ctx": (C.c_int, [c_p, C.c_int, C.c_int, C.c_int, C.c_int, c_p, c_p, c_p, c_p, f32, c_p, c_p]),
    "echo_conv_out_tanh_ctx": (C.c_int, [c_p, C.c_int, C.c_int, C.c_int, C.c_int, c_p, f32, c_p, c_p]),
    "echo_encode_kv_workspace_bytes": (i64, [C.POINTER(EncodeKVDesc)]),
    "echo_encode_kv": (C.c_int, [C.POINTER(EncodeKVDesc), c_p, i64, c_p]),
    "echo_ae_decode_workspace_bytes": (i64, [C.POINTER(AeDecoderDesc), C.c_int, C.c_int, C.c_int]),
    "echo_ae_decode": (C.c_int, [C.POINTER(AeDecoderDesc), c_p, C.c_int, C.c_int, c_p, c_p, i64, c_p]),
    "echo_ae_decode_stream_state_bytes": (i64, [C.POINTER(AeDecoderDesc)]),
    "echo_ae_decode_stream_init": (C.c_int, [C.POINTER(AeDecoderDesc), c_p, i64, c_p]),
    "echo_ae_decode_stream_step": (C.c_int, [C.POINTER(AeDecoderDesc), c_p, c_p, C.c_int, C.c_int, c_p, c_p, i64, c_p]),
    "echo_cast_snake": (C.c_int, [c_p, i64, C.c_int, c_p, c_p, c_p]),
    "echo_conv_in_snake": (C.c_int, [c_p, C.c_int, C.c_int, C.c_int, c_p, c_p, c_p, c_p, c_p, c_p]),
    "echo_snake_bf16": (C.c_int, [c_p, i64, C.c_int, c_p, c_p, c_p]),
    "echo_rmsnorm_f32": (C.c_int, [c_p, i64, C.c_int, C.c_int, f32, c_p, c_p, i64, c_p]),
    "echo_vq_nearest": (C.c_int, [c_p, C.c_int, C.c_int, c_p, c_p, C.c_int, c_p, c_p, c_p]),
    "echo_split_bf16x3": (C.c_int, [c_p, i64, i64, C.c_int, C.c_int, c_p, c_p, i64, c_p]),
    "echo_rope_f32": (C.c_int, [c_p, i64, i64, C.c_int, C.c_int, c_p, c_p, C.c_int, C.c_int, C.c_int, c_p]),
    "echo_attention_window_f32": (C.c_int, [c_p, c_p, c_p, i64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, f32, c_p, i64, c_p]),
    "echo_dwconv_layernorm_f32": (C.c_int, [c_p, C.c_int, C.c_int, C.c_int, c_p, c_p, c_p, c_p, f32, c_p, c_p]),
    "echo_resample_mono": (C.c_int, [c_p, C.c_int, i64, i64, C.c_int, C.c_int, C.c_int, c_p, c_p, c_p, c_p, i64, c_p, c_p]),
    "echo_peak_normalize": (C.c_int, [c_p, i64, c_p, f32, c_p]),
    "echo_flattening_point": (C.c_int, [c_p, C.c_int, C.c_int, C.c_int, f32, f32, c_p, c_p]),
}

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """dlopen the kernel library (built in-tree by ``__graft_entry__.build()``); fail loudly."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: the CUDA extension is required (no CPU fallback). "
                "Run `python __graft_entry__.py build` (needs nvcc, sm_100a).")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in EXPORTS.items():
            fn = getattr(lib, name)          # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise RuntimeError(f"echo_b200 error {rc}: {load().echo_last_error().decode()}")


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    """Device pointer of a CUDA tensor (None -> NULL).  Non-CUDA tensors are refused: no CPU path."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("echo_b200: expected a CUDA tensor (there is no CPU fallback)")
    return t.data_ptr()


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream
