"""ctypes binding of the C-ABI library (include/tha_b200.h).

This is the binding a Python host would add; tests and bench.py call the
CUDA path exclusively through it.  There is NO fallback: if the library is missing or a
symbol is absent, importing / calling fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os
import re

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
LIB_PATH = os.path.join(PKG, "lib", "libtha_b200.so")
HEADER = os.path.join(ROOT, "include", "tha_b200.h")

f32p = C.POINTER(C.c_float)
i32p = C.POINTER(C.c_int)
i8p = C.POINTER(C.c_int8)
vp = C.c_void_p


class Config(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("dim", "hidden_dim", "n_layers", "n_heads", "n_kv_heads", "vocab_size", "seq_len")]


class WeightsF32(C.Structure):
    _fields_ = [(n, vp) for n in ("token_embedding_table", "rms_att_weight", "rms_ffn_weight", "wq", "wk", "wv", "wo",
                                  "w1", "w2", "w3", "rms_final_weight", "wcls")]


class State(C.Structure):
    _fields_ = [(n, vp) for n in ("x", "xb", "xb2", "hb", "hb2", "q", "k", "v", "att", "logits", "key_cache",
                                  "value_cache", "key_matmul", "value_matmul", "key_layer_cache", "value_layer_cache")]


class QTensor(C.Structure):
    _fields_ = [("q", vp), ("s", vp)]


class WeightsQ8(C.Structure):
    _fields_ = ([(n, vp) for n in ("token_embedding_table", "rms_att_weight", "rms_ffn_weight", "rms_final_weight")] +
                [(n, QTensor) for n in ("wq", "wk", "wv", "wo", "w1", "w2", "w3", "wcls")] + [("group_size", C.c_int)])


class WeightsBF16(C.Structure):
    _fields_ = [(n, vp) for n in ("token_embedding_table", "rms_att_weight", "rms_ffn_weight", "rms_final_weight", "wq", "wk",
                                  "wv", "wo", "w1", "w2", "w3", "wcls")]


def declared_symbols(header: str = HEADER) -> list[str]:
    """Every function name include/tha_b200.h declares."""
    text = open(header).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tha_[a-z0-9_]+)\s*\(", text)))


class ThaError(RuntimeError):
    pass


class Lib:
    """Loaded C-ABI library.  `lib.check(lib.c.tha_xxx(...))` raises on a non-zero status."""

    def __init__(self, path: str = LIB_PATH):
        if not os.path.exists(path):
            raise ThaError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
        self.path = path
        self.c = C.CDLL(path)
        c = self.c
        c.tha_last_error.restype = C.c_char_p
        c.tha_set_option.argtypes = [C.c_char_p, C.c_int]
        I, F = C.c_int, C.c_float
        sig = {
            "tha_svds_f32": [vp, I, vp, vp, F],
            "tha_vecaddvec_f32": [vp, vp, vp, I],
            "tha_matmul_batch_f32": [vp, I, vp, vp, vp, I, I, I, I, vp, I, I],
            "tha_matmulvec_f32": [vp, vp, vp, vp, I, I],
            "tha_matmul_reduction_f32": [vp, vp, vp, vp, I, I, I],
            "tha_matmul_f32": [vp, I, I, I, vp, vp, vp],
            "tha_rmsnorm_batch_f32": [vp, I, vp, vp, vp, I, I],
            "tha_rope_f32": [vp, I, I, I, I, vp, vp],
            "tha_softmax_f32": [vp, vp, I],
            "tha_swiglu_f32": [vp, vp, vp, I],
            "tha_mha_scores_v1_f32": [vp, I, vp, I, I, vp, vp, vp, I, I, I, I, I, I],
            "tha_mha_softmax_f32": [vp, I, vp, vp, I, I],
            "tha_mha_weighted_v1_f32": [vp, I, vp, I, vp, vp, vp, I, I, I, I, I, I, I],
            "tha_mha_scores_v2_f32": [vp, I, vp, I, vp, vp, vp, I, I, I, I, I],
            "tha_mha_weighted_v2_f32": [vp, I, vp, I, vp, vp, vp, I, I, I, I, I],
            "tha_mha_fused_v1_f32": [vp, I, vp, I, I, I, vp, vp, vp, vp, I, I, I],
            "tha_mha_fused_v1_f32_static": [vp, I, vp, I, I, I, vp, vp, vp, vp, I, I, I],
            "tha_quantize_q8": [vp, vp, vp, vp, I, I],
            "tha_matmul_batch_q8": [vp, I, vp, I, vp, vp, vp, vp, I, I, I],
            "tha_forward_batch_f32": [vp, I, C.POINTER(Config), C.POINTER(WeightsF32), C.POINTER(State), vp, vp, vp],
            "tha_decode_step_f32": [vp, I, C.POINTER(Config), C.POINTER(WeightsF32), C.POINTER(State), vp, vp, vp, vp],
            "tha_decode_greedy_f32": [vp, I, C.POINTER(Config), C.POINTER(WeightsF32), C.POINTER(State), vp, vp, vp, vp],
            "tha_forward_batch_q8": [vp, I, C.POINTER(Config), C.POINTER(WeightsQ8), C.POINTER(State), vp, vp, vp],
            "tha_decode_step_q8": [vp, I, C.POINTER(Config), C.POINTER(WeightsQ8), C.POINTER(State), vp, vp, vp, vp],
            "tha_gemm_tf32x3": [vp, vp, C.c_longlong, vp, I, I, vp, I, C.c_longlong],
            "tha_gemm_bf16": [vp, vp, C.c_longlong, vp, I, I, vp, I, C.c_longlong],
            "tha_convert_bf16": [vp, vp, vp, C.c_size_t],
            "tha_forward_batch_bf16": [vp, I, C.POINTER(Config), C.POINTER(WeightsBF16), C.POINTER(State), vp, vp, vp],
            "tha_decode_step_bf16": [vp, I, C.POINTER(Config), C.POINTER(WeightsBF16), C.POINTER(State), vp, vp, vp, vp],
            "tha_forward_greedy_bf16": [vp, I, C.POINTER(Config), C.POINTER(WeightsBF16), C.POINTER(State), vp, vp, vp],
            "tha_forward_sample_bf16": [vp, I, C.POINTER(Config), C.POINTER(WeightsBF16), C.POINTER(State), vp, vp, vp, vp, F, F, vp],
            "tha_prefill_bf16": [vp, C.POINTER(Config), C.POINTER(WeightsBF16), C.POINTER(State), I, vp, vp, vp, vp, vp, vp],
            "tha_forward_greedy_f32": [vp, I, C.POINTER(Config), C.POINTER(WeightsF32), C.POINTER(State), vp, vp, vp],
            "tha_forward_greedy_q8": [vp, I, C.POINTER(Config), C.POINTER(WeightsQ8), C.POINTER(State), vp, vp, vp],
            "tha_prefill_f32": [vp, C.POINTER(Config), C.POINTER(WeightsF32), C.POINTER(State), I, vp, vp, vp, vp, vp, vp],
            "tha_prefill_supported": [C.POINTER(Config)],
            "tha_prefill_q8": [vp, C.POINTER(Config), C.POINTER(WeightsQ8), C.POINTER(State), I, vp, vp, vp, vp, vp, vp],
            "tha_matmul_q8_tc": [vp, vp, C.c_longlong, vp, vp, I, I, I, vp, vp, I, I],
            "tha_argmax_batch_f32": [vp, I, vp, I, vp],
            "tha_kv_pool_register": [vp, vp, vp, I, I, I],
            "tha_kv_pool_unregister": [vp],
            "tha_sample": [vp, I, vp, I, F, F, vp, vp, vp, vp, vp],
            "tha_forward_sample_f32": [vp, I, C.POINTER(Config), C.POINTER(WeightsF32), C.POINTER(State), vp, vp, vp, vp, F, F, vp],
            "tha_forward_sample_q8": [vp, I, C.POINTER(Config), C.POINTER(WeightsQ8), C.POINTER(State), vp, vp, vp, vp, F, F, vp],
            "tha_advance_slots": [vp, I, vp, vp, vp, I],
            "tha_tp_create": [C.POINTER(vp), I, I, I, C.POINTER(C.c_int), C.c_size_t],
            "tha_tp_export": [vp, vp],
            "tha_tp_local_ptr": [vp, C.POINTER(vp)],
            "tha_tp_connect_ipc": [vp, C.c_char_p],
            "tha_tp_connect_ptr": [vp, C.POINTER(vp)],
            "tha_tp_check_error": [vp],
            "tha_tp_destroy": [vp],
            "tha_tp_set_mode": [vp, I],
            "tha_tp_select_device": [vp],
            "tha_tp_decode_step_f32": [vp, vp, I, C.POINTER(Config), C.POINTER(WeightsF32), C.POINTER(State), vp, vp, vp, vp, I],
            "tha_tp_decode_step_q8": [vp, vp, I, C.POINTER(Config), C.POINTER(WeightsQ8), C.POINTER(State), vp, vp, vp, vp, I],
            "tha_tp_prefill_f32": [vp, vp, C.POINTER(Config), C.POINTER(WeightsF32), C.POINTER(State), I, vp, vp, vp, vp, vp, vp],
            "tha_tp_prefill_q8": [vp, vp, C.POINTER(Config), C.POINTER(WeightsQ8), C.POINTER(State), I, vp, vp, vp, vp, vp, vp],
            "tha_malloc": [C.POINTER(vp), C.c_size_t],
            "tha_free": [vp],
            "tha_stream_create": [C.POINTER(vp)],
            "tha_stream_sync": [vp],
        }
        for name, argtypes in sig.items():
            fn = getattr(c, name)  # AttributeError = missing symbol = loud failure
            fn.argtypes = argtypes
            fn.restype = C.c_int

    def check(self, status: int, what: str = "") -> None:
        if status != 0:
            raise ThaError(f"{what or 'tha call'} failed with status {status}: {self.c.tha_last_error().decode()}")

    def missing_symbols(self) -> list[str]:
        return [s for s in declared_symbols() if not hasattr(self.c, s)]


_lib = None


def lib() -> Lib:
    global _lib
    if _lib is None:
        _lib = Lib()
    return _lib
