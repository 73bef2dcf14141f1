"""ctypes binding of include/b200llm.h (libb200llm.so, built by csrc/Makefile).

There is no CPU fallback: if the shared library is missing the import of this module
raises, and every entry point returns B200LLM_ENODEV without an sm_100 GPU.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200LLM_LIB: another build of the same sources (the trace build of scripts/tcd_trace.py); there is no non-CUDA substitute
LIB_PATH = os.environ.get("B200LLM_LIB") or os.path.join(_HERE, "libb200llm.so")

ABI_VERSION = 2
OK, EINVAL, ENODEV, ECUDA, ESTATE, ENCCL, ENOMEM = 0, -1, -2, -3, -4, -5, -6

W_EMBED, W_LM_HEAD, W_FINAL_NORM, W_INV_FREQ, W_ATTN_NORM, W_QKV, W_O, W_MLP_NORM, W_GATE_UP, W_DOWN = range(10)
W_QKV_ST, W_O_ST, W_GATE_UP_ST, W_DOWN_ST, W_LM_HEAD_ST = range(10, 15)
W_QKV_TM, W_O_TM, W_GATE_UP_TM, W_DOWN_TM, W_LM_HEAD_TM = range(15, 20)
EPI_STORE, EPI_RESIDUAL, EPI_SWIGLU, EPI_QKV_ROPE, EPI_LOGITS = range(5)


class Config(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("d", C.c_int32), ("n_layers", C.c_int32), ("n_q", C.c_int32),
                ("n_kv", C.c_int32), ("head_dim", C.c_int32), ("ffn", C.c_int32), ("vocab", C.c_int32),
                ("rms_eps", C.c_float), ("page_tokens", C.c_int32), ("max_lanes", C.c_int32),
                ("max_pages_per_lane", C.c_int32), ("max_rows", C.c_int32), ("tp_rank", C.c_int32),
                ("tp_size", C.c_int32), ("device", C.c_int32), ("use_graphs", C.c_int32), ("use_pdl", C.c_int32),
                ("attn_splits", C.c_int32), ("use_megakernel", C.c_int32), ("use_tc", C.c_int32), ("stream", C.c_void_p), ("nccl_unique_id", C.c_void_p)]


class TensorDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("layer", C.c_int32), ("dev", C.c_void_p), ("rows", C.c_int64),
                ("cols", C.c_int64)]


class Sampling(C.Structure):
    _fields_ = [("temperature", C.c_float), ("top_k", C.c_int32), ("top_p", C.c_float),
                ("repeat_penalty", C.c_float), ("seed", C.c_uint64)]


class Timing(C.Structure):
    _fields_ = [("prefill_ms", C.c_float), ("decode_ms", C.c_float), ("prefill_tokens", C.c_int32),
                ("decode_steps", C.c_int32), ("kernel_launches", C.c_int64)]


class GemvArgs(C.Structure):
    _fields_ = [("w", C.c_void_p), ("x", C.c_void_p), ("x_stride", C.c_int64), ("norm_gain", C.c_void_p),
                ("rms_eps", C.c_float), ("out", C.c_void_p), ("out_stride", C.c_int64), ("N", C.c_int32),
                ("K", C.c_int32), ("bs", C.c_int32), ("epi", C.c_int32), ("n_q", C.c_int32), ("n_kv", C.c_int32),
                ("head_dim", C.c_int32), ("page_tokens", C.c_int32), ("max_pages_per_lane", C.c_int32),
                ("inv_freq", C.c_void_p), ("row_lane", C.c_void_p), ("row_pos", C.c_void_p),
                ("page_table", C.c_void_p), ("k_pages", C.c_void_p), ("v_pages", C.c_void_p),
                ("part_max", C.c_void_p), ("part_idx", C.c_void_p), ("logits_idx_offset", C.c_int32),
                ("grid_hint", C.c_int32), ("kv_page_heads", C.c_int32), ("reserved0", C.c_int32)]


class AttnArgs(C.Structure):
    _fields_ = [("q", C.c_void_p), ("out", C.c_void_p), ("k_pages", C.c_void_p), ("v_pages", C.c_void_p),
                ("row_lane", C.c_void_p), ("row_pos", C.c_void_p), ("page_table", C.c_void_p), ("bs", C.c_int32),
                ("n_q", C.c_int32), ("n_kv", C.c_int32), ("head_dim", C.c_int32), ("page_tokens", C.c_int32),
                ("max_pages_per_lane", C.c_int32), ("n_splits", C.c_int32), ("kv_page_heads", C.c_int32), ("scratch", C.c_void_p),
                ("scratch_bytes", C.c_size_t)]


class AttnPrefillArgs(C.Structure):
    _fields_ = [("q", C.c_void_p), ("out", C.c_void_p), ("k_pages", C.c_void_p), ("v_pages", C.c_void_p),
                ("page_table_row", C.c_void_p), ("n_ptab", C.c_int32), ("rows", C.c_int32), ("pos0", C.c_int32),
                ("n_q", C.c_int32), ("n_kv", C.c_int32), ("head_dim", C.c_int32), ("n_pages", C.c_int32),
                ("kv_page_heads", C.c_int32)]


class SampleArgs(C.Structure):
    _fields_ = [("logits", C.c_void_p), ("bs", C.c_int32), ("vocab", C.c_int32), ("part_max", C.c_void_p),
                ("part_idx", C.c_void_p), ("n_part", C.c_int32), ("params", C.c_void_p), ("history", C.c_void_p),
                ("step", C.c_uint64), ("out_token", C.c_void_p)]


# every symbol include/b200llm.h declares: name -> (restype, argtypes)
_P = C.POINTER
SYMBOLS = {
    "b200llm_last_error": (C.c_char_p, []),
    "b200llm_abi_version": (C.c_int, []),
    "b200llm_create": (C.c_int, [_P(Config), _P(C.c_void_p)]),
    "b200llm_destroy": (C.c_int, [C.c_void_p]),
    "b200llm_workspace_bytes": (C.c_int, [_P(Config), _P(C.c_size_t)]),
    "b200llm_kv_page_bytes": (C.c_int, [_P(Config), _P(C.c_size_t)]),
    "b200llm_bind_weights": (C.c_int, [C.c_void_p, _P(TensorDesc), C.c_int32]),
    "b200llm_bind_kv_pool": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "b200llm_bind_kv_pool_shared": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t]),
    "b200llm_bind_workspace": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "b200llm_set_page_table": (C.c_int, [C.c_void_p, C.c_int32, _P(C.c_int32), C.c_int32]),
    "b200llm_set_sampling": (C.c_int, [C.c_void_p, C.c_int32, _P(Sampling)]),
    "b200llm_set_lane": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32]),
    "b200llm_set_history": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.c_int32, C.c_int32]),
    "b200llm_prefill": (C.c_int, [C.c_void_p, C.c_int32, _P(C.c_int32), C.c_int32, C.c_int32]),
    "b200llm_decode": (C.c_int, [C.c_void_p, _P(C.c_int32), C.c_int32, C.c_int32]),
    "b200llm_read_tokens": (C.c_int, [C.c_void_p, _P(C.c_int32), C.c_int32, C.c_int32]),
    "b200llm_read_lane": (C.c_int, [C.c_void_p, C.c_int32, _P(C.c_int32), _P(C.c_int32)]),
    "b200llm_step_logits": (C.c_int, [C.c_void_p, C.c_int32, _P(C.c_float), C.c_int32]),
    "b200llm_sync": (C.c_int, [C.c_void_p]),
    "b200llm_last_timing": (C.c_int, [C.c_void_p, _P(Timing)]),
    "b200llm_stream": (C.c_void_p, [C.c_void_p]),
    "b200llm_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "b200llm_tp_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200llm_tp_import": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32]),
    "b200llm_debug_read_trace": (C.c_int, [C.c_void_p, _P(C.c_uint64), C.c_int32]),
    "b200llm_debug_tp_allreduce_us": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, _P(C.c_float)]),
    "b200llm_op_gemv": (C.c_int, [_P(GemvArgs), C.c_void_p]),
    "b200llm_op_gemv_grid": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P(C.c_int32)]),
    "b200llm_op_gemv_tm": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "b200llm_op_attn_decode": (C.c_int, [_P(AttnArgs), C.c_void_p]),
    "b200llm_op_attn_prefill": (C.c_int, [_P(AttnPrefillArgs), C.c_void_p]),
    "b200llm_op_attn_scratch_bytes": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                                _P(C.c_size_t)]),
    "b200llm_op_attn_decode_static": (C.c_int, [_P(AttnArgs), C.c_void_p]),
    "b200llm_op_gemm_tc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                     C.c_int32, C.c_void_p]),
    "b200llm_op_sample": (C.c_int, [_P(SampleArgs), C.c_void_p]),
    "b200llm_op_fill_weights": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_float, C.c_float, C.c_void_p]),
    "b200llm_op_dequant_bf16": (C.c_int, [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p]),
}


class B200LLMError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"b200llm error {code}: {message}")
        self.code = code
        self.message = message


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make -C llm-based-apache-spark-optimization_b200/csrc` "
            "(or __graft_entry__.build()). b200llm has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.b200llm_abi_version() != ABI_VERSION:
        raise ImportError(f"libb200llm.so ABI {lib.b200llm_abi_version()} != binding ABI {ABI_VERSION}")
    return lib


lib = _load()


def check(rc: int) -> None:
    if rc != OK:
        raise B200LLMError(rc, (lib.b200llm_last_error() or b"").decode("utf-8", "replace"))
