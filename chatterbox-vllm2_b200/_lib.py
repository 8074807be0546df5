"""ctypes binding of ``csrc/libt3b200.so`` (the C ABI declared in ``include/t3b200.h``).

There is no fallback: ``lib()`` raises :class:`T3BLibraryError` when the shared library is missing
or does not export every symbol the header declares, and every wrapper raises it when an entry
point returns a ``T3B_ERR_*`` code (message from ``t3b_last_error``).
"""
import ctypes as C
import os
import re
from typing import List

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "csrc", "libt3b200.so")
HEADER_PATH = os.path.join(os.path.dirname(HERE), "include", "t3b200.h")

T3B_OK = 0
T3B_W_BF16, T3B_W_AWQ = 0, 1
T3B_IMPL_AUTO, T3B_IMPL_GEMV, T3B_IMPL_UMMA = 0, 1, 2
T3B_COMBINE_BF16, T3B_COMBINE_FP32 = 0, 1
T3B_EPI_NONE, T3B_EPI_SILU_MUL = 0, 1
T3B_KV_BF16, T3B_KV_FP8 = 0, 1
T3B_LIN_GATE_UP_INTERLEAVED = 1
ERR_NAMES = {-1: "T3B_ERR_ARG", -2: "T3B_ERR_SHAPE", -3: "T3B_ERR_ALIGN", -4: "T3B_ERR_CUDA", -5: "T3B_ERR_UNSUPPORTED"}


class T3BLibraryError(RuntimeError):
    """libt3b200.so is missing / incomplete, or one of its entry points failed."""

    def __init__(self, msg, code=None):
        super().__init__(msg)
        self.code = code


vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float
p_i32 = C.POINTER(C.c_int32)


class Linear(C.Structure):          # t3b_linear_t
    _fields_ = [("kind", i32), ("n", i32), ("k", i32), ("w", vp), ("scales", vp), ("zeros", vp), ("flags", i32)]


class Layer(C.Structure):           # t3b_layer_t
    _fields_ = [("qkv", Linear), ("o", Linear), ("gate_up", Linear), ("down", Linear), ("ln1", vp), ("ln2", vp)]


class Model(C.Structure):           # t3b_model_t
    _fields_ = [("n_layers", i32), ("hidden", i32), ("n_heads", i32), ("head_dim", i32), ("ffn", i32),
                ("speech_vocab", i32), ("speech_pos_rows", i32), ("max_pos", i32), ("rms_eps", f32),
                ("layers", C.POINTER(Layer)), ("final_norm", vp), ("head", Linear), ("speech_emb", vp),
                ("speech_pos_emb", vp), ("rope_cos_sin", vp)]


class KV(C.Structure):              # t3b_kv_t
    _fields_ = [("k", vp), ("v", vp), ("layer_stride", i64), ("num_blocks", i32), ("dtype", i32)]


class State(C.Structure):           # t3b_state_t
    _fields_ = [("tokens", vp), ("speech_pos", vp), ("rope_pos", vp), ("ctx_lens", vp), ("block_tables", vp),
                ("max_blocks", i32), ("history", vp), ("hist_len", vp), ("hist_stride", i32), ("step", vp),
                ("active", vp), ("stop_tokens", vp), ("max_ctx", i32), ("row_order", vp)]


class Sampling(C.Structure):        # t3b_sampling_t
    _fields_ = [("cfg_scale", vp), ("repetition_penalty", vp), ("temperature", vp), ("min_p", vp), ("top_p", vp),
                ("seed", vp), ("analyzer_text_tokens", vp), ("combine", i32)]


class Options(C.Structure):         # t3b_options_t
    _fields_ = [("gemv_max_rows", i32), ("gemv_no_norm", i32), ("umma_stages", i32), ("umma_row_chunk", i32),
                ("umma_split", i32 * 4), ("attn_team_warps", i32), ("weight_prefetch", i32), ("umma_big_min_rows", i32), ("norm_fusion", i32), ("gemv_mma", i32)]


p_opt = C.POINTER(Options)


def make_options(gemv_max_rows=0, gemv_no_norm=0, umma_stages=0, umma_row_chunk=0, umma_split=(0, 0, 0, 0), attn_team_warps=0,
                 weight_prefetch=0, umma_big_min_rows=0, norm_fusion=0, gemv_mma=0) -> Options:
    return Options(gemv_max_rows, gemv_no_norm, umma_stages, umma_row_chunk, (i32 * 4)(*umma_split), attn_team_warps, weight_prefetch,
                   umma_big_min_rows, norm_fusion, gemv_mma)

_PROTOS = {
    "t3b_last_error": (C.c_char_p, []),
    "t3b_version": (i32, []),
    "t3b_check_device": (i32, []),
    "t3b_workspace_bytes": (C.c_size_t, [C.POINTER(Model), i32]),
    "t3b_kv_slice_bytes": (i32, [i32]),
    "t3b_embed_gather": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp]),
    "t3b_add_rmsnorm": (i32, [vp, vp, i32, vp, vp, vp, i32, i32, f32, i32, vp]),
    "t3b_linear_bf16": (i32, [vp, C.POINTER(Linear), i32, vp, vp, i32, p_i32, i32, i32, p_opt, vp]),
    "t3b_linear_awq": (i32, [vp, C.POINTER(Linear), i32, vp, vp, i32, p_i32, i32, i32, p_opt, vp]),
    "t3b_repack_weights_awq": (i32, [vp, vp, vp, i32, i32, i32, vp, vp, vp, vp]),
    "t3b_gemv_frag_bytes": (C.c_size_t, [i32, i32, i32]),
    "t3b_repack_weights_gemv": (i32, [vp, i32, i32, i32, vp, vp]),
    "t3b_attn_decode_fused": (i32, [vp, i32, vp, vp, vp, i32, vp, i32, vp, vp, i32, i32, vp, p_opt, vp]),
    "t3b_rope_kv_write": (i32, [vp, vp, i32, i32, vp, vp, vp, vp, i32, vp, i32, vp, vp, i32, i32, vp, vp]),
    "t3b_paged_attn_decode": (i32, [vp, i32, vp, vp, vp, i32, vp, vp, i32, i32, vp, p_opt, vp]),
    "t3b_prefill_attn": (i32, [vp, i32, vp, vp, vp, i32, vp, vp, i32, i32, i32, vp, vp]),
    "t3b_prefill_attn_max_len": (i32, [i32]),
    "t3b_prefill_embed": (i32, [vp, vp, vp, vp, vp, i32, vp, vp, i32, i32, vp, vp, i32, i32, i32, vp, vp, vp, vp]),
    "t3b_silu_mul": (i32, [vp, vp, i32, i32, i32, vp, vp]),
    "t3b_forward_tokens": (i32, [C.POINTER(Model), C.POINTER(KV), vp, i32, vp, vp, vp, vp, i32, i32, vp, vp, i32, vp, vp, C.c_size_t,
                                 p_opt, vp]),
    "t3b_decode_step": (i32, [C.POINTER(Model), C.POINTER(KV), C.POINTER(State), vp, i32, vp, vp, C.c_size_t, p_opt, vp]),
    "t3b_head_cfg_sample": (i32, [C.POINTER(Model), vp, i32, C.POINTER(Sampling), C.POINTER(State), i32, vp, vp, vp,
                                  C.c_size_t, p_opt, vp]),
    "t3b_head_cfg_sample_static": (i32, [C.POINTER(Model), vp, i32, C.POINTER(Sampling), C.POINTER(State), i32, vp, vp, p_opt, vp]),
    "t3b_sample_logits": (i32, [vp, i32, i32, C.POINTER(Sampling), C.POINTER(State), vp, vp]),
    "t3b_cfg_sample_head_logits": (i32, [vp, i32, i32, C.POINTER(Sampling), C.POINTER(State), i32, vp, vp, vp]),
    "t3b_filter_speech_tokens": (i32, [vp, i32, i32, i32, vp, vp, vp]),
}

_lib = None


def header_symbols() -> List[str]:
    """Every function name ``include/t3b200.h`` declares (used by the symbol-export test)."""
    src = open(HEADER_PATH).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(t3b_[a-z0-9_]+)\s*\(", src)))


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise T3BLibraryError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  There is no CPU or PyTorch fallback for the T3 decode path.")
    try:
        L = C.CDLL(LIB_PATH)
    except OSError as e:
        raise T3BLibraryError(f"cannot load {LIB_PATH}: {e}") from e
    for name, (res, args) in _PROTOS.items():
        try:
            fn = getattr(L, name)
        except AttributeError as e:
            raise T3BLibraryError(f"{LIB_PATH} does not export {name}; rebuild it") from e
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def check(rc: int, what: str = "") -> None:
    if rc != T3B_OK:
        msg = lib().t3b_last_error().decode("utf-8", "replace")
        raise T3BLibraryError(f"{what or 'libt3b200'}: {ERR_NAMES.get(rc, rc)}: {msg}", rc)


def ptr(t) -> int:
    """Device (or host) address of a torch tensor, or None."""
    return None if t is None else t.data_ptr()


def stream_ptr(stream=None) -> int:
    import torch
    s = torch.cuda.current_stream() if stream is None else stream
    return s.cuda_stream
