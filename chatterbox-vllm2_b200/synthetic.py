"""Seeded synthetic T3 checkpoint and prompts.

There is no network in the build image and the real ``t3_mtl23ls_v2.safetensors`` is pulled
from the HF hub by the reference (tts.py:220-247), so parity and benchmarks run on a
deterministic random-init checkpoint of the exact architecture, using the reference's
safetensors key names (t3.py:300-322) so the same loader serves the real file.
"""
from typing import Dict, List, Tuple

import torch

from .constants import CONDITIONING_SIZE, T3Dims

CKPT_SEED = 20260101


def synthetic_checkpoint(dims: T3Dims = T3Dims(), seed: int = CKPT_SEED,
                         dtype: torch.dtype = torch.bfloat16) -> Dict[str, torch.Tensor]:
    """Linears N(0, 1/fan_in), norm weights 1 + 0.1 N(0,1), embeddings N(0, 0.02^2)
    (learned_pos_emb.py:12), head N(0, 1/hidden).  Returned in ``dtype`` (bf16 like the real file)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    D, F = dims.hidden, dims.ffn
    qkv = dims.n_heads * dims.head_dim

    def lin(n, k):
        return (torch.randn(n, k, generator=g) * (k ** -0.5)).to(dtype)

    sd: Dict[str, torch.Tensor] = {}
    for i in range(dims.n_layers):
        p = f"tfmr.layers.{i}."
        sd[p + "self_attn.q_proj.weight"] = lin(qkv, D)
        sd[p + "self_attn.k_proj.weight"] = lin(qkv, D)
        sd[p + "self_attn.v_proj.weight"] = lin(qkv, D)
        sd[p + "self_attn.o_proj.weight"] = lin(D, qkv)
        sd[p + "mlp.gate_proj.weight"] = lin(F, D)
        sd[p + "mlp.up_proj.weight"] = lin(F, D)
        sd[p + "mlp.down_proj.weight"] = lin(D, F)
        sd[p + "input_layernorm.weight"] = (1 + 0.1 * torch.randn(D, generator=g)).to(dtype)
        sd[p + "post_attention_layernorm.weight"] = (1 + 0.1 * torch.randn(D, generator=g)).to(dtype)
    sd["tfmr.norm.weight"] = (1 + 0.1 * torch.randn(D, generator=g)).to(dtype)
    sd["speech_emb.weight"] = (0.02 * torch.randn(dims.speech_vocab, D, generator=g)).to(dtype)
    sd["speech_pos_emb.emb.weight"] = (0.02 * torch.randn(dims.speech_pos_rows, D, generator=g)).to(dtype)
    sd["text_emb.weight"] = (0.02 * torch.randn(dims.text_vocab, D, generator=g)).to(dtype)
    sd["text_pos_emb.emb.weight"] = (0.02 * torch.randn(dims.text_pos_rows, D, generator=g)).to(dtype)
    sd["speech_head.weight"] = lin(dims.speech_vocab, D)
    return sd


def synthetic_request(i: int, dims: T3Dims = T3Dims(), n_text: int = 64) -> Tuple[torch.Tensor, List[int]]:
    """Request i (seed 1000+i): conditioning prefix 0.3 N(0,1) [34, hidden] bf16 and n_text text ids."""
    g = torch.Generator(device="cpu").manual_seed(1000 + i)
    cond = (0.3 * torch.randn(CONDITIONING_SIZE, dims.hidden, generator=g)).to(torch.bfloat16)
    ids = torch.randint(0, dims.text_vocab, (n_text,), generator=g).tolist()
    return cond, ids
