"""CPU definition of the optional FP8 KV-cache format T3B_KV_FP8 (TEST INFRASTRUCTURE - see oracle/__init__.py).

Not in the reference ("KV-traffic reducers"; the reference's vLLM would use ``kv_cache_dtype="fp8"``
with one per-tensor scale).  Definition used by csrc/attention.cu and csrc/elementwise.cu:

    for every (token, head) vector x of 64 values (K after RoPE, V as projected):
        scale = max|x| / 448        (448 = largest finite e4m3; scale = 1 for an all-zero vector)
        q     = e4m3_rn(x / scale)  (round to nearest even, saturating - never reached because |x / scale| <= 448)
        value read back = float(q) * scale

parity unpinned (there is nothing in the reference to pin it to); torch's float8_e4m3fn conversion is the rounding authority.
"""
import torch


def quantize(x: torch.Tensor):
    """x fp32 [..., 64] -> (uint8 e4m3 bytes [..., 64], fp32 scales [...])."""
    amax = x.abs().amax(dim=-1)
    scale = torch.where(amax > 0, amax / 448.0, torch.ones_like(amax))
    q = (x / scale.unsqueeze(-1)).to(torch.float8_e4m3fn)
    return q.view(torch.uint8), scale


def dequantize(q: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    return q.view(torch.float8_e4m3fn).to(torch.float32) * scale.unsqueeze(-1)


def roundtrip(x: torch.Tensor) -> torch.Tensor:
    q, s = quantize(x)
    return dequantize(q, s)
