"""CPU restatement of the AWQ W4A16 (group 128) tensor format and linear layer the reference switches on
with ``quantization="awq"`` (TEST INFRASTRUCTURE - see oracle/__init__.py).

parity unpinned by the reference: it ships no quantised checkpoint and no test for this path
(quantization.py:161-169 only forwards ``{"quantization": "awq"}`` to vLLM, tts.py:161-169).  The format
is restated from the installed vLLM (a third-party dependency, pinned upstream at vllm==0.10.0,
pyproject.toml:29) and pinned on the GPU box against vLLM's own ``awq_dequantize_triton``
(tests/test_gpu_awq.py):

* tensors          vllm@installed:model_executor/layers/quantization/awq.py:214-260
                   qweight int32 [K, N/8], qzeros int32 [K/128, N/8], scales fp16 [K/128, N]
* nibble order     vllm@installed:model_executor/layers/quantization/awq_triton.py:56-62 - output column
                   8j+i sits at bits 4*[0,4,1,5,2,6,3,7][i] of word j (AutoAWQ packs with order
                   [0,2,4,6,1,3,5,7])
* dequantisation   W[k, n] = (q[k, n] - z[k // 128, n]) * s[k // 128, n]   (awq_triton.py:64-98)
* linear           y = x @ W  (awq.py:268-283: dequantise-then-matmul when tokens >= 256, fused kernel below)

Build semantics: activations are bf16 (the reference's AWQ path is fp16, awq.py:71-72), scales
are converted fp16 -> bf16 once at load and every dequantised weight is rounded to bf16 before the
fp32-accumulated product: W_deq = bf16_rn((q - z) * bf16(s)).
"""
from __future__ import annotations

from typing import Dict, Tuple

import numpy as np
import torch

GROUP = 128
AWQ_ORDER = [0, 2, 4, 6, 1, 3, 5, 7]          # nibble position i of a word holds column 8j + AWQ_ORDER[i]
AWQ_REVERSE = [0, 4, 1, 5, 2, 6, 3, 7]        # column 8j + i sits at nibble position AWQ_REVERSE[i]


def quantize(w: torch.Tensor) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Asymmetric 4-bit, group 128 along K (AutoAWQ pseudo_quantize_tensor, zero_point=True, without the
    activation-aware scale search - there is no calibration data offline).  ``w`` [N, K] float.
    Returns q uint8 [K, N], z uint8 [K/128, N], s fp16 [K/128, N] (numpy)."""
    N, K = w.shape
    assert K % GROUP == 0
    g = w.detach().float().reshape(N, K // GROUP, GROUP)
    mx, mn = g.amax(-1), g.amin(-1)
    s = ((mx - mn).clamp(min=1e-5) / 15.0).to(torch.float16)
    s_used = s.to(torch.bfloat16).float()                    # what the kernels multiply by
    z = (-torch.round(mn / s_used)).clamp(0, 15)
    q = (torch.round(g / s_used[..., None]) + z[..., None]).clamp(0, 15)
    return (q.reshape(N, K).t().contiguous().to(torch.uint8).numpy(), z.t().contiguous().to(torch.uint8).numpy(),
            s.t().contiguous().numpy())


def pack(q: np.ndarray) -> np.ndarray:
    """uint8 [rows, N] values 0..15 -> int32 [rows, N/8] in AutoAWQ nibble order."""
    rows, N = q.shape
    assert N % 8 == 0
    v = q.reshape(rows, N // 8, 8).astype(np.uint32)
    out = np.zeros((rows, N // 8), dtype=np.uint32)
    for i in range(8):
        out |= v[:, :, AWQ_ORDER[i]] << np.uint32(4 * i)
    return out.view(np.int32)


def unpack(p: np.ndarray) -> np.ndarray:
    """int32 [rows, N/8] -> uint8 [rows, N]  (awq_triton.py:56-62 reverse order)."""
    u = p.view(np.uint32)
    rows, n8 = u.shape
    out = np.zeros((rows, n8, 8), dtype=np.uint8)
    for i in range(8):
        out[:, :, i] = ((u >> np.uint32(4 * AWQ_REVERSE[i])) & np.uint32(15)).astype(np.uint8)
    return out.reshape(rows, n8 * 8)


def dequantize_fp16(qweight: np.ndarray, qzeros: np.ndarray, scales: np.ndarray) -> torch.Tensor:
    """The reference engine's arithmetic: fp16 [K, N] = (q - z) * s evaluated in fp16 (awq_triton.py:64-98)."""
    q, z = unpack(qweight).astype(np.float32), unpack(qzeros).astype(np.float32)
    zs = np.repeat(z, GROUP, axis=0)
    ss = np.repeat(scales.astype(np.float16), GROUP, axis=0)
    d = torch.from_numpy(q - zs).to(torch.float16)
    return d * torch.from_numpy(ss)


def dequantize_build(qweight: np.ndarray, qzeros: np.ndarray, scales: np.ndarray) -> torch.Tensor:
    """The build's arithmetic: fp32 tensor [N, K] holding bf16_rn((q - z) * bf16(s)) (nn.Linear layout)."""
    q, z = unpack(qweight).astype(np.float32), unpack(qzeros).astype(np.float32)
    s = torch.from_numpy(scales.astype(np.float32)).to(torch.bfloat16).float()
    d = torch.from_numpy(q - np.repeat(z, GROUP, axis=0))
    w = (d * s.repeat_interleave(GROUP, dim=0)).to(torch.bfloat16).float()
    return w.t().contiguous()


def quantize_state_dict(sd: Dict[str, torch.Tensor]) -> Tuple[Dict[str, torch.Tensor], Dict[str, Tuple[np.ndarray, np.ndarray, np.ndarray]]]:
    """Quantise every projection of the backbone (q/k/v/o/gate/up/down; not the head, norms, embeddings -
    SURVEY.md 8(a) row Q).  Returns (state dict with those weights replaced by their dequantised bf16
    values - what a bf16-mode T3Oracle then computes with - and the packed AutoAWQ tensors per key)."""
    out, packed = dict(sd), {}
    for k, w in sd.items():
        if k.startswith("tfmr.layers.") and k.endswith("_proj.weight"):
            q, z, s = quantize(w)
            qw, qz = pack(q), pack(z)
            packed[k] = (qw, qz, s)
            out[k] = dequantize_build(qw, qz, s).to(torch.bfloat16)
    return out, packed
