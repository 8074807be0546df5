"""AWQ W4A16 (group 128) weights for the T3 projections: AutoAWQ "GEMM" tensors in, kernel layout out.

The reference selects AWQ with ``quantization="awq"`` (quantization.py:161-169 -> tts.py:161-169) and
expects a pre-quantised checkpoint in AutoAWQ's tensor format (vllm awq.py:214-260).  It ships none, and
there is no network here, so ``quantize_linear`` makes those tensors from bf16 weights (asymmetric 4-bit,
group 128 - AutoAWQ's pseudo-quantiser without the calibration-driven scale search).  Both routes end in
``repack``: one call of the C ABI's ``t3b_repack_weights_awq`` (a one-off layout transform at load time).
"""
from typing import Dict

import torch

from . import _lib

GROUP = 128
_ORDER = (0, 2, 4, 6, 1, 3, 5, 7)      # nibble position i of an int32 holds column 8j + _ORDER[i] (AutoAWQ)


def _pack_cols(v: torch.Tensor) -> torch.Tensor:
    """uint8 [rows, N] (0..15) -> int32 [rows, N/8]."""
    rows, n = v.shape
    v = v.reshape(rows, n // 8, 8).to(torch.int64)
    out = torch.zeros(rows, n // 8, dtype=torch.int64, device=v.device)
    for i, o in enumerate(_ORDER):
        out |= v[:, :, o] << (4 * i)
    return ((out + (1 << 31)) % (1 << 32) - (1 << 31)).to(torch.int32)


def quantize_linear(w: torch.Tensor) -> Dict[str, torch.Tensor]:
    """bf16/fp32 [N, K] -> {"qweight": int32 [K, N/8], "qzeros": int32 [K/128, N/8], "scales": fp16 [K/128, N]}."""
    n, k = w.shape
    if k % GROUP or n % 8:
        raise ValueError(f"AWQ needs K % 128 == 0 and N % 8 == 0 (got N={n}, K={k})")
    g = w.detach().float().reshape(n, k // GROUP, GROUP)
    mx, mn = g.amax(-1), g.amin(-1)
    s = ((mx - mn).clamp(min=1e-5) / 15.0).to(torch.float16)
    s_used = s.to(torch.bfloat16).float()
    z = (-torch.round(mn / s_used)).clamp(0, 15)
    q = (torch.round(g / s_used[..., None]) + z[..., None]).clamp(0, 15)
    return {"qweight": _pack_cols(q.reshape(n, k).t().contiguous().to(torch.uint8)),
            "qzeros": _pack_cols(z.t().contiguous().to(torch.uint8)), "scales": s.t().contiguous()}


def repack(qweight: torch.Tensor, qzeros: torch.Tensor, scales: torch.Tensor, device, gate_up: bool = False) -> Dict[str, torch.Tensor]:
    """AutoAWQ tensors -> {"w": uint8 [N*K/2], "zeros": uint8 [K/128, N], "scales": bf16 [K/128, N]} on ``device``."""
    L = _lib.lib()
    k, n8 = qweight.shape
    n = n8 * 8
    qw = qweight.to(device=device, dtype=torch.int32).contiguous()
    qz = qzeros.to(device=device, dtype=torch.int32).contiguous()
    sc = scales.to(device=device, dtype=torch.float16).contiguous()
    w = torch.empty(n * k // 2, dtype=torch.uint8, device=device)
    zeros = torch.empty(k // GROUP, n, dtype=torch.uint8, device=device)
    sb = torch.empty(k // GROUP, n, dtype=torch.bfloat16, device=device)
    _lib.check(L.t3b_repack_weights_awq(qw.data_ptr(), qz.data_ptr(), sc.data_ptr(), n, k, 1 if gate_up else 0, w.data_ptr(),
                                        zeros.data_ptr(), sb.data_ptr(), _lib.stream_ptr()), "t3b_repack_weights_awq")
    torch.cuda.current_stream().synchronize()      # qw / qz / sc are released on return
    return {"w": w, "zeros": zeros, "scales": sb}


def quantize_and_repack(w: torch.Tensor, gate_up: bool = False) -> Dict[str, torch.Tensor]:
    t = quantize_linear(w)
    return repack(t["qweight"], t["qzeros"], t["scales"], w.device, gate_up)
