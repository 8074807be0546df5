"""CPU restatement of the sampler vLLM runs after T3VllmModel.compute_logits
(TEST INFRASTRUCTURE - see oracle/__init__.py; parity unpinned by the reference, pinned against
the installed vLLM's own pure-torch functions in tests/test_oracle_pins.py).

Order (vllm@installed:v1/sample/sampler.py:67-130 forward, :238-307 sample):
  fp32 cast -> repetition penalty (model_executor/layers/utils.py:51-89; formula
  _custom_ops.py apply_repetition_penalties: l>0 ? l/rho : l*rho on ids seen in prompt U output)
  -> greedy argmax (first max index) if T < 1e-5
  -> l /= T -> min-p (v1/sample/logits_processor/builtin.py:100-114)
  -> top-p (v1/sample/ops/topk_topp_sampler.py:328-369: ascending sort, softmax, cumsum <= 1-p
     masked, last kept) -> softmax -> q ~ Exp(1) -> argmax(p / q)  (:395-416)

History = generated speech tokens only (prompt ids < 2500 fall in the -inf pad of
t3.py:669-672, BOS is injected as an embedding).  Ids are un-offset
(0..8193); the caller adds SPEECH_TOKEN_OFFSET to what it emits.

RNG contract (torch's CUDA Philox stream is not a stable contract, so the
build defines its own, implemented identically here and in csrc/sampler.cu):
  x = Philox4x32-10(counter=(vocab_index, step, 0, 0), key=(seed_lo, seed_hi))[0]
  u = ((x >> 9) + 0.5) * 2^-23        (exact in fp32, 0 < u < 1)
  q = -log(u)                          (fp32)
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

SAMPLING_EPS = 1e-5  # vllm@installed:v1/sample/sampler.py _SAMPLING_EPS

_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10 (Salmon et al. 2011).  All args uint32 arrays/scalars."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint32) for c in (c0, c1, c2, c3))
    k0, k1 = np.uint32(k0), np.uint32(k1)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = _M0 * c0.astype(np.uint64)
            p1 = _M1 * c2.astype(np.uint64)
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), p0.astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), p1.astype(np.uint32)
            c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
            k0 = np.uint32(k0 + _W0)
            k1 = np.uint32(k1 + _W1)
    return c0, c1, c2, c3


def exp_race_q(vocab: int, seed: int, step: int) -> np.ndarray:
    """q[i] ~ Exp(1), i in [0, vocab): the contract in the module docstring."""
    idx = np.arange(vocab, dtype=np.uint32)
    z = np.zeros(vocab, dtype=np.uint32)
    x0, _, _, _ = philox4x32_10(idx, np.full(vocab, step, np.uint32), z, z,
                                seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    u = ((x0 >> np.uint32(9)).astype(np.float32) + np.float32(0.5)) * np.float32(2.0 ** -23)
    return (-np.log(u)).astype(np.float32)


def apply_repetition_penalty(logits: torch.Tensor, history: Sequence[Sequence[int]], rho: Sequence[float]) -> torch.Tensor:
    out = logits.clone()
    for b, h in enumerate(history):
        if len(h) == 0 or rho[b] == 1.0:
            continue
        ids = torch.unique(torch.as_tensor(list(h), dtype=torch.long))
        v = out[b, ids]
        r = torch.tensor(rho[b], dtype=torch.float32)
        out[b, ids] = torch.where(v > 0, v / r, v * r)
    return out


def min_p_mask(logits: torch.Tensor, min_p: float) -> torch.Tensor:
    """builtin.py:100-114 for one row."""
    if min_p <= 0.0:
        return logits
    probs = torch.softmax(logits, dim=-1)
    thr = probs.max() * torch.tensor(min_p, dtype=torch.float32)
    return logits.masked_fill(probs < thr, float("-inf"))


def top_p_mask(logits: torch.Tensor, top_p: float) -> torch.Tensor:
    """topk_topp_sampler.py:328-369 for one row (stable ascending sort: ties keep index order)."""
    if top_p >= 1.0:
        return logits
    srt, idx = torch.sort(logits, dim=-1, descending=False, stable=True)
    cs = torch.cumsum(torch.softmax(srt, dim=-1), dim=-1)
    mask = cs <= (1.0 - torch.tensor(top_p, dtype=torch.float32))
    mask[-1] = False
    srt = srt.masked_fill(mask, float("-inf"))
    return torch.empty_like(logits).scatter_(0, idx, srt)


def sample_ref(logits, history: Sequence[Sequence[int]], *, repetition_penalty, temperature,
               min_p, top_p, seeds, steps, info: Optional[List[Dict]] = None) -> np.ndarray:
    """logits [B,V] fp32 (already CFG-combined).  Per-request scalars as sequences of length B.
    Returns int32 [B] un-offset token ids.  ``info`` (if a list) receives per-request decision
    margins so tests can separate genuine mismatches from fp32-order near-ties."""
    lg = torch.as_tensor(np.asarray(logits, dtype=np.float32)).clone()
    B, V = lg.shape
    lg = apply_repetition_penalty(lg, history, repetition_penalty)
    out = np.zeros(B, dtype=np.int32)
    for b in range(B):
        row = lg[b]
        rec: Dict = {}
        if temperature[b] < SAMPLING_EPS:
            out[b] = int(torch.argmax(row))
            top2 = torch.topk(row, 2).values
            rec["greedy_margin"] = float(top2[0] - top2[1])
        else:
            row = row / torch.tensor(temperature[b], dtype=torch.float32)
            if min_p[b] > 0:
                probs = torch.softmax(row, -1)
                thr = probs.max() * torch.tensor(min_p[b], dtype=torch.float32)
                rec["minp_margin"] = float(((probs - thr).abs() / thr).min())
            row = min_p_mask(row, float(min_p[b]))
            if top_p[b] < 1.0:
                srt, _ = torch.sort(row, stable=True)
                cs = torch.cumsum(torch.softmax(srt, -1), -1)
                rec["topp_margin"] = float((cs - (1.0 - top_p[b])).abs().min())
            row = top_p_mask(row, float(top_p[b]))
            probs = torch.softmax(row, dim=-1)
            q = torch.from_numpy(exp_race_q(V, int(seeds[b]), int(steps[b])))
            score = probs / q
            out[b] = int(torch.argmax(score))
            t2 = torch.topk(score, 2).values
            rec["race_margin"] = float((t2[0] - t2[1]) / t2[0])
            rec["n_kept"] = int(torch.isfinite(row).sum())
        if info is not None:
            info.append(rec)
    return out
