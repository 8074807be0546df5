"""Shared test drivers: the oracle's request-level decode loop (CPU) and small utilities."""
import json
import os
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from oracle import sampler_ref
from oracle.t3_ref import START_SPEECH_TOKEN, T3Oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def oracle_generate(orc: T3Oracle, prompts: Sequence, n_steps: int, *, rho: float = 2.0, temperature: float = 0.0,
                    min_p: float = 0.0, top_p: float = 1.0, seeds: Optional[Sequence[int]] = None,
                    forced: Optional[np.ndarray] = None) -> Dict[str, np.ndarray]:
    """Run B requests in lock-step through the oracle: prefill (t3.py:542-561), then ``n_steps`` sampled
    tokens each (the first comes from the BOS hidden state).  ``prompts`` = [(cond_emb [34,D], text_ids)].
    ``forced`` [B, n_steps] (optional) teacher-forces the token fed back at every step.
    Returns tokens [B, n_steps] (what the sampler chose), logits [n_steps, B, V] (combined, pre-penalty)
    and margins [n_steps, B] (greedy top-2 gap after penalty)."""
    B = len(prompts)
    rows, hs = [], []
    for cond, ids in prompts:
        c, u = orc.prefill_embeds(cond.float(), ids)
        rc, ru, h = orc.prefill_request(c, u)
        rows += [rc, ru]
        hs.append(h)
    hidden = torch.cat(hs, 0)
    hist: List[List[int]] = [[] for _ in range(B)]
    toks = np.zeros((B, n_steps), dtype=np.int32)
    logits_all, margins = [], np.zeros((n_steps, B), dtype=np.float32)
    seeds = list(seeds) if seeds is not None else list(range(B))
    for t in range(n_steps):
        lg = orc.compute_logits(hidden)
        logits_all.append(lg.numpy().copy())
        info: List[Dict] = []
        tok = sampler_ref.sample_ref(lg, hist, repetition_penalty=[rho] * B, temperature=[temperature] * B,
                                     min_p=[min_p] * B, top_p=[top_p] * B, seeds=seeds, steps=[t] * B, info=info)
        toks[:, t] = tok
        margins[t] = [r.get("greedy_margin", r.get("race_margin", 0.0)) for r in info]
        fed = forced[:, t] if forced is not None else tok
        for b in range(B):
            hist[b].append(int(fed[b]))
        if t + 1 < n_steps:
            hidden = orc.decode_step(rows, [int(x) for x in fed], [t + 1] * B)
    return {"tokens": toks, "logits": np.stack(logits_all), "margins": margins}


def first_divergence(a: Sequence[int], b: Sequence[int]) -> int:
    """Index of the first differing element, or min(len) if none."""
    n = min(len(a), len(b))
    for i in range(n):
        if int(a[i]) != int(b[i]):
            return i
    return n


def record(name: str, payload: dict) -> None:
    """Append a parity record to gpurun_out/parity.jsonl (travels back from the GPU box)."""
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "parity.jsonl"), "a") as f:
        f.write(json.dumps({"test": name, **payload}) + "\n")
