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
    minp_m, topp_m = np.ones((n_steps, B), dtype=np.float32), np.ones((n_steps, B), dtype=np.float32)
    seeds = list(seeds) if seeds is not None else list(range(B))
    for t in range(n_steps):
        lg = orc.compute_logits(hidden)
        logits_all.append(lg.numpy().copy())
        info: List[Dict] = []
        tok = sampler_ref.sample_ref(lg, hist, repetition_penalty=[rho] * B, temperature=[temperature] * B,
                                     min_p=[min_p] * B, top_p=[top_p] * B, seeds=seeds, steps=[t] * B, info=info)
        toks[:, t] = tok
        margins[t] = [r.get("greedy_margin", r.get("race_margin", 0.0)) for r in info]
        minp_m[t] = [r.get("minp_margin", 1.0) for r in info]
        topp_m[t] = [r.get("topp_margin", 1.0) for r in info]
        fed = forced[:, t] if forced is not None else tok
        for b in range(B):
            hist[b].append(int(fed[b]))
        if t + 1 < n_steps:
            hidden = orc.decode_step(rows, [int(x) for x in fed], [t + 1] * B)
    return {"tokens": toks, "logits": np.stack(logits_all), "margins": margins, "minp_margins": minp_m, "topp_margins": topp_m}


def random_state(dims, rows, lens, seed=0, device="cuda:0"):
    """KV pool filled with random bf16 K/V and a decode state whose row r has lens[r] cached tokens (kernel-level tests)."""
    from chatterbox_vllm2_b200.model import DecodeState, KVPool
    g = torch.Generator().manual_seed(seed)
    max_blocks = (max(lens) + 16) // 16 + 1
    nblk = rows * max_blocks + 1
    kv = KVPool(dims, nblk, device)
    for which in ("k", "v"):
        t = kv.as_bf16(which)
        t.copy_((torch.randn(t.shape, generator=g) * 0.5).to(torch.bfloat16))
    st = DecodeState(rows // 2, max_blocks, 64, max(lens) + 8, device)
    perm = torch.randperm(nblk - 1, generator=g)[: rows * max_blocks] + 1
    st.block_tables.copy_(perm.reshape(rows, max_blocks).to(torch.int32))
    st.ctx_lens.copy_(torch.tensor(lens, dtype=torch.int32))
    st.rope_pos.copy_(torch.tensor(lens, dtype=torch.int32))
    st.tokens.copy_(torch.randint(0, dims.speech_vocab, (rows // 2,), generator=g).to(torch.int32))
    st.speech_pos.copy_(torch.randint(0, 300, (rows // 2,), generator=g).to(torch.int32))
    return kv, st


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
