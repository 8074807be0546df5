"""Prompt packing and the prefill branches of the reference's T3 plugin, host side.

What the reference does around the prefill (all in ``models/t3/t3.py``):

* ``T3MultiModalProcessor.apply`` (t3.py:143-249) rewrites a tokenised text prompt into
  ``[695, ids[0] x 32, 696, *ids, 697]`` and stretches the ``[34, 1024]`` conditioning tensor to one row per
  prompt token: the 34 conditioning rows, then a lower-triangular ones matrix whose row sum - 1 is the position of
  each text token (t3.py:94-102,212-221), then a zero row for the start-of-speech token.
* ``split_prefill_decode`` (t3.py:340-421) cuts the engine's flat ``input_ids`` into prefill blocks (ids < 2500,
  a new block at every 695) and decode runs (ids >= 2500), handing each prefill block its slice of the stretched rows.
* ``get_input_embeddings`` (t3.py:542-639) turns a prefill block into cond / uncond embeddings, with four cases for
  blocks the engine's chunked prefill cut in two.

This module restates that control flow; the embedding arithmetic itself (table gathers + adds) runs in the CUDA
library through ``T3B200Model.embed``.  ``T3B200Engine`` does not go through these functions for its own prompts
(it assembles whole prompts in one kernel, ``t3b_prefill_embed``); they exist so a caller that speaks the
reference's plugin protocol - the vLLM model runner - gets the same rows.
"""
from typing import List, Optional, Sequence, Tuple

import torch

from .constants import (CONDITIONING_SIZE, PREFILL_COND_END_TOKEN, PREFILL_COND_START_TOKEN, PREFILL_END_TOKEN,
                        SPEECH_TOKEN_OFFSET)


def create_triangular_matrix(m: int, n: int) -> torch.Tensor:
    """t3.py:94-102: row i has ones in columns 0..i."""
    return (torch.arange(n).unsqueeze(0) <= torch.arange(m).unsqueeze(1)).float()


def pack_prompt(prompt_ids: Sequence[int], conditionals: torch.Tensor) -> Tuple[List[int], torch.Tensor]:
    """t3.py:189-221: (final_prompt_ids, stretched conditionals [len(final_prompt_ids), D])."""
    prompt_ids = list(prompt_ids)
    if len(prompt_ids) == 0:
        raise ValueError("pack_prompt: empty text prompt (the reference replicates prompt_ids[0], t3.py:193)")
    if conditionals.shape[0] != CONDITIONING_SIZE:
        raise ValueError(f"Conditionals must be {CONDITIONING_SIZE} tokens long")          # t3.py:210
    final = [PREFILL_COND_START_TOKEN, *([prompt_ids[0]] * (CONDITIONING_SIZE - 2)), PREFILL_COND_END_TOKEN, *prompt_ids,
             PREFILL_END_TOKEN]
    d = conditionals.shape[1]
    rows = torch.cat([conditionals, create_triangular_matrix(len(prompt_ids), d).to(conditionals.dtype),
                      torch.zeros(1, d, dtype=conditionals.dtype)], dim=0)
    assert rows.shape[0] == len(final)
    return final, rows


def unpack_prompt(final_prompt_ids: Sequence[int]) -> List[int]:
    """The text ids inside a complete packed prompt."""
    ids = list(final_prompt_ids)
    if len(ids) < CONDITIONING_SIZE + 1 or ids[0] != PREFILL_COND_START_TOKEN or ids[-1] != PREFILL_END_TOKEN:
        raise ValueError("not a complete packed prompt")
    return ids[CONDITIONING_SIZE:-1]


def split_prefill_decode(input_ids: Sequence[int], multimodal_embeddings: Sequence[torch.Tensor]):
    """t3.py:340-421.  Returns [(ids list, rows | None)]: rows = that prefill block's slice of the stretched
    conditionals, None for a decode run."""
    ids = [int(i) for i in input_ids]
    if not ids:
        return []
    remaining = torch.cat(list(multimodal_embeddings), dim=0) if len(multimodal_embeddings) else torch.zeros(0, 0)
    out, buf = [], []
    in_prefill = ids[0] < SPEECH_TOKEN_OFFSET

    def flush():
        nonlocal remaining, buf
        if not buf:
            return
        if in_prefill:
            out.append((buf, remaining[:len(buf)]))
            remaining = remaining[len(buf):]
        else:
            out.append((buf, None))
        buf = []

    for t in ids:
        if (in_prefill != (t < SPEECH_TOKEN_OFFSET)) or t == PREFILL_COND_START_TOKEN:
            flush()
            in_prefill = t < SPEECH_TOKEN_OFFSET
        buf.append(t)
    flush()
    return out


def prefill_block_plan(ids: Sequence[int], rows: Optional[torch.Tensor]):
    """The four cases of t3.py:562-639 as a plan: (n_cond_rows taken from `rows`, text ids, text positions, has_bos).
    The block's embedding is [rows[:n_cond] | text_emb[ids] + text_pos_emb[pos] (zeros in the uncond half) | BOS if has_bos]."""
    ids = list(ids)
    n = len(ids)
    if ids[0] == PREFILL_COND_START_TOKEN and ids[-1] == PREFILL_END_TOKEN:            # t3.py:562-581 the whole block
        text = ids[CONDITIONING_SIZE:-1]
        return CONDITIONING_SIZE, text, list(range(len(text))), True
    if ids[0] == PREFILL_COND_START_TOKEN:                                              # t3.py:582-601 the start of a block
        text = ids[CONDITIONING_SIZE:]
        return min(CONDITIONING_SIZE, n), text, list(range(len(text))), False
    if ids[-1] == PREFILL_END_TOKEN:
        if PREFILL_COND_END_TOKEN in ids:                                               # t3.py:611-627 tail of the prefix + all text
            k = ids.index(PREFILL_COND_END_TOKEN)
            text = ids[k + 1:-1]
            return k + 1, text, list(range(len(text))), True
        text = ids[:-1]                                                                 # t3.py:628-646 text tail only
        if rows is None or rows.shape[0] < len(text):
            raise ValueError("prefill tail block without its positional rows")
        pos = (rows[:len(text)].float().sum(dim=1) - 1).round().to(torch.int64).tolist()   # t3.py:621 row sum - 1
        return 0, text, pos, True
    raise ValueError(f"Unknown prefill block: {ids}")                                   # t3.py:648-652
