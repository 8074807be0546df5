"""CPU restatement of the token heuristics of the reference's ``AlignmentStreamAnalyzer.step``
(src/chatterbox_vllm/models/t3/inference/alignment_stream_analyzer.py:143-195, repetition test :203-213), run IN the decode
loop (TEST INFRASTRUCTURE - see oracle/__init__.py).

Pinned: tests/test_oracle_pins.py drives the reference's own class (importable on CPU, loaded by file path) and this
restatement over the same token streams and requires identical decisions; tests/golden/analyzer_cases.json holds decisions
recorded from the reference class.

Calling convention (what "in the loop" means here, the one the CUDA sampler implements): the analyzer's n-th ``step`` call
(n = 1, 2, ...) shapes the logits of token i = n - 1 and is handed ``next_token`` = token i - 1 (None for i = 0), exactly how the
class is written to be used (``step(logits, next_token=<the token that was just generated>)``).  Its state is then a pure
function of (i, text_tokens_count, the last three tokens):

    frame          = i + 1                                     (curr_frame_pos after the increment, :152)
    text_position  = min(frame // 2, n_text - 1)               (:156-157)
    complete from the first frame with text_position >= n_text - 3  = frame max(1, 2 (n_text - 3))   (:160-163)
    repetition     = tokens[i-1] == tokens[i-2] == tokens[i-3]  (:166, :203-213; the class keeps the last 8 tokens)
    long_tail      = frame - completed_at >= 10                (:169-172)
    force EOS      if long_tail or repetition: every logit -2^15, EOS +2^15   (:188-193)
    suppress EOS   else if text_position < n_text - 3 and n_text > 5: EOS logit = -2^15   (:183-186)
"""
from typing import Optional, Sequence

FORCE, SUPPRESS = "force_eos", "suppress_eos"


def decision(i: int, n_text: int, tokens: Sequence[int]) -> Optional[str]:
    """What the analyzer does to the logits of token i given tokens[0..i-1]; None = untouched."""
    if n_text <= 0:
        return None
    frame = i + 1
    text_position = min(frame // 2, n_text - 1)
    completed_at = max(1, 2 * (n_text - 3))
    repetition = i >= 3 and tokens[i - 1] == tokens[i - 2] == tokens[i - 3]
    long_tail = frame - completed_at >= 10
    if long_tail or repetition:
        return FORCE
    if text_position < n_text - 3 and n_text > 5:
        return SUPPRESS
    return None


def apply(logits, i: int, n_text: int, tokens: Sequence[int], eos: int):
    """logits: 1-D fp32 tensor/array (modified copy returned)."""
    d = decision(i, n_text, tokens)
    out = logits.clone() if hasattr(logits, "clone") else logits.copy()
    if d == FORCE:
        out[:] = -(2.0 ** 15)
        out[eos] = 2.0 ** 15
    elif d == SUPPRESS:
        out[eos] = -(2.0 ** 15)
    return out


def run_reference_class(path: str, n_text: int, tokens: Sequence[int], eos: int, vocab: int = 8194):
    """Decisions of the reference's own class for every i in [0, len(tokens)] (used to pin this file and to mint fixtures)."""
    import importlib.util
    import torch
    spec = importlib.util.spec_from_file_location("ref_alignment_stream_analyzer", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.logger.disabled = True
    an = mod.AlignmentStreamAnalyzer(text_tokens_count=n_text, eos_token_id=eos, device="cpu")
    out = []
    for i in range(len(tokens) + 1):
        lg = an.step(torch.zeros(1, vocab), next_token=None if i == 0 else torch.tensor(tokens[i - 1]))
        e = float(lg[0, eos])
        other = float(lg[0, (eos + 1) % vocab])
        out.append(FORCE if e > 2 ** 14 else (SUPPRESS if (e < -2 ** 14 and other == 0.0) else None))
    return out
