"""Mint the golden fixtures under tests/golden/ (TEST INFRASTRUCTURE - see oracle/__init__.py).

The reference has no golden vectors for this path, so these are minted from the
oracle AFTER it has been pinned against transformers.LlamaModel and the installed vLLM's torch ops
(tests/test_oracle_pins.py, which imports those libraries and the reference's importable modules
in the build container).  Run from the repo root:  python -m oracle.make_golden

Files (small; committed):
  t3_small_greedy.npz   2-layer model, 3 requests, 16 greedy steps, rho=2: tokens, margins, top-8 logits per step
  t3_small_sampled.npz  same model, T=0.8 min_p=0.05 top_p=0.8 rho=2, seeds 42+i (Philox contract)
  sampler_cases.npz     sampler-only cases on seeded logits (greedy+penalty, min-p/top-p/race)
  rope_table.npz        llama3 cos/sin rows 0,1,99,100,1199 of the full-size table (pinned to vLLM's)
"""
import hashlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from chatterbox_vllm2_b200 import T3Dims                                   # noqa: E402
from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint, synthetic_request   # noqa: E402
from helpers import oracle_generate                                        # noqa: E402
from oracle import sampler_ref                                             # noqa: E402
from oracle.t3_ref import T3Oracle, rope_cos_sin_table                     # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
SMALL = T3Dims().small(n_layers=2, max_pos=512)
N_REQ, N_TEXT, N_STEPS = 3, 12, 16


def top8(logits):
    idx = np.argsort(-logits, axis=-1, kind="stable")[..., :8]
    return idx.astype(np.int32), np.take_along_axis(logits, idx, -1).astype(np.float32)


def main():
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    sd = synthetic_checkpoint(SMALL)
    prompts = [synthetic_request(i, SMALL, n_text=N_TEXT) for i in range(N_REQ)]

    orc = T3Oracle(sd, SMALL, mode="bf16")
    g = oracle_generate(orc, prompts, N_STEPS, rho=2.0)
    ti, tv = top8(g["logits"])
    np.savez_compressed(os.path.join(GOLD, "t3_small_greedy.npz"), tokens=g["tokens"], margins=g["margins"], top_idx=ti, top_val=tv,
                        logits_sha256=np.array(hashlib.sha256(g["logits"].tobytes()).hexdigest()))

    orc = T3Oracle(sd, SMALL, mode="bf16")
    s = oracle_generate(orc, prompts, N_STEPS, rho=2.0, temperature=0.8, min_p=0.05, top_p=0.8, seeds=[42 + i for i in range(N_REQ)])
    np.savez_compressed(os.path.join(GOLD, "t3_small_sampled.npz"), tokens=s["tokens"], margins=s["margins"])

    gen = torch.Generator().manual_seed(77)
    lg = (2.5 * torch.randn(8, 8194, generator=gen)).numpy()
    hist = [torch.randint(0, 8194, (5 * b,), generator=gen).tolist() for b in range(8)]
    cases = {}
    for name, kw in {"greedy": dict(temperature=0.0, min_p=0.0, top_p=1.0),
                     "cfg3": dict(temperature=0.8, min_p=0.05, top_p=0.8),
                     "topp_only": dict(temperature=1.0, min_p=0.0, top_p=0.5)}.items():
        cases[name] = sampler_ref.sample_ref(lg, hist, repetition_penalty=[2.0] * 8, temperature=[kw["temperature"]] * 8,
                                             min_p=[kw["min_p"]] * 8, top_p=[kw["top_p"]] * 8, seeds=[42 + b for b in range(8)],
                                             steps=[3 * b for b in range(8)])
    hl = np.array([len(h) for h in hist], dtype=np.int32)
    hp = np.zeros((8, 40), dtype=np.int32)
    for b, h in enumerate(hist):
        hp[b, :len(h)] = h
    np.savez_compressed(os.path.join(GOLD, "sampler_cases.npz"), logits_seed=np.array(77), hist=hp, hist_len=hl,
                        q_seed42_step0=sampler_ref.exp_race_q(8194, 42, 0)[:16], **{"tok_" + k: v for k, v in cases.items()})

    full = rope_cos_sin_table(T3Dims())
    np.savez_compressed(os.path.join(GOLD, "rope_table.npz"), rows=np.array([0, 1, 99, 100, 1199]),
                        table=full[[0, 1, 99, 100, 1199]].numpy())
    print("golden fixtures written to", GOLD)


if __name__ == "__main__":
    main()
