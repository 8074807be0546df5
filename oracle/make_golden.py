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
  analyzer_cases.json   decisions of the REFERENCE's own AlignmentStreamAnalyzer.step on seeded token streams (needs /root/reference)
  t3_full_greedy300.npz / t3_full_sampled300.npz  (python -m oracle.make_golden --full; ~10 min of CPU) the real 30-layer shape,
                        2 requests x 300 steps: tokens, decision margins and, for the greedy stream, the top-8 logits of every step
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
    mint_analyzer()
    if "--batch" in sys.argv:
        mint_batch_fixture()
        mint_awq_fixture()
        return
    if "--full" in sys.argv:
        mint_full_streams()
        mint_full_streams(n_req=1, n_steps=1000, sampled=False)      # BASELINE.json configs[3]: context to ~1100
    print("golden fixtures written to", GOLD)


REF_ANALYZER = "/root/reference/src/chatterbox_vllm/models/t3/inference/alignment_stream_analyzer.py"


def mint_analyzer():
    """Token streams through the reference's own class (importable on CPU): the fixture pins oracle/analyzer_ref.py and the CUDA
    sampler's in-loop heuristics to the reference itself."""
    import json
    import random
    if not os.path.exists(REF_ANALYZER):
        print("reference not present: analyzer_cases.json left as committed")
        return
    from oracle import analyzer_ref
    rng = random.Random(3)
    cases = []
    for n_text in [1, 2, 3, 4, 5, 6, 7, 10, 20, 40, 128]:
        for trial in range(3):
            T = min(2 * n_text + 30, 120)
            toks = [rng.randrange(0, 6561) for _ in range(T)]
            if trial == 1:
                k = rng.randrange(3, T - 3)
                toks[k] = toks[k + 1] = toks[k + 2] = 77
            if trial == 2:
                k = rng.randrange(0, T - 2)
                toks[k] = toks[k + 1] = 5
            cases.append({"n_text": n_text, "tokens": toks, "decisions": analyzer_ref.run_reference_class(REF_ANALYZER, n_text, toks, 6562)})
    with open(os.path.join(GOLD, "analyzer_cases.json"), "w") as f:
        json.dump({"eos": 6562, "source": "AlignmentStreamAnalyzer.step of the reference (alignment_stream_analyzer.py:114-195) run on CPU "
                                          "by oracle/make_golden.py", "cases": cases}, f)


def mint_full_streams(n_req=2, n_steps=300, sampled=True):
    full = T3Dims()
    sd = synthetic_checkpoint(full)
    prompts = [synthetic_request(i, full, n_text=64) for i in range(n_req)]
    orc = T3Oracle(sd, full, mode="bf16")
    g = oracle_generate(orc, prompts, n_steps, rho=2.0)
    ti, tv = top8(g["logits"])
    np.savez_compressed(os.path.join(GOLD, f"t3_full_greedy{n_steps}.npz"), tokens=g["tokens"], margins=g["margins"], top_idx=ti, top_val=tv)
    if not sampled:
        return
    orc = T3Oracle(sd, full, mode="bf16")
    s = oracle_generate(orc, prompts, n_steps, rho=2.0, temperature=0.8, min_p=0.05, top_p=0.8, seeds=[42 + i for i in range(n_req)])
    np.savez_compressed(os.path.join(GOLD, f"t3_full_sampled{n_steps}.npz"), tokens=s["tokens"], race_margins=s["margins"],
                        minp_margins=s["minp_margins"], topp_margins=s["topp_margins"])


SAMPLE_IDX = np.arange(0, 8194, 64, dtype=np.int64)          # a fixed 129-column sample of every logit row (cosine / max-abs)


def mint_batch_fixture(n_req=32, n_steps=8):
    """BASELINE.json configs[1] at its own shape: 30 layers, 32 requests = 64 CFG rows, P = 99, greedy, rho = 2; the first
    ``n_steps`` tokens of every request with the top-8 logits and a fixed column sample of every step."""
    full = T3Dims()
    sd = synthetic_checkpoint(full)
    prompts = [synthetic_request(i, full, n_text=64) for i in range(n_req)]
    g = oracle_generate(T3Oracle(sd, full, mode="bf16"), prompts, n_steps, rho=2.0)
    ti, tv = top8(g["logits"])
    np.savez_compressed(os.path.join(GOLD, f"t3_full_batch{n_req}.npz"), tokens=g["tokens"], margins=g["margins"], top_idx=ti, top_val=tv,
                        sample_idx=SAMPLE_IDX, sample_val=g["logits"][..., SAMPLE_IDX].astype(np.float32))


def mint_awq_fixture(n_req=64, n_steps=4):
    """BASELINE.json configs[2] at full depth: AWQ W4A16 projections (oracle on the dequantised weights), 64 requests = 128 CFG
    rows, temperature 0.8 / top-p 0.8 / min-p 0.05 / rho 2, seeds 42 + i.  Also the same steps (teacher-forced with the same
    tokens) under the reference engine's fp16 dequant definition (awq_ref.dequantize_fp16, vllm awq.py:71-72): its top-8 /
    sampled logits, so the GPU test can report the distance of the bf16-dequant build to the fp16 definition."""
    from oracle import awq_ref
    full = T3Dims()
    sd = synthetic_checkpoint(full)
    sd_deq, packed = awq_ref.quantize_state_dict(sd)
    prompts = [synthetic_request(i, full, n_text=64) for i in range(n_req)]
    seeds = [42 + i for i in range(n_req)]
    s = oracle_generate(T3Oracle(sd_deq, full, mode="bf16"), prompts, n_steps, rho=2.0, temperature=0.8, min_p=0.05, top_p=0.8, seeds=seeds)
    ti, tv = top8(s["logits"])
    sd16 = dict(sd)
    for k, (qw, qz, sc) in packed.items():
        sd16[k] = awq_ref.dequantize_fp16(qw, qz, sc).float().t().contiguous()
    f = oracle_generate(T3Oracle(sd16, full, mode="bf16"), prompts, n_steps, rho=2.0, temperature=0.8, min_p=0.05, top_p=0.8, seeds=seeds,
                        forced=s["tokens"])
    np.savez_compressed(os.path.join(GOLD, f"t3_full_awq{n_req}_sampled.npz"), tokens=s["tokens"], race_margins=s["margins"],
                        minp_margins=s["minp_margins"], topp_margins=s["topp_margins"], top_idx=ti, top_val=tv, sample_idx=SAMPLE_IDX,
                        sample_val=s["logits"][..., SAMPLE_IDX].astype(np.float32),
                        fp16_sample_val=f["logits"][..., SAMPLE_IDX].astype(np.float32),
                        fp16_top_val=np.take_along_axis(f["logits"], ti.astype(np.int64), -1).astype(np.float32))


if __name__ == "__main__":
    main()
