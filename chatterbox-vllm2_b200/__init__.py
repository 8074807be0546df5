"""chatterbox-vllm2_b200: B200-native (sm_100a) T3 CFG decode step for Chatterbox-on-vLLM.

Drop-in for one path of groxaxo/chatterbox-vllm2: the T3 speech-token decode step
(models/t3/t3.py get_input_embeddings / forward / compute_logits + the sampler vLLM runs after
it).  Host code is Python; all arithmetic runs in hand-written CUDA behind the C ABI declared in
``include/t3b200.h`` (``csrc/libt3b200.so``).  There is no CPU fallback: anything that computes
raises ``T3BLibraryError`` if the library is missing.
"""
from .constants import (CONDITIONING_SIZE, KV_BLOCK, PREFILL_COND_END_TOKEN, PREFILL_COND_START_TOKEN,
                        PREFILL_END_TOKEN, SPEECH_TOKEN_OFFSET, T3Config, T3Dims)

__all__ = ["CONDITIONING_SIZE", "KV_BLOCK", "PREFILL_COND_END_TOKEN", "PREFILL_COND_START_TOKEN",
           "PREFILL_END_TOKEN", "SPEECH_TOKEN_OFFSET", "T3Config", "T3Dims",
           "T3B200Model", "T3B200Engine", "SamplingParams"]


def __getattr__(name):
    if name == "T3B200Model":
        from .model import T3B200Model
        return T3B200Model
    if name in ("T3B200Engine", "SamplingParams"):
        from . import engine
        return getattr(engine, name)
    raise AttributeError(name)
