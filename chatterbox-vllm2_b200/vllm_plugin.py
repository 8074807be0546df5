"""Adapters either side of the path.

(i) vLLM plugin shim.  The reference registers its model class with the engine at import time
    (models/t3/__init__.py:5-7: ``ModelRegistry.register_model("ChatterboxT3", T3VllmModel)``; the architecture name comes from
    t3-model*/config.json:2-4) and the engine then drives the class through fixed hook names.  The reference was written
    against vllm==0.10.0 (``get_multimodal_embeddings`` / ``get_input_embeddings`` / ``compute_logits(h, sampling_metadata)``,
    t3.py:335,424,650); current engines (0.22: ``v1/worker/gpu_model_runner.py``) call ``embed_multimodal`` /
    ``embed_input_ids`` / ``compute_logits(h)``.  ``T3B200ForVllm`` answers to both sets of names and forwards to
    ``T3B200Model`` (same arguments, same tensor shapes: the packed [N, 2048] cond | uncond rows of t3.py:260-263).
    ``register()`` is what a maintainer calls in place of the reference's registration line.

    What the shim does NOT do: vLLM owns its KV cache and attention metadata; the B200 kernels use their own pool layout
    ([block][head][16][64]) and device-resident request state.  ``forward`` therefore runs on the pool / state bound
    with ``bind()`` (what ``T3B200Engine`` does); mapping vLLM's block tables into ``t3b_state_t`` is the integration work left
    to the engine side and is out of this path's scope.

(ii) Token hand-off to S3Gen (tts.py:492-514): ``speech_tokens_for_s3gen`` = subtract SPEECH_TOKEN_OFFSET, keep 0 <= t < 6561, in
    order (the GPU form is ``T3B200Model.filter_speech_tokens`` / ``t3b_filter_speech_tokens``).
"""
from typing import Iterable, List, Optional, Sequence, Set, Tuple

import torch
from torch import nn

from .constants import SPEECH_TOKEN_OFFSET, T3Config
from .model import T3B200Model

ARCHITECTURE = "ChatterboxT3"                      # t3-model*/config.json "architectures"
SPEECH_VOCAB_SIZE = T3Config().start_speech_token  # 6561 valid S3 codes (tts.py:506-514)


class T3B200ForVllm(nn.Module):
    """``T3VllmModel`` surface (t3.py:255-716) for the engine's model runner, old (0.10) and new (0.22) hook names."""

    supports_multimodal = True                     # SupportsMultiModal marker attributes the runner checks
    is_text_generation_model = True

    def __init__(self, *, vllm_config=None, prefix: str = "", device: str = "cuda:0", **kw):
        super().__init__()
        self.vllm_config = vllm_config
        self.impl = T3B200Model(vllm_config=vllm_config, prefix=prefix, device=device, **kw)

    # ---- weights (t3.py:300-332)
    def load_weights(self, weights: Iterable[Tuple[str, torch.Tensor]]) -> Set[str]:
        return self.impl.load_weights(weights)

    # ---- multimodal embeddings: conditionals pass straight through (t3.py:335-337)
    def get_multimodal_embeddings(self, **kwargs) -> List[torch.Tensor]:
        return self.impl.get_multimodal_embeddings(**kwargs)

    embed_multimodal = get_multimodal_embeddings                 # vLLM >= 0.11 name

    # ---- input embeddings (t3.py:424-647)
    def get_input_embeddings(self, input_ids: torch.Tensor, multimodal_embeddings=None, **kw) -> torch.Tensor:
        return self.impl.get_input_embeddings(input_ids, multimodal_embeddings, speech_pos=kw.get("speech_pos"))

    def embed_input_ids(self, input_ids: torch.Tensor, multimodal_embeddings=None, *, is_multimodal=None, **kw) -> torch.Tensor:
        return self.impl.get_input_embeddings(input_ids, multimodal_embeddings, speech_pos=kw.get("speech_pos"))

    # ---- backbone (t3.py:676-713)
    def bind(self, kv, state) -> None:
        self.impl.bind(kv, state)

    def forward(self, input_ids, positions: torch.Tensor, intermediate_tensors=None, inputs_embeds: Optional[torch.Tensor] = None,
                **kw) -> torch.Tensor:
        return self.impl.forward(input_ids, positions, intermediate_tensors, inputs_embeds)

    # ---- logits (t3.py:650-673); 0.22 passes hidden states only
    def compute_logits(self, hidden_states: torch.Tensor, sampling_metadata=None) -> torch.Tensor:
        return self.impl.compute_logits(hidden_states, sampling_metadata)

    def get_language_model(self):                                # t3.py:715
        return self.impl


def register(arch: str = ARCHITECTURE) -> None:
    """Replace the reference's ``ModelRegistry.register_model("ChatterboxT3", T3VllmModel)`` (models/t3/__init__.py:5).  Lazy
    form ("module:Class"), so registering does not import torch.cuda or load the CUDA library."""
    from vllm import ModelRegistry
    ModelRegistry.register_model(arch, "chatterbox_vllm2_b200.vllm_plugin:T3B200ForVllm")


def speech_tokens_for_s3gen(token_ids: Sequence[int]) -> List[int]:
    """tts.py:492-514 on host lists: offset ids (what ``generate`` returns) -> S3 codes S3Gen accepts."""
    return [t - SPEECH_TOKEN_OFFSET for t in token_ids if 0 <= t - SPEECH_TOKEN_OFFSET < SPEECH_VOCAB_SIZE]
