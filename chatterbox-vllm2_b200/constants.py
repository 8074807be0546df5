"""Constants of the T3 decode path, mirroring the reference's names.

Sources (relative to the reference repo): models/t3/modules/t3_config.py:1-38 (T3Config),
models/t3/t3.py:37-49 (prefill markers, CONDITIONING_SIZE, SPEECH_TOKEN_OFFSET),
t3-model-multilingual/config.json (backbone shapes; ``hidden_size`` there is the 2048 packing
hack of t3.py:260-263 - the real width is 1024).
"""
from dataclasses import dataclass, replace

PREFILL_COND_START_TOKEN = 695   # t3.py:37
PREFILL_COND_END_TOKEN = 696     # t3.py:38
PREFILL_END_TOKEN = 697          # t3.py:39
CONDITIONING_SIZE = 34           # t3.py:42
SPEECH_TOKEN_OFFSET = 2500       # t3.py:49
KV_BLOCK = 16                    # paged-KV block size (vLLM default; BASELINE.json configs[3])


class T3Config:
    """Same fields as the reference's T3Config (t3_config.py:1-38)."""

    def __init__(self, text_tokens_dict_size=704):
        self.start_text_token = 255
        self.stop_text_token = 0
        self.text_tokens_dict_size = text_tokens_dict_size
        self.max_text_tokens = 2048
        self.start_speech_token = 6561
        self.stop_speech_token = 6562
        self.speech_tokens_dict_size = 8194
        self.max_speech_tokens = 4096
        self.llama_config_name = "Llama_520M"
        self.input_pos_emb = "learned"
        self.speech_cond_prompt_len = 150
        self.encoder_type = "voice_encoder"
        self.speaker_embed_size = 256
        self.use_perceiver_resampler = True
        self.emotion_adv = True

    @property
    def n_channels(self):
        return 1024

    @property
    def is_multilingual(self):
        return self.text_tokens_dict_size > 704

    @classmethod
    def english_only(cls):
        return cls(text_tokens_dict_size=704)

    @classmethod
    def multilingual(cls):
        return cls(text_tokens_dict_size=2454)


@dataclass(frozen=True)
class T3Dims:
    """Backbone shapes.  Kernels require head_dim == 64, hidden % 128 == 0, ffn % 128 == 0."""
    n_layers: int = 30
    hidden: int = 1024
    n_heads: int = 16
    head_dim: int = 64
    ffn: int = 4096
    speech_vocab: int = 8194
    speech_pos_rows: int = 4100      # max_speech_tokens + 2 + 2 (t3.py:283)
    text_vocab: int = 2454
    text_pos_rows: int = 2050        # max_text_tokens + 2 (t3.py:280)
    rms_eps: float = 1e-5
    rope_theta: float = 500000.0
    rope_factor: float = 8.0
    rope_low: float = 1.0
    rope_high: float = 4.0
    rope_orig_max: int = 8192
    max_pos: int = 2048              # rows of the RoPE cos/sin table (>= max_model_len)

    def small(self, **kw):
        return replace(self, **kw)
