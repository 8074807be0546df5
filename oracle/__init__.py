"""CPU oracle for the Chatterbox T3 decode step.

THIS PACKAGE IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it.  The product package
(``chatterbox-vllm2_b200``) never imports anything from here and fails loudly
when its CUDA library is missing.

PARITY STATUS: **parity unpinned by the reference.**  groxaxo/chatterbox-vllm2
ships no tests, golden vectors or fixtures for this path,
its T3 class cannot be imported in this image (it needs vllm==0.10.0, the image
has 0.22.0) and the arithmetic lives in that third-party wheel.  What pins the
oracle instead (see tests/test_oracle_pins.py and oracle/make_golden.py):

* ``transformers.LlamaModel`` (5.5.0, fp32) built from the reference's own
  ``t3-model-multilingual/config.json`` with hidden_size=1024 - backbone math,
  llama3 RoPE, RMSNorm, SiLU-gated MLP, causal attention with KV cache;
* the installed vLLM 0.22.0 pure-torch ops the reference's engine runs on
  (``ir/ops/layernorm.py`` fused_add_rms_norm, ``rotary_embedding`` llama3 +
  ``ApplyRotaryEmb.forward_static``, ``apply_top_k_top_p_pytorch``,
  ``MinPLogitsProcessor.apply`` arithmetic, repetition-penalty formula);
* the reference's importable modules (``t3_config.py``, ``learned_pos_emb.py``)
  for constants and the positional table gather;
* golden vectors minted from those imports and committed under tests/golden/.
"""
