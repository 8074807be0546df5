"""Pin the CPU oracle against independent implementations (the reference itself has no tests or
golden vectors for this path - SURVEY.md 4 / 8(c)):

* transformers.LlamaModel fp32 built from the reference's config.json shapes  -> backbone
* installed vLLM pure-torch ops (layernorm, llama3 rope, top-p, min-p)          -> op semantics
* the reference's importable modules (t3_config.py, learned_pos_emb.py)        -> constants
* committed golden fixtures (tests/golden/*.npz, minted by oracle/make_golden.py)
"""
import importlib.util
import json
import os

import numpy as np
import pytest
import torch

from chatterbox_vllm2_b200 import T3Config, T3Dims
from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint, synthetic_request
from oracle import sampler_ref
from oracle.t3_ref import T3Oracle, llama3_inv_freq, rope_cos_sin_table

REF = "/root/reference"
GOLD = os.path.join(os.path.dirname(__file__), "golden")
SMALL = T3Dims().small(n_layers=2, max_pos=256)


def _hf_llama(dims, sd):
    from transformers import LlamaConfig, LlamaModel
    cfg = LlamaConfig(hidden_size=dims.hidden, intermediate_size=dims.ffn, num_hidden_layers=dims.n_layers,
                      num_attention_heads=dims.n_heads, num_key_value_heads=dims.n_heads, head_dim=dims.head_dim,
                      rms_norm_eps=dims.rms_eps, rope_theta=dims.rope_theta, vocab_size=8,
                      max_position_embeddings=131072, attention_bias=False, mlp_bias=False,
                      rope_scaling={"factor": 8.0, "high_freq_factor": 4.0, "low_freq_factor": 1.0,
                                    "original_max_position_embeddings": 8192, "rope_type": "llama3"},
                      attn_implementation="eager")
    m = LlamaModel(cfg).to(torch.float32).eval()
    own = m.state_dict()
    for k, v in sd.items():
        if k.startswith("tfmr.") and k[5:] in own:
            own[k[5:]].copy_(v.float())
    return m


def test_backbone_matches_transformers_llama_fp32():
    sd = synthetic_checkpoint(SMALL)
    orc = T3Oracle(sd, SMALL, mode="fp32")
    hf = _hf_llama(SMALL, sd)
    g = torch.Generator().manual_seed(3)
    P = 23
    x = 0.3 * torch.randn(P, SMALL.hidden, generator=g)
    (row,) = orc.new_rows(1)
    h_prefill = orc.forward_tokens(x, [row] * P, range(P))
    with torch.no_grad():
        out = hf(inputs_embeds=x[None], use_cache=True)
    assert torch.allclose(h_prefill, out.last_hidden_state[0], atol=2e-4, rtol=1e-4)
    # three decode steps through the KV cache
    past = out.past_key_values
    for t in range(3):
        xt = 0.3 * torch.randn(1, SMALL.hidden, generator=g)
        h = orc.forward_tokens(xt, [row], [P + t])
        with torch.no_grad():
            o = hf(inputs_embeds=xt[None], past_key_values=past, use_cache=True)
        past = o.past_key_values
        assert torch.allclose(h, o.last_hidden_state[0], atol=2e-4, rtol=1e-4)


def test_prefill_equals_token_by_token():
    sd = synthetic_checkpoint(SMALL)
    a, b = T3Oracle(sd, SMALL, mode="bf16"), T3Oracle(sd, SMALL, mode="bf16")
    x = (0.3 * torch.randn(9, SMALL.hidden, generator=torch.Generator().manual_seed(5))).bfloat16().float()
    (ra,), (rb,) = a.new_rows(1), b.new_rows(1)
    ha = a.forward_tokens(x, [ra] * 9, range(9))
    hb = torch.cat([b.forward_tokens(x[t:t + 1], [rb], [t]) for t in range(9)])
    assert torch.allclose(ha, hb, atol=5e-2)           # bf16 ulps from BLAS summation order only
    assert (ha - hb).abs().mean() < 5e-3               # < 1 bf16 ulp (2^-8 at |x|~1) on average; BLAS-build dependent


def test_rmsnorm_matches_vllm_ir_op():
    pytest.importorskip("vllm")
    spec = importlib.util.spec_from_file_location(
        "_vllm_ln", os.path.join(os.path.dirname(importlib.util.find_spec("vllm").origin), "ir/ops/layernorm.py"))
    src = open(spec.origin).read()
    # the file only needs `register_op`; evaluate its two functions without importing vllm.ir
    ns = {"torch": torch, "Tensor": torch.Tensor}
    body = src.split("@register_op\ndef rms_norm")[1]
    exec("def rms_norm" + body.split("@rms_norm.register_input_generator")[0], ns)
    fused = src.split("@register_op(allow_inplace=True)\ndef fused_add_rms_norm")[1]
    exec("def fused_add_rms_norm" + fused.split("# fused_add_rms_norm has similar")[0], ns)
    orc = T3Oracle(synthetic_checkpoint(SMALL), SMALL, mode="bf16")
    g = torch.Generator().manual_seed(0)
    x = torch.randn(5, 1024, generator=g).bfloat16()
    r = torch.randn(5, 1024, generator=g).bfloat16()
    w = (1 + 0.1 * torch.randn(1024, generator=g)).bfloat16()
    y_ref, res_ref = ns["fused_add_rms_norm"](x, r, w, 1e-5)
    y, res = orc.add_rmsnorm(x.float(), r.float(), w.float())
    assert torch.equal(y.bfloat16(), y_ref) and torch.equal(res.bfloat16(), res_ref)
    y0_ref = ns["rms_norm"](x, w, 1e-5)
    y0, res0 = orc.add_rmsnorm(x.float(), None, w.float())
    assert torch.equal(y0.bfloat16(), y0_ref) and torch.equal(res0.bfloat16(), x)


def test_rope_matches_vllm_llama3():
    pytest.importorskip("vllm")
    from vllm.model_executor.layers.rotary_embedding.common import ApplyRotaryEmb
    from vllm.model_executor.layers.rotary_embedding.llama3_rope import Llama3RotaryEmbedding
    d = T3Dims()
    tmpl = Llama3RotaryEmbedding.__new__(Llama3RotaryEmbedding)
    tmpl.scaling_factor, tmpl.low_freq_factor, tmpl.high_freq_factor = d.rope_factor, d.rope_low, d.rope_high
    tmpl.orig_max_position, tmpl.rotary_dim, tmpl.base = d.rope_orig_max, d.head_dim, d.rope_theta
    tmpl.max_position_embeddings = d.max_pos
    inv = Llama3RotaryEmbedding._compute_inv_freq(tmpl, d.rope_theta)
    assert torch.equal(inv, llama3_inv_freq(d))
    cache = Llama3RotaryEmbedding._compute_cos_sin_cache(tmpl)
    assert torch.equal(cache, rope_cos_sin_table(d))
    orc = T3Oracle(synthetic_checkpoint(SMALL), SMALL, mode="bf16")
    g = torch.Generator().manual_seed(1)
    x = torch.randn(7, 16, 64, generator=g).bfloat16()
    pos = torch.tensor([0, 1, 5, 99, 100, 200, 255])
    cs = rope_cos_sin_table(SMALL).bfloat16()[pos]
    ref = ApplyRotaryEmb.forward_static(x, cs[:, :32], cs[:, 32:], True)
    assert torch.equal(orc.rope(x.float(), pos).bfloat16(), ref)


def test_sampler_matches_vllm_torch_ops():
    pytest.importorskip("vllm")
    from vllm.v1.sample.ops.topk_topp_sampler import apply_top_k_top_p_pytorch
    g = torch.Generator().manual_seed(2)
    lg = 3 * torch.randn(4, 8194, generator=g)
    p = torch.tensor([0.8, 0.5, 0.95, 0.2])
    ref = apply_top_k_top_p_pytorch(lg.clone(), None, p)
    mine = torch.stack([sampler_ref.top_p_mask(lg[i], float(p[i])) for i in range(4)])
    assert torch.equal(torch.isinf(ref), torch.isinf(mine)) and torch.equal(ref, mine)
    # min-p: builtin.py:100-114 restated on a [B,V] tensor
    mp = torch.tensor([[0.05], [0.2], [0.0], [0.5]])
    probs = torch.softmax(lg, -1)
    ref2 = lg.clone().masked_fill_(probs < probs.amax(-1, keepdim=True) * mp, float("-inf"))
    mine2 = torch.stack([sampler_ref.min_p_mask(lg[i], float(mp[i])) for i in range(4)])
    assert torch.equal(ref2, mine2)
    # repetition penalty: utils.py:51-89 mask semantics + l>0 ? l/rho : l*rho
    hist = [[5, 5, 9], [], [8193], [0, 1, 2, 3]]
    pen = sampler_ref.apply_repetition_penalty(lg, hist, [2.0, 2.0, 1.5, 1.0])
    assert pen[0, 5] == (lg[0, 5] / 2 if lg[0, 5] > 0 else lg[0, 5] * 2)
    assert torch.equal(pen[1], lg[1]) and torch.equal(pen[3], lg[3])
    assert pen[2, 8193] == (lg[2, 8193] / 1.5 if lg[2, 8193] > 0 else lg[2, 8193] * 1.5)


def test_philox_known_answer():
    # Random123 kat_vectors: philox4x32-10, ctr=0, key=0 and ctr=key=0xffffffff, and the pi vector
    out = sampler_ref.philox4x32_10(0, 0, 0, 0, 0, 0)
    assert [int(x) for x in out] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    f = 0xFFFFFFFF
    out = sampler_ref.philox4x32_10(f, f, f, f, f, f)
    assert [int(x) for x in out] == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    out = sampler_ref.philox4x32_10(0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344, 0xa4093822, 0x299f31d0)
    assert [int(x) for x in out] == [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]
    q = sampler_ref.exp_race_q(8194, 42, 7)
    assert q.dtype == np.float32 and (q > 0).all() and np.isfinite(q).all()
    assert 0.9 < q.mean() < 1.1


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree only exists in the build container")
def test_constants_match_reference_modules():
    def load(path, name):
        spec = importlib.util.spec_from_file_location(name, os.path.join(REF, path))
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        return m
    rc = load("src/chatterbox_vllm/models/t3/modules/t3_config.py", "_ref_t3cfg").T3Config
    for ctor in ("english_only", "multilingual"):
        a, b = getattr(rc, ctor)(), getattr(T3Config, ctor)()
        assert vars(a) == vars(b) and a.n_channels == b.n_channels and a.is_multilingual == b.is_multilingual
    cfg = json.load(open(os.path.join(REF, "t3-model-multilingual/config.json")))
    d = T3Dims()
    assert (cfg["num_hidden_layers"], cfg["num_attention_heads"], cfg["head_dim"], cfg["intermediate_size"]) == \
        (d.n_layers, d.n_heads, d.head_dim, d.ffn)
    assert cfg["hidden_size"] == 2 * d.hidden and cfg["rms_norm_eps"] == d.rms_eps
    assert cfg["rope_theta"] == d.rope_theta and cfg["rope_scaling"]["factor"] == d.rope_factor
    # learned positional table gather (learned_pos_emb.py:21-32) == plain row gather
    lpe = load("src/chatterbox_vllm/models/t3/modules/learned_pos_emb.py", "_ref_lpe").LearnedPositionEmbeddings
    m = lpe(d.speech_pos_rows, 64)
    idx = torch.tensor([0, 1, 7, 4099])
    assert torch.equal(m.get_fixed_embedding(idx)[0], m.emb.weight[idx])
    src = open(os.path.join(REF, "src/chatterbox_vllm/models/t3/t3.py")).read()
    assert "SPEECH_TOKEN_OFFSET = 2500" in src and "CONDITIONING_SIZE = 34" in src


def test_decode_embedding_and_cfg_semantics():
    sd = synthetic_checkpoint(SMALL)
    orc = T3Oracle(sd, SMALL, mode="bf16", cfg_scale=0.5)
    e = orc.embed_decode([10, 6561], [3, 0])
    ref = (sd["speech_emb.weight"][[10, 6561]] + sd["speech_pos_emb.emb.weight"][[3, 0]])  # torch bf16 add
    assert torch.equal(e.bfloat16(), ref)
    lc = torch.randn(3, 8194).bfloat16()
    lu = torch.randn(3, 8194).bfloat16()
    ref = lc + 0.5 * (lc - lu)                                   # t3.py:662 evaluated by torch in bf16
    assert torch.equal(orc.cfg_combine(lc.float(), lu.float()).bfloat16(), ref)
    cond, ids = synthetic_request(0, SMALL, n_text=5)
    c, u = orc.prefill_embeds(cond.float(), ids)
    assert c.shape == (34 + 5 + 1, 1024) and torch.equal(c[:34], u[:34]) and torch.equal(c[-1], u[-1])
    assert u[34:39].abs().sum() == 0


def test_gemv_fragment_order_forms_the_matmul():
    """oracle/gemv_frag_ref.py: the fragment-ordered weight copy of the tensor-core GEMV, with the k permutation of its mma.sync
    operands, pairs every W[n][k] with x[r][k] - plain projections (incl. a last tile that is only partly valid, like the 8194-wide
    head) and gate_up tiles (8 gate + the 8 matching up features).  The GPU test compares t3b_repack_weights_gemv with this layout."""
    from oracle import gemv_frag_ref as gf
    rng = np.random.default_rng(0)
    for N, K, gu in ((48, 64, False), (50, 128, False), (64, 64, True), (130, 1024, False), (96, 4096, False)):
        w, x = rng.standard_normal((N, K)), rng.standard_normal((5, K))
        f = gf.frag_order(w, gu)
        assert f.shape == ((N // 16 if gu else (N + 15) // 16), K // 32, 2, 32, 8)
        assert np.abs(gf.mma_emulate(f, x, N, gu) - x @ w.T).max() < 1e-10
