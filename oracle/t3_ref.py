"""CPU restatement of the Chatterbox T3 CFG decode path (TEST INFRASTRUCTURE - see oracle/__init__.py).

parity unpinned by the reference: it has no tests/golden vectors for this path; this
restatement is pinned against transformers.LlamaModel and the installed vLLM's pure-torch
ops instead (tests/test_oracle_pins.py).

What it follows (paths relative to /root/reference unless prefixed ``vllm@installed:``):

* decode embedding         src/chatterbox_vllm/models/t3/t3.py:440-486 (intended semantics,
                           explicit positional index - SURVEY.md 0.3-2), table gather
                           models/t3/modules/learned_pos_emb.py:21-32
* prefill cond/uncond/BOS  t3.py:542-561
* row pairing, positions   t3.py:676-713  (rows 2i = cond, 2i+1 = uncond of request i here)
* head x2, CFG, offset     t3.py:650-673
* backbone                 vllm@installed:model_executor/models/llama.py:117-121,223-233,316-333,395-432
* fused add + RMSNorm      vllm@installed:ir/ops/layernorm.py:40-59
* llama3 RoPE              vllm@installed:model_executor/layers/rotary_embedding/llama3_rope.py:33-55,
                           base.py:69-92 (cache fp32 -> model dtype), common.py:144-174 (neox rotate)
* SiLU*mul                 vllm@installed:model_executor/layers/activation.py:138-141

Two numeric modes:
  ``fp32``  no rounding anywhere (weights are the bf16-representable values, held in fp32);
  ``bf16``  rounds to bf16 at every point where the reference's bf16 pipeline materialises a
            tensor (GEMM outputs, norm output and its pre-weight cast, residual, each RoPE
            product/sum, attention output, SiLU output and product, head output, each CFG
            combine op).  Reductions (GEMM accumulate, variance, softmax, P.V) stay fp32.
            The CUDA kernels round at exactly these points, so the only remaining difference
            is fp32 summation order.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, replace
from typing import Dict, List, Optional, Sequence, Tuple

import torch

SPEECH_TOKEN_OFFSET = 2500   # t3.py:49
CONDITIONING_SIZE = 34       # t3.py:42
START_SPEECH_TOKEN = 6561    # t3_config.py:8
STOP_SPEECH_TOKEN = 6562     # t3_config.py:9


@dataclass(frozen=True)
class T3Dims:
    """Shapes of the T3 backbone (t3-model-multilingual/config.json)."""
    n_layers: int = 30
    hidden: int = 1024
    n_heads: int = 16
    head_dim: int = 64
    ffn: int = 4096
    speech_vocab: int = 8194
    speech_pos_rows: int = 4100
    text_vocab: int = 2454
    text_pos_rows: int = 2050
    rms_eps: float = 1e-5
    rope_theta: float = 500000.0
    rope_factor: float = 8.0
    rope_low: float = 1.0
    rope_high: float = 4.0
    rope_orig_max: int = 8192
    max_pos: int = 2048

    def small(self, **kw) -> "T3Dims":
        return replace(self, **kw)


def _rnd_bf16(x: torch.Tensor) -> torch.Tensor:
    return x.to(torch.bfloat16).to(torch.float32)


def llama3_inv_freq(d: T3Dims) -> torch.Tensor:
    """vllm@installed:.../rotary_embedding/base.py:69-79 + llama3_rope.py:33-55."""
    inv = 1.0 / (d.rope_theta ** (torch.arange(0, d.head_dim, 2, dtype=torch.float) / d.head_dim))
    low_wl = d.rope_orig_max / d.rope_low
    high_wl = d.rope_orig_max / d.rope_high
    wl = 2 * math.pi / inv
    smooth = (d.rope_orig_max / wl - d.rope_low) / (d.rope_high - d.rope_low)
    return torch.where(
        wl < high_wl, inv,
        torch.where(wl > low_wl, inv / d.rope_factor,
                    (1 - smooth) * inv / d.rope_factor + smooth * inv))


def rope_cos_sin_table(d: T3Dims) -> torch.Tensor:
    """[max_pos, head_dim] fp32, columns = cos(32) | sin(32)  (base.py:81-92)."""
    t = torch.arange(d.max_pos, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, llama3_inv_freq(d))
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1)


class T3Oracle:
    """Two-row-per-request CFG decode of the T3 Llama backbone with per-row KV."""

    def __init__(self, state_dict: Dict[str, torch.Tensor], dims: T3Dims = T3Dims(),
                 mode: str = "bf16", cfg_scale: float = 0.5, combine: str = "bf16",
                 max_len: int = 1280):
        assert mode in ("fp32", "bf16") and combine in ("fp32", "bf16")
        self.d, self.mode, self.cfg_scale, self.combine = dims, mode, float(cfg_scale), combine
        self.max_len = max_len
        f = lambda k: state_dict[k].detach().to(torch.float32).contiguous()
        L = dims.n_layers
        p = "tfmr.layers.%d."
        self.wqkv = [torch.cat([f(p % i + "self_attn.q_proj.weight"), f(p % i + "self_attn.k_proj.weight"),
                                f(p % i + "self_attn.v_proj.weight")], 0) for i in range(L)]
        self.wo = [f(p % i + "self_attn.o_proj.weight") for i in range(L)]
        self.wgu = [torch.cat([f(p % i + "mlp.gate_proj.weight"), f(p % i + "mlp.up_proj.weight")], 0)
                    for i in range(L)]
        self.wd = [f(p % i + "mlp.down_proj.weight") for i in range(L)]
        self.ln1 = [f(p % i + "input_layernorm.weight") for i in range(L)]
        self.ln2 = [f(p % i + "post_attention_layernorm.weight") for i in range(L)]
        self.lnf = f("tfmr.norm.weight")
        self.speech_emb = f("speech_emb.weight")
        self.speech_pos = f("speech_pos_emb.emb.weight")
        self.head = f("speech_head.weight")
        self.text_emb = f("text_emb.weight") if "text_emb.weight" in state_dict else None
        self.text_pos = f("text_pos_emb.emb.weight") if "text_pos_emb.emb.weight" in state_dict else None
        cs = rope_cos_sin_table(dims)
        self.cos_sin = self.rnd(cs)              # base.py:59-63: cache cast to the model dtype
        self.kc: List[List[torch.Tensor]] = []   # [row][layer] -> [max_len, H, d]
        self.vc: List[List[torch.Tensor]] = []
        self.len: List[int] = []

    # ---------------------------------------------------------------- helpers
    def rnd(self, x: torch.Tensor) -> torch.Tensor:
        return _rnd_bf16(x) if self.mode == "bf16" else x

    def new_rows(self, n: int) -> List[int]:
        d = self.d
        first = len(self.len)
        for _ in range(n):
            self.kc.append([torch.zeros(self.max_len, d.n_heads, d.head_dim) for _ in range(d.n_layers)])
            self.vc.append([torch.zeros(self.max_len, d.n_heads, d.head_dim) for _ in range(d.n_layers)])
            self.len.append(0)
        return list(range(first, first + n))

    def add_rmsnorm(self, x: torch.Tensor, res: Optional[torch.Tensor], w: torch.Tensor):
        """vllm@installed:ir/ops/layernorm.py:40-59 (res=None: rms_norm, :9-21)."""
        s = x if res is None else x + res
        res_out = self.rnd(s)
        var = s.pow(2).mean(dim=-1, keepdim=True)
        y = s * torch.rsqrt(var + self.d.rms_eps)
        y = self.rnd(self.rnd(y) * w)            # x.to(weight.dtype) * weight
        return y, res_out

    def rope(self, x: torch.Tensor, pos: torch.Tensor) -> torch.Tensor:
        """x [n, H, d]; neox halves; common.py:144-174 evaluated in the model dtype."""
        half = self.d.head_dim // 2
        cs = self.cos_sin[pos]                   # [n, d]
        cos, sin = cs[:, None, :half], cs[:, None, half:]
        x1, x2 = x[..., :half], x[..., half:]
        o1 = self.rnd(self.rnd(x1 * cos) - self.rnd(x2 * sin))
        o2 = self.rnd(self.rnd(x2 * cos) + self.rnd(x1 * sin))
        return torch.cat((o1, o2), dim=-1)

    # ---------------------------------------------------------------- backbone
    @torch.no_grad()
    def forward_tokens(self, x: torch.Tensor, rows: Sequence[int], pos: Sequence[int]) -> torch.Tensor:
        """Run n tokens through the 30 layers + final norm.

        ``rows[j]`` is the KV row token j belongs to, ``pos[j]`` its RoPE position.  Tokens of the
        same row must be contiguous and in position order; each token attends to everything
        already in its row's cache plus the tokens of this call up to and including itself.
        Returns the final-norm output [n, hidden].
        """
        d = self.d
        n = x.shape[0]
        H, hd = d.n_heads, d.head_dim
        post = torch.as_tensor(list(pos), dtype=torch.long)
        segs: List[Tuple[int, int, int]] = []    # (row, a, b)
        a = 0
        for j in range(1, n + 1):
            if j == n or rows[j] != rows[a]:
                segs.append((rows[a], a, j))
                a = j
        res = None
        h = x
        for li in range(d.n_layers):
            y, res = self.add_rmsnorm(h, res, self.ln1[li])
            qkv = self.rnd(y @ self.wqkv[li].t())
            q = self.rope(qkv[:, :d.hidden].reshape(n, H, hd), post)
            k = self.rope(qkv[:, d.hidden:2 * d.hidden].reshape(n, H, hd), post)
            v = qkv[:, 2 * d.hidden:].reshape(n, H, hd)
            att = torch.empty(n, H, hd)
            for (row, a, b) in segs:
                L0 = self.len[row]
                t = b - a
                self.kc[row][li][L0:L0 + t] = k[a:b]
                self.vc[row][li][L0:L0 + t] = v[a:b]
                K = self.kc[row][li][:L0 + t]
                V = self.vc[row][li][:L0 + t]
                s = torch.einsum("thd,lhd->htl", q[a:b], K) * (hd ** -0.5)
                if t > 1:
                    key_idx = torch.arange(L0 + t)[None, None, :]
                    q_idx = (L0 + torch.arange(t))[None, :, None]
                    s = s.masked_fill(key_idx > q_idx, float("-inf"))
                pr = torch.softmax(s, dim=-1)
                att[a:b] = torch.einsum("htl,lhd->thd", pr, V)
            att = self.rnd(att.reshape(n, d.hidden))
            o = self.rnd(att @ self.wo[li].t())
            y, res = self.add_rmsnorm(o, res, self.ln2[li])
            gu = self.rnd(y @ self.wgu[li].t())
            g, u = gu[:, :d.ffn], gu[:, d.ffn:]
            act = self.rnd(self.rnd(torch.nn.functional.silu(g)) * u)
            h = self.rnd(act @ self.wd[li].t())
        for (row, a, b) in segs:
            self.len[row] += b - a
        y, _ = self.add_rmsnorm(h, res, self.lnf)
        return y

    # ---------------------------------------------------------------- embeddings
    def embed_decode(self, tokens: Sequence[int], speech_pos: Sequence[int]) -> torch.Tensor:
        """t3.py:440,475-478: speech_emb[id - 2500] + speech_pos_emb[p]  (ids here are un-offset)."""
        tok = torch.as_tensor(list(tokens), dtype=torch.long)
        p = torch.as_tensor(list(speech_pos), dtype=torch.long) % self.speech_pos.shape[0]  # t3.py:472
        return self.rnd(self.speech_emb[tok] + self.speech_pos[p])

    def prefill_embeds(self, cond_emb: torch.Tensor, text_ids: Sequence[int]):
        """t3.py:542-561 -> (cond [P,D], uncond [P,D]), P = 34 + len(text) + 1."""
        ids = torch.as_tensor(list(text_ids), dtype=torch.long)
        text = self.rnd(self.text_emb[ids] + self.text_pos[: len(ids)])
        bos = self.embed_decode([START_SPEECH_TOKEN], [0])
        ce = self.rnd(cond_emb.to(torch.float32))
        cond = torch.cat([ce, text, bos], 0)
        uncond = torch.cat([ce, torch.zeros_like(text), bos], 0)
        return cond, uncond

    # ---------------------------------------------------------------- head + CFG
    @torch.no_grad()
    def head_logits(self, hidden: torch.Tensor) -> torch.Tensor:
        return self.rnd(hidden @ self.head.t())

    def cfg_combine(self, lc: torch.Tensor, lu: torch.Tensor, w: Optional[float] = None) -> torch.Tensor:
        """t3.py:662: cond + w * (cond - uncond).  'bf16' = torch's bf16 tensor arithmetic."""
        w = self.cfg_scale if w is None else float(w)
        if self.mode == "bf16" and self.combine == "bf16":
            return _rnd_bf16(lc + _rnd_bf16(w * _rnd_bf16(lc - lu)))
        return lc + w * (lc - lu)

    @torch.no_grad()
    def compute_logits(self, hidden: torch.Tensor) -> torch.Tensor:
        """hidden [R,D] (rows interleaved cond/uncond) -> combined logits [B, 8194] fp32 (no -inf pad:
        t3.py:666-672's 2500 -inf columns only shift the emitted id)."""
        l = self.head_logits(hidden)
        return self.cfg_combine(l[0::2], l[1::2])

    # ---------------------------------------------------------------- request-level driver
    @torch.no_grad()
    def prefill_request(self, cond_embeds: torch.Tensor, uncond_embeds: torch.Tensor) -> Tuple[int, int, torch.Tensor]:
        """Allocate two rows, run the whole prompt.  Returns (cond_row, uncond_row, hidden[2,D] of the
        last prompt token = BOS) so the caller can sample the first speech token."""
        rc, ru = self.new_rows(2)
        P = cond_embeds.shape[0]
        hc = self.forward_tokens(cond_embeds, [rc] * P, range(P))
        hu = self.forward_tokens(uncond_embeds, [ru] * P, range(P))
        return rc, ru, torch.stack([hc[-1], hu[-1]], 0)

    @torch.no_grad()
    def decode_step(self, rows: Sequence[int], tokens: Sequence[int], speech_pos: Sequence[int]) -> torch.Tensor:
        """One decode step for B requests: rows = [c0,u0,c1,u1,...], tokens/speech_pos per request.
        Returns hidden [2B, D]."""
        e = self.embed_decode(tokens, speech_pos)
        x = e.repeat_interleave(2, dim=0)                 # t3.py:480 same embedding for both rows
        pos = [self.len[r] for r in rows]
        return self.forward_tokens(x, list(rows), pos)
