"""T3B200Model: the reference's T3 model surface (models/t3/t3.py:255-716 ``T3VllmModel``) over
libt3b200.so.

Same method names and argument meaning as the reference class - ``load_weights``,
``get_multimodal_embeddings``, ``get_input_embeddings``, ``forward``, ``compute_logits``,
``get_language_model`` - plus ``sample`` (the reference leaves sampling to the vLLM engine;
BASELINE.json's north star puts it on the operator surface) and the native-row entry points the
engine drives (``forward_tokens``, ``decode_step``, ``head_cfg_sample``).

Layout note: the reference packs cond‖uncond side by side into one 2048-wide row per request
(t3.py:260-263,480,697,712).  A bf16 [N, 2048] tensor is bit-for-bit a [2N, 1024] tensor whose
rows 2i / 2i+1 are request i's cond / uncond rows - the native row order of the kernels - so the
packed adapters are views, not copies.

All arithmetic happens in the CUDA library; torch is used for device memory and streams only.
"""
import ctypes as C
import os
from typing import Dict, Iterable, List, Optional, Sequence, Set, Tuple

import torch

from . import _lib
from ._lib import T3BLibraryError
from .constants import CONDITIONING_SIZE, KV_BLOCK, SPEECH_TOKEN_OFFSET, T3Config, T3Dims

START_SPEECH_TOKEN = 6561   # t3_config.py:8
STOP_SPEECH_TOKEN = 6562    # t3_config.py:9


def llama3_rope_table(d: T3Dims) -> torch.Tensor:
    """cos|sin cache [max_pos, 64] fp32 (vllm rotary_embedding/base.py:69-92, llama3_rope.py:33-55).
    One-off host precompute at load time; the kernels read the bf16 cast (base.py:59-63)."""
    import math
    inv = 1.0 / (d.rope_theta ** (torch.arange(0, d.head_dim, 2, dtype=torch.float) / d.head_dim))
    low_wl, high_wl = d.rope_orig_max / d.rope_low, d.rope_orig_max / d.rope_high
    wl = 2 * math.pi / inv
    smooth = (d.rope_orig_max / wl - d.rope_low) / (d.rope_high - d.rope_low)
    inv = torch.where(wl < high_wl, inv, torch.where(wl > low_wl, inv / d.rope_factor,
                                                     (1 - smooth) * inv / d.rope_factor + smooth * inv))
    freqs = torch.einsum("i,j -> ij", torch.arange(d.max_pos, dtype=torch.float), inv)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1)


class KVPool:
    """Paged KV pool (include/t3b200.h t3b_kv_t): K and V as bytes [n_layers, num_blocks, n_heads, slice] where a
    (block, head) slice holds 16 tokens - 2048 B of bf16, or for ``dtype="fp8"`` 1024 B of e4m3 + 16 fp32 scales.
    Block 0 is the scratch block idle rows write to; the allocator hands out 1..num_blocks-1.  Blocks are reference
    counted so that rows can share the blocks of a common prefix (t3.py:554-556: the cond and uncond rows of a request
    start with the same 34 conditioning embeddings)."""

    def __init__(self, dims: T3Dims, num_blocks: int, device, dtype: str = "bf16"):
        if dtype not in ("bf16", "fp8"):
            raise ValueError(f"KV dtype {dtype!r} not supported (bf16 or fp8)")
        self.dtype = dtype
        self.kv_dtype = _lib.T3B_KV_FP8 if dtype == "fp8" else _lib.T3B_KV_BF16
        self.slice_bytes = 1088 if dtype == "fp8" else 2048
        shape = (dims.n_layers, num_blocks, dims.n_heads, self.slice_bytes)
        self.k = torch.zeros(shape, dtype=torch.uint8, device=device)
        self.v = torch.zeros(shape, dtype=torch.uint8, device=device)
        self.num_blocks = num_blocks
        self.c = _lib.KV(self.k.data_ptr(), self.v.data_ptr(), num_blocks * dims.n_heads * self.slice_bytes, num_blocks, self.kv_dtype)
        self.free: List[int] = list(range(num_blocks - 1, 0, -1))
        self.refs: Dict[int, int] = {}

    @staticmethod
    def bytes_per_block(dims: T3Dims, dtype: str = "bf16") -> int:
        return 2 * dims.n_layers * dims.n_heads * (1088 if dtype == "fp8" else 2048)

    def alloc(self, n: int) -> List[int]:
        if n > len(self.free):
            raise MemoryError(f"KV pool exhausted: need {n} blocks, {len(self.free)} free")
        out = self.free[-n:] if n else []
        if n:
            del self.free[-n:]
        for b in out:
            self.refs[b] = 1
        return out

    def share(self, blocks: Sequence[int]) -> None:
        """One more row names these blocks in its table."""
        for b in blocks:
            self.refs[b] += 1

    def release(self, blocks: Sequence[int]) -> None:
        for b in blocks:
            r = self.refs.get(b, 1) - 1
            if r <= 0:
                self.refs.pop(b, None)
                self.free.append(b)
            else:
                self.refs[b] = r

    def as_bf16(self, which: str = "k") -> torch.Tensor:
        """bf16 view [n_layers, num_blocks, n_heads, 16, 64] of a bf16 pool (tests fill / compare it)."""
        assert self.dtype == "bf16"
        t = self.k if which == "k" else self.v
        return t.view(torch.bfloat16).view(t.shape[0], t.shape[1], t.shape[2], KV_BLOCK, 64)


class DecodeState:
    """Device-resident per-request decode state (t3b_state_t) for up to ``max_req`` requests."""

    def __init__(self, max_req: int, max_blocks: int, hist_stride: int, max_ctx: int, device,
                 stop_token: int = STOP_SPEECH_TOKEN):
        z = lambda *s: torch.zeros(*s, dtype=torch.int32, device=device)
        self.max_req, self.max_blocks, self.hist_stride, self.max_ctx = max_req, max_blocks, hist_stride, max_ctx
        # the per-step counters live in ONE buffer (views below), so a snapshot / restore of a decode position is one copy
        self.small = z(9 * max_req)
        v = lambda a, n: self.small[a * max_req:(a + n) * max_req]
        self.tokens, self.speech_pos, self.hist_len, self.step, self.active = v(0, 1), v(1, 1), v(2, 1), v(3, 1), v(4, 1)
        self.rope_pos, self.ctx_lens = v(5, 2), v(7, 2)
        self.block_tables = z(2 * max_req, max_blocks)
        self.history = z(max_req, hist_stride)
        self.stop_tokens = torch.full((max_req,), stop_token, dtype=torch.int32, device=device)   # per request; < 0 = never stops
        self.row_order = torch.arange(2 * max_req, dtype=torch.int32, device=device)              # attention deal order (a hint)
        self.c = self.view(0, ordered=True)

    def view(self, slot0: int, ordered: bool = False) -> _lib.State:
        """t3b_state_t whose request 0 is slot ``slot0`` (``ordered``: with the row_order hint, slot0 = 0 only)."""
        p = lambda t, n: t.data_ptr() + 4 * n
        return _lib.State(p(self.tokens, slot0), p(self.speech_pos, slot0), p(self.rope_pos, 2 * slot0), p(self.ctx_lens, 2 * slot0),
                          p(self.block_tables, 2 * slot0 * self.max_blocks), self.max_blocks, p(self.history, slot0 * self.hist_stride),
                          p(self.hist_len, slot0), self.hist_stride, p(self.step, slot0), p(self.active, slot0),
                          p(self.stop_tokens, slot0), self.max_ctx, self.row_order.data_ptr() if (ordered and slot0 == 0) else None)


class SamplingTensors:
    """Per-request sampling knobs on the device (t3b_sampling_t; tts.py:455-464 SamplingParams)."""
    FIELDS = ("cfg_scale", "repetition_penalty", "temperature", "min_p", "top_p", "seed", "analyzer_text_tokens")

    def __init__(self, max_req: int, device, combine: str = "bf16"):
        f = lambda v: torch.full((max_req,), v, dtype=torch.float32, device=device)
        self.cfg_scale, self.repetition_penalty, self.temperature = f(0.5), f(1.0), f(0.0)
        self.min_p, self.top_p = f(0.0), f(1.0)
        self.seed = torch.zeros(max_req, dtype=torch.int64, device=device)
        self.analyzer_text_tokens = torch.zeros(max_req, dtype=torch.int32, device=device)     # 0 = heuristics off
        self.combine = _lib.T3B_COMBINE_BF16 if combine == "bf16" else _lib.T3B_COMBINE_FP32
        self.c = self.view(0)

    def view(self, slot0: int) -> _lib.Sampling:
        p = lambda t: t.data_ptr() + t.element_size() * slot0
        return _lib.Sampling(p(self.cfg_scale), p(self.repetition_penalty), p(self.temperature), p(self.min_p), p(self.top_p),
                             p(self.seed), p(self.analyzer_text_tokens), self.combine)

    def set(self, slot: int, *, cfg_scale, repetition_penalty, temperature, min_p, top_p, seed, analyzer_text_tokens=0) -> None:
        self.cfg_scale[slot], self.repetition_penalty[slot], self.temperature[slot] = cfg_scale, repetition_penalty, temperature
        self.min_p[slot], self.top_p[slot] = min_p, top_p
        self.seed[slot] = seed & 0x7FFFFFFFFFFFFFFF
        self.analyzer_text_tokens[slot] = analyzer_text_tokens

    def set_many(self, rows: Sequence[dict]) -> None:
        """Slots 0..len(rows)-1 from per-request dicts with the keys of ``set``: one host->device copy per field instead of seven
        scalar stores per request (32 admitted requests: 224 tiny launches, ~3 ms of host time)."""
        n = len(rows)
        dev = self.cfg_scale.device
        for name in ("cfg_scale", "repetition_penalty", "temperature", "min_p", "top_p"):
            getattr(self, name)[:n] = torch.tensor([float(r[name]) for r in rows], dtype=torch.float32).to(dev, non_blocking=True)
        self.seed[:n] = torch.tensor([int(r["seed"]) & 0x7FFFFFFFFFFFFFFF for r in rows], dtype=torch.int64).to(dev, non_blocking=True)
        self.analyzer_text_tokens[:n] = torch.tensor([int(r.get("analyzer_text_tokens", 0)) for r in rows], dtype=torch.int32).to(dev, non_blocking=True)

    def fill(self, n: int, *, cfg_scale, repetition_penalty, temperature, min_p, top_p, seeds: Sequence[int]) -> None:
        self.cfg_scale[:n], self.repetition_penalty[:n], self.temperature[:n] = cfg_scale, repetition_penalty, temperature
        self.min_p[:n], self.top_p[:n] = min_p, top_p
        self.seed[:n] = torch.tensor([s & 0x7FFFFFFFFFFFFFFF for s in seeds], dtype=torch.int64)


class T3B200Model:
    """Drop-in for ``T3VllmModel`` (t3.py:255-716) on one B200."""

    def __init__(self, *, vllm_config=None, prefix: str = "", dims: T3Dims = T3Dims(), t3_config: Optional[T3Config] = None,
                 device="cuda:0", cfg_scale: Optional[float] = None, combine: str = "bf16", quantization: Optional[str] = None,
                 options: Optional[_lib.Options] = None):
        # t3.py:258-263: the reference reads the Llama shapes from vllm_config.model_config.hf_config and forces hidden 1024
        # (the config file says 2048, the cond|uncond packing hack).  The same fields are honoured here when a config is given.
        self.prefix = prefix
        if vllm_config is not None:
            dims, quantization = self._dims_from_vllm_config(vllm_config, dims, quantization)
        self.dims = dims
        self.cfg = t3_config or T3Config.multilingual()
        self.device = torch.device(device)
        # t3.py:296: CFG scale comes from the environment, default 0.5
        self.cfg_scale = float(os.environ.get("CHATTERBOX_CFG_SCALE", "0.5")) if cfg_scale is None else float(cfg_scale)
        self.combine = combine
        if quantization not in (None, "awq"):
            raise ValueError(f"quantization {quantization!r} not supported (None or 'awq')")
        self.quantization = quantization
        self.options = options                      # t3b_options_t (None = library defaults)
        self.w: Dict[str, torch.Tensor] = {}        # device tensors the C structs point into
        self.aux_weights: Dict[str, torch.Tensor] = {}   # cond_enc.* etc: not on the decode path, kept on CPU
        self._layers = None
        self.c: Optional[_lib.Model] = None
        self._ws: Optional[torch.Tensor] = None
        self._bound: Optional[Tuple[KVPool, DecodeState]] = None

    @staticmethod
    def _dims_from_vllm_config(vllm_config, dims: T3Dims, quantization: Optional[str]):
        """Shapes from a vLLM config object (``vllm_config.model_config.hf_config``, t3.py:259-266) or a plain dict
        with the keys of t3-model*/config.json."""
        hf = getattr(getattr(vllm_config, "model_config", None), "hf_config", None)
        if hf is None:
            hf = vllm_config.get("hf_config", vllm_config) if isinstance(vllm_config, dict) else None
        if hf is None:
            return dims, quantization
        get = (lambda k, d: hf.get(k, d)) if isinstance(hf, dict) else (lambda k, d: getattr(hf, k, d))
        rs = get("rope_scaling", None) or {}
        rs_get = (lambda k, d: rs.get(k, d)) if isinstance(rs, dict) else (lambda k, d: getattr(rs, k, d))
        heads = int(get("num_attention_heads", dims.n_heads))
        new = dims.small(n_layers=int(get("num_hidden_layers", dims.n_layers)), n_heads=heads, hidden=heads * dims.head_dim,
                         ffn=int(get("intermediate_size", dims.ffn)), rms_eps=float(get("rms_norm_eps", dims.rms_eps)),
                         rope_theta=float(get("rope_theta", dims.rope_theta)), rope_factor=float(rs_get("factor", dims.rope_factor)),
                         rope_low=float(rs_get("low_freq_factor", dims.rope_low)), rope_high=float(rs_get("high_freq_factor", dims.rope_high)),
                         rope_orig_max=int(rs_get("original_max_position_embeddings", dims.rope_orig_max)))
        qc = getattr(vllm_config, "quant_config", None)
        if quantization is None and qc is not None and "awq" in type(qc).__name__.lower():
            quantization = "awq"
        return new, quantization

    # ------------------------------------------------------------------ weights
    def load_weights(self, weights: Iterable[Tuple[str, torch.Tensor]]) -> Set[str]:
        """t3.py:300-332.  Accepts the reference's safetensors key names; ``tfmr.*`` uses the HF Llama
        names and q/k/v, gate/up are stacked like vLLM does (llama.py:437-443).  Unknown prefixes
        (e.g. ``text_head.``) are skipped silently like the reference's ``hasattr`` check (:316-319)."""
        d = self.dims
        loaded: Set[str] = set()
        raw: Dict[str, torch.Tensor] = {}
        for name, t in weights:
            head = name.split(".", 1)[0]
            if head == "tfmr" or name in ("speech_emb.weight", "speech_pos_emb.emb.weight", "speech_head.weight",
                                          "text_emb.weight", "text_pos_emb.emb.weight"):
                raw[name] = t
                loaded.add(name)
            elif head == "cond_enc":
                self.aux_weights[name] = t.detach().cpu()
                loaded.add(name)
        dev = self.device
        bf = lambda t: t.detach().to(device=dev, dtype=torch.bfloat16).contiguous()
        need = lambda k: raw[k] if k in raw else (_ for _ in ()).throw(KeyError(f"load_weights: missing {k}"))
        layers = (_lib.Layer * d.n_layers)()
        self._gemv_frag = False
        for i in range(d.n_layers):
            p = f"tfmr.layers.{i}."
            qkv = bf(torch.cat([need(p + "self_attn.q_proj.weight"), need(p + "self_attn.k_proj.weight"),
                                need(p + "self_attn.v_proj.weight")], 0))
            o = bf(need(p + "self_attn.o_proj.weight"))
            gu = bf(torch.cat([need(p + "mlp.gate_proj.weight"), need(p + "mlp.up_proj.weight")], 0))
            dn = bf(need(p + "mlp.down_proj.weight"))
            ln1, ln2 = bf(need(p + "input_layernorm.weight")), bf(need(p + "post_attention_layernorm.weight"))
            assert qkv.shape == (3 * d.hidden, d.hidden) and gu.shape == (2 * d.ffn, d.hidden) and dn.shape == (d.hidden, d.ffn)
            for nm, t in (("qkv", qkv), ("o", o), ("gate_up", gu), ("down", dn)):
                self.w[f"{i}.{nm}"] = t
                setattr(layers[i], nm, self._linear_struct(f"{i}.{nm}", t, gate_up=(nm == "gate_up")))
            self.w[f"{i}.ln1"], self.w[f"{i}.ln2"] = ln1, ln2
            layers[i].ln1, layers[i].ln2 = ln1.data_ptr(), ln2.data_ptr()
        self.w["norm"] = bf(need("tfmr.norm.weight"))
        self.w["speech_emb"] = bf(need("speech_emb.weight"))
        self.w["speech_pos_emb"] = bf(need("speech_pos_emb.emb.weight"))     # precomputed table of t3.py:324-330
        self.w["head"] = bf(need("speech_head.weight"))
        if "text_emb.weight" in raw:
            self.w["text_emb"] = bf(raw["text_emb.weight"])
            self.w["text_pos_emb"] = bf(raw["text_pos_emb.emb.weight"])
        self.w["rope"] = llama3_rope_table(d).to(device=dev, dtype=torch.bfloat16).contiguous()
        assert self.w["speech_emb"].shape == (d.speech_vocab, d.hidden) and self.w["head"].shape == (d.speech_vocab, d.hidden)
        self._layers = layers
        head = _lib.Linear(_lib.T3B_W_BF16, d.speech_vocab, d.hidden, self.w["head"].data_ptr(), None, None, 0)
        self.c = _lib.Model(d.n_layers, d.hidden, d.n_heads, d.head_dim, d.ffn, d.speech_vocab, self.w["speech_pos_emb"].shape[0],
                            d.max_pos, d.rms_eps, layers, self.w["norm"].data_ptr(), head, self.w["speech_emb"].data_ptr(),
                            self.w["speech_pos_emb"].data_ptr(), self.w["rope"].data_ptr())
        return loaded

    def _linear_struct(self, key: str, w: torch.Tensor, gate_up: bool = False) -> _lib.Linear:
        """bf16 [N, K] -> t3b_linear_t.  quantization="awq": the weight is quantised here (no pre-quantised T3
        checkpoint exists; load_awq_tensors() takes AutoAWQ tensors directly) and repacked for the kernel."""
        if self.quantization == "awq":
            from .awq import quantize_and_repack
            return self._awq_struct(key, quantize_and_repack(w, gate_up), w.shape[0], w.shape[1], gate_up)
        return _lib.Linear(_lib.T3B_W_BF16, w.shape[0], w.shape[1], w.data_ptr(), None, None, 0)

    def _awq_struct(self, key: str, q: Dict[str, torch.Tensor], n: int, k: int, gate_up: bool) -> _lib.Linear:
        self.w[key] = q["w"]
        self.w[key + ".scales"], self.w[key + ".zeros"] = q["scales"], q["zeros"]
        return _lib.Linear(_lib.T3B_W_AWQ, n, k, q["w"].data_ptr(), q["scales"].data_ptr(), q["zeros"].data_ptr(),
                           _lib.T3B_LIN_GATE_UP_INTERLEAVED if gate_up else 0)

    def load_awq_tensors(self, layer: int, name: str, qweight: torch.Tensor, qzeros: torch.Tensor, scales: torch.Tensor) -> None:
        """Replace projection ``name`` ("qkv" | "o" | "gate_up" | "down") of ``layer`` by AutoAWQ GEMM tensors
        (qweight int32 [K, N/8], qzeros int32 [K/128, N/8], scales fp16 [K/128, N]; vllm awq.py:214-260), with
        q/k/v and gate/up already concatenated along N like vLLM's stacked parameters."""
        from .awq import repack
        if self._layers is None:
            raise RuntimeError("load_weights() first")
        k, n = qweight.shape[0], qweight.shape[1] * 8
        q = repack(qweight, qzeros, scales, self.device, gate_up=(name == "gate_up"))
        setattr(self._layers[layer], name, self._awq_struct(f"{layer}.{name}", q, n, k, name == "gate_up"))

    def enable_gemv_mma(self) -> bool:
        """Fragment-ordered copies of the bf16 projections and the speech head (t3b_repack_weights_gemv) for the tensor-core GEMV
        of decode steps of <= 8 rows (t3b_options_t.gemv_mma): a second copy of the weights (+1.0 GB for T3; a step still streams
        one copy), made once.  False (nothing done) for quantised models and shapes the kernel does not take."""
        L = self._require()
        if self.quantization is not None:
            return False
        if getattr(self, "_gemv_frag", False):
            return True
        d = self.dims
        jobs = [(f"{i}.{nm}", getattr(self._layers[i], nm), nm == "gate_up") for i in range(d.n_layers) for nm in ("qkv", "o", "gate_up", "down")]
        jobs.append(("head", self.c.head, False))
        if any(L.t3b_gemv_frag_bytes(lin.n, lin.k, int(gu)) == 0 for _, lin, gu in jobs):
            return False
        stream = torch.cuda.current_stream(self.device).cuda_stream
        for key, lin, gu in jobs:
            out = torch.empty(L.t3b_gemv_frag_bytes(lin.n, lin.k, int(gu)), dtype=torch.uint8, device=self.device)
            _lib.check(L.t3b_repack_weights_gemv(self.w[key].data_ptr(), lin.n, lin.k, int(gu), out.data_ptr(), stream), "t3b_repack_weights_gemv")
            self.w["frag." + key] = out
            lin.scales = out.data_ptr()
        self._gemv_frag = True
        return True

    def weight_bytes(self) -> int:
        """Bytes streamed from HBM per decode step (layers + final norm + head)."""
        keep = lambda k: k[0].isdigit() or k in ("norm", "head")
        return sum(t.numel() * t.element_size() for k, t in self.w.items() if keep(k))

    # ------------------------------------------------------------------ plumbing
    def _require(self):
        if self.c is None:
            raise RuntimeError("T3B200Model: load_weights() has not been called")
        if self.device.type != "cuda":
            raise T3BLibraryError("T3B200Model computes on a CUDA device only (no CPU fallback)")
        return _lib.lib()

    def workspace(self, rows: int) -> torch.Tensor:
        """The model's own scratch for direct calls (grows on demand).  An engine never uses it: captured CUDA graphs hold
        the workspace address, so ``T3B200Engine`` owns one buffer sized once for its largest launch."""
        L = self._require()
        need = L.t3b_workspace_bytes(C.byref(self.c), rows)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._ws

    def workspace_bytes(self, rows: int) -> int:
        return int(self._require().t3b_workspace_bytes(C.byref(self.c), rows))

    def _opt(self, opt=None):
        """t3b_options_t* of a call: the caller's (an engine fixes its launch shapes once), else the model's, else NULL."""
        o = opt if opt is not None else self.options
        return C.byref(o) if o is not None else None

    def bind(self, kv: KVPool, state: DecodeState) -> None:
        """Attach the KV pool / decode state that ``forward`` uses (vLLM passes these through its
        forward context; the reference never sees them, t3.py:703-708)."""
        self._bound = (kv, state)

    # ------------------------------------------------------------------ native-row entry points
    def embed(self, table: str, pos_table: str, ids: torch.Tensor, pos: torch.Tensor, dup: int = 1) -> torch.Tensor:
        """E1: out[dup*b+j] = table[ids[b]] + pos_table[pos[b] % rows]  (t3.py:440,475-480)."""
        L = self._require()
        emb, pe = self.w[table], self.w[pos_table]
        n = ids.numel()
        out = torch.empty(n * dup, self.dims.hidden, dtype=torch.bfloat16, device=self.device)
        _lib.check(L.t3b_embed_gather(emb.data_ptr(), pe.data_ptr(), ids.data_ptr(), pos.data_ptr(), n, self.dims.hidden,
                                      emb.shape[0], pe.shape[0], dup, out.data_ptr(), _lib.stream_ptr()), "t3b_embed_gather")
        return out

    def forward_tokens(self, kv: KVPool, x: torch.Tensor, tok_row: torch.Tensor, tok_idx: torch.Tensor,
                       rope_pos: torch.Tensor, block_tables: torch.Tensor, seq_start: Optional[torch.Tensor] = None,
                       seq_row: Optional[torch.Tensor] = None, max_seq_len: int = 0, ws: Optional[torch.Tensor] = None,
                       out: Optional[torch.Tensor] = None, opt=None) -> torch.Tensor:
        """F2 over (row, index) tokens: the prefill of new rows.  ``seq_start`` / ``seq_row`` (whole prompts starting at index 0)
        switch the attention to the read-the-prompt-once kernel (t3b_prefill_attn)."""
        L = self._require()
        n = x.shape[0]
        ws = self.workspace(n) if ws is None else ws
        if out is None:
            out = torch.empty(n, self.dims.hidden, dtype=torch.bfloat16, device=self.device)
        n_seq = 0 if seq_start is None else seq_start.numel() - 1
        _lib.check(L.t3b_forward_tokens(C.byref(self.c), C.byref(kv.c), x.data_ptr(), n, tok_row.data_ptr(), tok_idx.data_ptr(),
                                        rope_pos.data_ptr(), block_tables.data_ptr(), block_tables.shape[1], n_seq,
                                        _lib.ptr(seq_start), _lib.ptr(seq_row), max_seq_len, out.data_ptr(),
                                        ws.data_ptr(), ws.numel(), self._opt(opt), _lib.stream_ptr()), "t3b_forward_tokens")
        return out

    def decode_step(self, kv: KVPool, state: DecodeState, rows: int, hidden_out: torch.Tensor,
                    x_in: Optional[torch.Tensor] = None, ws: Optional[torch.Tensor] = None, opt=None) -> torch.Tensor:
        """F1+F2 for one token per CFG row; embeddings gathered from ``state`` unless ``x_in`` is given."""
        L = self._require()
        ws = self.workspace(rows) if ws is None else ws
        _lib.check(L.t3b_decode_step(C.byref(self.c), C.byref(kv.c), C.byref(state.c), _lib.ptr(x_in), rows, hidden_out.data_ptr(),
                                     ws.data_ptr(), ws.numel(), self._opt(opt), _lib.stream_ptr()), "t3b_decode_step")
        return hidden_out

    def cfg_sample_head_logits(self, head_logits: torch.Tensor, rows: int, sampling_c, state_c, out_tokens_ptr: int, advance: bool = True,
                               logits_out: Optional[torch.Tensor] = None) -> None:
        L = self._require()
        _lib.check(L.t3b_cfg_sample_head_logits(head_logits.data_ptr(), rows, self.dims.speech_vocab, C.byref(sampling_c), C.byref(state_c),
                                                1 if advance else 0, out_tokens_ptr, _lib.ptr(logits_out), _lib.stream_ptr()),
                   "t3b_cfg_sample_head_logits")

    def decode_and_sample(self, kv: KVPool, state: DecodeState, sampling: "SamplingTensors", n_slots: int, hidden: torch.Tensor,
                          out_tokens: torch.Tensor, ws: torch.Tensor, opt=None) -> None:
        """One whole decode step (F1+F2, C1, S1..S4, state advance) for slots [0, n_slots) on the current stream."""
        L = self._require()
        rows = 2 * n_slots
        s = _lib.stream_ptr()
        _lib.check(L.t3b_decode_step(C.byref(self.c), C.byref(kv.c), C.byref(state.c), None, rows, hidden.data_ptr(), ws.data_ptr(),
                                     ws.numel(), self._opt(opt), s), "t3b_decode_step")
        _lib.check(L.t3b_head_cfg_sample(C.byref(self.c), hidden.data_ptr(), rows, C.byref(sampling.c), C.byref(state.c), 1,
                                         out_tokens.data_ptr(), None, ws.data_ptr(), ws.numel(), self._opt(opt), s), "t3b_head_cfg_sample")

    def head_cfg_sample(self, hidden: torch.Tensor, rows: int, sampling: SamplingTensors, state: Optional[DecodeState],
                        out_tokens: Optional[torch.Tensor], advance: bool = False,
                        logits_out: Optional[torch.Tensor] = None, ws: Optional[torch.Tensor] = None, opt=None) -> Optional[torch.Tensor]:
        """C1 + S1..S4 fused: speech head on both rows, CFG combine, penalties, sampling."""
        L = self._require()
        ws = self.workspace(rows) if ws is None else ws
        _lib.check(L.t3b_head_cfg_sample(C.byref(self.c), hidden.data_ptr(), rows, C.byref(sampling.c),
                                         C.byref(state.c) if state is not None else None, 1 if advance else 0,
                                         _lib.ptr(out_tokens), _lib.ptr(logits_out), ws.data_ptr(), ws.numel(), self._opt(opt),
                                         _lib.stream_ptr()), "t3b_head_cfg_sample")
        return out_tokens

    def prefill_embeds(self, cond_emb: torch.Tensor, text_ids: Sequence[int]) -> Tuple[torch.Tensor, torch.Tensor]:
        """t3.py:542-561: cond = [cond_emb(34) | text_emb+text_pos | BOS], uncond = [cond_emb | zeros | BOS]."""
        x, _, _, starts = self.prefill_assemble([cond_emb], [text_ids], [0])
        return x[starts[0]:starts[1]], x[starts[1]:starts[2]]

    def prefill_assemble(self, conds: Sequence[torch.Tensor], text_ids: Sequence[Sequence[int]], slots: Sequence[int]):
        """Prefill rows of n requests in ONE kernel launch (t3b_prefill_embed; layout t3.py:542-561): two host->device copies
        (conditionals; ids + offsets + slots in one int32 buffer), no per-request device work.
        Returns (x bf16 [n_tok, D], tok_row int32 [n_tok], tok_idx int32 [n_tok], starts: host list of 2n+1 sequence offsets);
        sequence 2r / 2r+1 = the cond / uncond row of request r, KV row 2 * slots[r] (+1).  ``self.last_seq`` holds the device
        copies (seq_start int32 [2n+1], seq_row int32 [2n]) that ``forward_tokens`` takes."""
        L = self._require()
        if "text_emb" not in self.w:
            raise RuntimeError("prefill needs text_emb.weight / text_pos_emb.emb.weight in the checkpoint")
        dev, D, C34 = self.device, self.dims.hidden, CONDITIONING_SIZE
        n = len(conds)
        lens = [len(t) for t in text_ids]
        starts, tstart = [0], [0]
        for l in lens:
            starts += [starts[-1] + C34 + l + 1, starts[-1] + 2 * (C34 + l + 1)]
            tstart.append(tstart[-1] + l)
        n_tok = starts[-1]
        flat = [t for ids in text_ids for t in ids] or [0]
        seq_rows = [2 * sl + half for sl in slots for half in (0, 1)]
        meta = torch.tensor(flat + tstart + starts + list(slots) + seq_rows, dtype=torch.int32).pin_memory().to(dev, non_blocking=True)
        o_t, o_s, o_o = len(flat), len(flat) + n + 1, len(flat) + n + 1 + 2 * n + 1
        self.last_seq = (meta[o_s:o_o], meta[o_o + n:o_o + 3 * n])
        for c in conds:
            if c.shape[0] < C34 or c.shape[1] != D:
                raise ValueError(f"conditionals must be [>= {C34}, {D}], got {tuple(c.shape)}")
        ce = torch.stack([c[:C34].to(torch.bfloat16) for c in conds])
        if ce.device.type == "cpu":
            ce = ce.pin_memory()
        ce = ce.to(dev, non_blocking=True).contiguous()
        x = torch.empty(n_tok, D, dtype=torch.bfloat16, device=dev)
        tok_row = torch.empty(n_tok, dtype=torch.int32, device=dev)
        tok_idx = torch.empty(n_tok, dtype=torch.int32, device=dev)
        p = meta.data_ptr()
        _lib.check(L.t3b_prefill_embed(ce.data_ptr(), p, p + 4 * o_t, p + 4 * o_s, p + 4 * o_o, n, self.w["text_emb"].data_ptr(),
                                       self.w["text_pos_emb"].data_ptr(), self.w["text_emb"].shape[0], self.w["text_pos_emb"].shape[0],
                                       self.w["speech_emb"].data_ptr(), self.w["speech_pos_emb"].data_ptr(), START_SPEECH_TOKEN, D, C34,
                                       x.data_ptr(), tok_row.data_ptr(), tok_idx.data_ptr(), _lib.stream_ptr()), "t3b_prefill_embed")
        return x, tok_row, tok_idx, starts

    def prefill_embeds_batch(self, conds: Sequence[torch.Tensor], text_ids: Sequence[Sequence[int]]) -> List[Tuple[torch.Tensor, torch.Tensor]]:
        """Per-request (cond, uncond) views of ``prefill_assemble``."""
        x, _, _, st = self.prefill_assemble(conds, text_ids, list(range(len(conds))))
        return [(x[st[2 * i]:st[2 * i + 1]], x[st[2 * i + 1]:st[2 * i + 2]]) for i in range(len(conds))]

    def filter_speech_tokens(self, token_ids: torch.Tensor) -> torch.Tensor:
        """S3Gen hand-off (tts.py:492-514): offset ids -> un-offset ids in [0, 6561), order kept."""
        L = self._require()
        ids = token_ids.to(device=self.device, dtype=torch.int32).contiguous()
        out = torch.empty_like(ids)
        n_out = torch.zeros(1, dtype=torch.int32, device=self.device)
        _lib.check(L.t3b_filter_speech_tokens(ids.data_ptr(), ids.numel(), SPEECH_TOKEN_OFFSET, self.cfg.start_speech_token, out.data_ptr(),
                                              n_out.data_ptr(), _lib.stream_ptr()), "t3b_filter_speech_tokens")
        return out[:int(n_out.item())]

    # ------------------------------------------------------------------ the reference's operator surface
    def get_language_model(self):                                 # t3.py:715
        return self

    def get_multimodal_embeddings(self, **kwargs) -> List[torch.Tensor]:      # t3.py:335-337
        cond = kwargs.get("conditionals")
        return [] if cond is None else [c for c in cond]

    def get_input_embeddings(self, input_ids: torch.Tensor, multimodal_embeddings=None, *,
                             speech_pos: Optional[torch.Tensor] = None) -> torch.Tensor:
        """t3.py:424-647.  Decode ids are offset speech ids (>= 2500): bf16 [N, 2048] with the same embedding in both
        halves (t3.py:480).  ``speech_pos`` is the explicit positional index the reference fails to track (SURVEY.md
        0.3-2); None = the reference's literal fallback, index 0 for every token.
        With ``multimodal_embeddings`` (the stretched conditionals of ``packing.pack_prompt`` / t3.py:212-221) the batch
        may hold prefill blocks: it is cut like ``split_prefill_decode`` (t3.py:340-421) and every block gets the cond | uncond
        rows of t3.py:542-639, including the three partial-chunk cases."""
        from . import packing
        ids_all = input_ids.to("cpu").to(torch.int64).tolist() if input_ids.device.type != "cpu" else input_ids.to(torch.int64).tolist()
        if multimodal_embeddings is None or len(multimodal_embeddings) == 0:
            if any(t < SPEECH_TOKEN_OFFSET for t in ids_all):
                raise ValueError("get_input_embeddings: decode ids must be speech ids (>= SPEECH_TOKEN_OFFSET)")
            ids = torch.tensor([t - SPEECH_TOKEN_OFFSET for t in ids_all], dtype=torch.int32, device=self.device)
            pos = torch.zeros_like(ids) if speech_pos is None else speech_pos.to(device=self.device, dtype=torch.int32)
            return self.embed("speech_emb", "speech_pos_emb", ids, pos, dup=2).view(ids.numel(), 2 * self.dims.hidden)
        D, dev = self.dims.hidden, self.device
        out: List[torch.Tensor] = []
        done = 0
        for blk, rows in packing.split_prefill_decode(ids_all, multimodal_embeddings):
            n = len(blk)
            if rows is None:                                   # a decode run inside a mixed batch (t3.py:497-527)
                ids = torch.tensor([t - SPEECH_TOKEN_OFFSET for t in blk], dtype=torch.int32, device=dev)
                pos = torch.zeros_like(ids) if speech_pos is None else speech_pos[done:done + n].to(device=dev, dtype=torch.int32)
                out.append(self.embed("speech_emb", "speech_pos_emb", ids, pos, dup=2).view(n, 2 * D))
            else:
                n_cond, text, tpos, has_bos = packing.prefill_block_plan(blk, rows)
                parts_c, parts_u = [], []
                if n_cond:
                    ce = rows[:n_cond].to(device=dev, dtype=torch.bfloat16)
                    parts_c.append(ce); parts_u.append(ce)
                if text:
                    te = self.embed("text_emb", "text_pos_emb", torch.tensor(text, dtype=torch.int32, device=dev),
                                    torch.tensor(tpos, dtype=torch.int32, device=dev))
                    parts_c.append(te); parts_u.append(torch.zeros_like(te))          # t3.py:556
                if has_bos:
                    bos = self.embed("speech_emb", "speech_pos_emb", torch.tensor([START_SPEECH_TOKEN], dtype=torch.int32, device=dev),
                                     torch.zeros(1, dtype=torch.int32, device=dev))
                    parts_c.append(bos); parts_u.append(bos)
                out.append(torch.cat([torch.cat(parts_c, 0), torch.cat(parts_u, 0)], dim=1))    # t3.py:560 side by side
            done += n
        return torch.cat(out, dim=0)

    def forward(self, input_ids, positions: torch.Tensor, intermediate_tensors=None,
                inputs_embeds: Optional[torch.Tensor] = None) -> torch.Tensor:
        """t3.py:676-713 for a decode batch: ``inputs_embeds`` [N, 2048] -> hidden [N, 2048].  The N
        requests are slots 0..N-1 of the bound state; both CFG rows share ``positions`` (t3.py:705)."""
        if inputs_embeds is None:
            raise ValueError("forward: inputs_embeds is required (the reference passes input_ids=None, t3.py:704)")
        if self._bound is None:
            raise RuntimeError("forward: bind(kv, state) first")
        kv, st = self._bound
        n = inputs_embeds.shape[0]
        x = inputs_embeds.view(2 * n, self.dims.hidden)
        st.rope_pos[:2 * n] = positions.to(device=self.device, dtype=torch.int32).repeat_interleave(2)
        out = torch.empty(2 * n, self.dims.hidden, dtype=torch.bfloat16, device=self.device)
        self.decode_step(kv, st, 2 * n, out, x_in=x)
        return out.view(n, 2 * self.dims.hidden)

    def compute_logits(self, hidden_states: torch.Tensor, sampling_metadata=None, *, pad: bool = True) -> torch.Tensor:
        """t3.py:650-673: head on both halves, cond + w (cond - uncond), left-pad 2500 x -inf."""
        m = hidden_states.shape[0]
        sp = SamplingTensors(m, self.device, self.combine)
        sp.cfg_scale.fill_(self.cfg_scale)
        logits = torch.empty(m, self.dims.speech_vocab, dtype=torch.float32, device=self.device)
        self.head_cfg_sample(hidden_states.view(2 * m, self.dims.hidden), 2 * m, sp, None, None, False, logits)
        out = logits.to(torch.bfloat16)
        if not pad:
            return out
        padc = torch.full((m, SPEECH_TOKEN_OFFSET), float("-inf"), dtype=out.dtype, device=out.device)
        return torch.cat([padc, out], dim=1)

    def sample(self, hidden_states: torch.Tensor, sampling: SamplingTensors, state: Optional[DecodeState] = None,
               advance: bool = False) -> torch.Tensor:
        """compute_logits + the engine's sampler in one call; returns int32 [M] un-offset speech ids
        (add SPEECH_TOKEN_OFFSET for the ids the reference's caller sees, tts.py:474-492)."""
        m = hidden_states.shape[0] if hidden_states.shape[1] == 2 * self.dims.hidden else hidden_states.shape[0] // 2
        out = torch.empty(m, dtype=torch.int32, device=self.device)
        self.head_cfg_sample(hidden_states.reshape(2 * m, self.dims.hidden), 2 * m, sampling, state, out, advance)
        return out
