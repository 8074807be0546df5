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
    """Paged KV pool: K and V bf16 [n_layers, num_blocks, n_heads, 16, 64] (include/t3b200.h t3b_kv_t).
    Block 0 is the scratch block idle rows write to; the allocator hands out 1..num_blocks-1."""

    def __init__(self, dims: T3Dims, num_blocks: int, device):
        shape = (dims.n_layers, num_blocks, dims.n_heads, KV_BLOCK, dims.head_dim)
        self.k = torch.zeros(shape, dtype=torch.bfloat16, device=device)
        self.v = torch.zeros(shape, dtype=torch.bfloat16, device=device)
        self.num_blocks = num_blocks
        self.c = _lib.KV(self.k.data_ptr(), self.v.data_ptr(), num_blocks * dims.n_heads * KV_BLOCK * dims.head_dim, num_blocks)
        self.free: List[int] = list(range(num_blocks - 1, 0, -1))

    @staticmethod
    def bytes_per_block(dims: T3Dims) -> int:
        return 2 * dims.n_layers * dims.n_heads * KV_BLOCK * dims.head_dim * 2

    def alloc(self, n: int) -> List[int]:
        if n > len(self.free):
            raise MemoryError(f"KV pool exhausted: need {n} blocks, {len(self.free)} free")
        out = self.free[-n:]
        del self.free[-n:]
        return out

    def release(self, blocks: Sequence[int]) -> None:
        self.free.extend(blocks)


class DecodeState:
    """Device-resident per-request decode state (t3b_state_t) for up to ``max_req`` requests."""

    def __init__(self, max_req: int, max_blocks: int, hist_stride: int, max_ctx: int, device,
                 stop_token: int = STOP_SPEECH_TOKEN):
        z = lambda *s: torch.zeros(*s, dtype=torch.int32, device=device)
        self.max_req, self.max_blocks, self.hist_stride, self.max_ctx = max_req, max_blocks, hist_stride, max_ctx
        self.tokens, self.speech_pos = z(max_req), z(max_req)
        self.rope_pos, self.ctx_lens = z(2 * max_req), z(2 * max_req)
        self.block_tables = z(2 * max_req, max_blocks)
        self.history, self.hist_len = z(max_req, hist_stride), z(max_req)
        self.step, self.active = z(max_req), z(max_req)
        self.c = self.view(0, stop_token)

    def view(self, slot0: int, stop_token: Optional[int] = None) -> _lib.State:
        """t3b_state_t whose request 0 is slot ``slot0`` (a lane of the decode step works on a slice of slots)."""
        p = lambda t, n: t.data_ptr() + 4 * n
        stop = self.c.stop_token if stop_token is None else stop_token
        return _lib.State(p(self.tokens, slot0), p(self.speech_pos, slot0), p(self.rope_pos, 2 * slot0), p(self.ctx_lens, 2 * slot0),
                          p(self.block_tables, 2 * slot0 * self.max_blocks), self.max_blocks, p(self.history, slot0 * self.hist_stride),
                          p(self.hist_len, slot0), self.hist_stride, p(self.step, slot0), p(self.active, slot0), stop, self.max_ctx)


class SamplingTensors:
    """Per-request sampling knobs on the device (t3b_sampling_t; tts.py:455-464 SamplingParams)."""

    def __init__(self, max_req: int, device, combine: str = "bf16"):
        f = lambda v: torch.full((max_req,), v, dtype=torch.float32, device=device)
        self.cfg_scale, self.repetition_penalty, self.temperature = f(0.5), f(1.0), f(0.0)
        self.min_p, self.top_p = f(0.0), f(1.0)
        self.seed = torch.zeros(max_req, dtype=torch.int64, device=device)
        self.combine = _lib.T3B_COMBINE_BF16 if combine == "bf16" else _lib.T3B_COMBINE_FP32
        self.c = self.view(0)

    def view(self, slot0: int) -> _lib.Sampling:
        p = lambda t: t.data_ptr() + t.element_size() * slot0
        return _lib.Sampling(p(self.cfg_scale), p(self.repetition_penalty), p(self.temperature), p(self.min_p), p(self.top_p),
                             p(self.seed), self.combine)

    def set(self, slot: int, *, cfg_scale, repetition_penalty, temperature, min_p, top_p, seed) -> None:
        self.cfg_scale[slot], self.repetition_penalty[slot], self.temperature[slot] = cfg_scale, repetition_penalty, temperature
        self.min_p[slot], self.top_p[slot] = min_p, top_p
        self.seed[slot] = seed & 0x7FFFFFFFFFFFFFFF

    def fill(self, n: int, *, cfg_scale, repetition_penalty, temperature, min_p, top_p, seeds: Sequence[int]) -> None:
        self.cfg_scale[:n], self.repetition_penalty[:n], self.temperature[:n] = cfg_scale, repetition_penalty, temperature
        self.min_p[:n], self.top_p[:n] = min_p, top_p
        self.seed[:n] = torch.tensor([s & 0x7FFFFFFFFFFFFFFF for s in seeds], dtype=torch.int64)


class T3B200Model:
    """Drop-in for ``T3VllmModel`` (t3.py:255-716) on one B200."""

    def __init__(self, *, vllm_config=None, prefix: str = "", dims: T3Dims = T3Dims(), t3_config: Optional[T3Config] = None,
                 device="cuda:0", cfg_scale: Optional[float] = None, combine: str = "bf16", quantization: Optional[str] = None):
        # vllm_config / prefix are accepted for signature compatibility with t3.py:258 and ignored.
        self.dims = dims
        self.cfg = t3_config or T3Config.multilingual()
        self.device = torch.device(device)
        # t3.py:296: CFG scale comes from the environment, default 0.5
        self.cfg_scale = float(os.environ.get("CHATTERBOX_CFG_SCALE", "0.5")) if cfg_scale is None else float(cfg_scale)
        self.combine = combine
        if quantization not in (None, "awq"):
            raise ValueError(f"quantization {quantization!r} not supported (None or 'awq')")
        self.quantization = quantization
        self.w: Dict[str, torch.Tensor] = {}        # device tensors the C structs point into
        self.aux_weights: Dict[str, torch.Tensor] = {}   # cond_enc.* etc: not on the decode path, kept on CPU
        self._layers = None
        self.c: Optional[_lib.Model] = None
        self._ws: Optional[torch.Tensor] = None
        self._lane_ws: Dict[int, torch.Tensor] = {}
        self.mega: Optional[_lib.Mega] = None       # repacked weights of the persistent whole-step kernel (bf16 only)
        self._mega_ws: Optional[torch.Tensor] = None
        self.fused: Optional[torch.Tensor] = None   # weight blob of the fused step kernel (csrc/decode_fused.cu, bf16 only)
        self._fused_ws: Optional[torch.Tensor] = None
        self._bound: Optional[Tuple[KVPool, DecodeState]] = None

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
        self.mega = None
        self.fused = None
        if self.quantization is None and dev.type == "cuda":
            if os.environ.get("T3B_MEGA", "0") == "1":
                self._build_mega()
            # opt-in (T3B_FUSED=1): the fused persistent kernel hit a cluster-barrier timeout in the 300-step bench at
            # growing context (profiles/r01d_notes.md), so the PDL-chained kernels are the default decode path
            if os.environ.get("T3B_FUSED", "0") == "1":
                self._build_fused()
        return loaded

    def _build_fused(self) -> None:
        """Per-CTA weight streams of csrc/decode_fused.cu (t3b_fused_repack); None if the shape is not the T3 one."""
        L = _lib.lib()
        n = L.t3b_fused_weight_bytes(C.byref(self.c))
        if n == 0:
            return
        blob = torch.empty(n, dtype=torch.uint8, device=self.device)
        _lib.check(L.t3b_fused_repack(C.byref(self.c), blob.data_ptr(), _lib.stream_ptr()), "t3b_fused_repack")
        torch.cuda.synchronize(self.device)
        self.fused = blob

    def _build_mega(self) -> None:
        """Unit layout of csrc/decode_mega.cu: every projection as 16 KB blocks of 128 x 64 weights (t3b_repack_weights_bf16)."""
        L, d = _lib.lib(), self.dims
        s = _lib.stream_ptr()

        def rp(key: str, w: torch.Tensor, gate_up: bool = False) -> int:
            n, k = w.shape
            out = torch.empty(L.t3b_repacked_bytes(n, k), dtype=torch.uint8, device=self.device)
            _lib.check(L.t3b_repack_weights_bf16(w.data_ptr(), n, k, 1 if gate_up else 0, out.data_ptr(), s), "t3b_repack_weights_bf16")
            self.w["mega." + key] = out
            return out.data_ptr()

        table = torch.empty(d.n_layers, 6, dtype=torch.int64)
        for i in range(d.n_layers):
            table[i] = torch.tensor([rp(f"{i}.qkv", self.w[f"{i}.qkv"]), rp(f"{i}.o", self.w[f"{i}.o"]),
                                     rp(f"{i}.gate_up", self.w[f"{i}.gate_up"], True), rp(f"{i}.down", self.w[f"{i}.down"]),
                                     self.w[f"{i}.ln1"].data_ptr(), self.w[f"{i}.ln2"].data_ptr()], dtype=torch.int64)
        self.w["mega.table"] = table.to(self.device)
        self.mega = _lib.Mega(self.w["mega.table"].data_ptr(), rp("head", self.w["head"]))
        torch.cuda.synchronize(self.device)

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

    def weight_bytes(self) -> int:
        """Bytes streamed from HBM per decode step (layers + final norm + head)."""
        keep = lambda k: k[0].isdigit() or k in ("norm", "head")     # "mega.*" / fused copies are the same weights in another layout
        return sum(t.numel() * t.element_size() for k, t in self.w.items() if keep(k))

    # ------------------------------------------------------------------ plumbing
    def _require(self):
        if self.c is None:
            raise RuntimeError("T3B200Model: load_weights() has not been called")
        if self.device.type != "cuda":
            raise T3BLibraryError("T3B200Model computes on a CUDA device only (no CPU fallback)")
        return _lib.lib()

    def workspace(self, rows: int) -> torch.Tensor:
        L = self._require()
        need = L.t3b_workspace_bytes(C.byref(self.c), rows)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._ws

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

    def forward_tokens(self, kv: KVPool, x: torch.Tensor, tok_row: Optional[torch.Tensor], tok_idx: torch.Tensor,
                       rope_pos: torch.Tensor, block_tables: torch.Tensor) -> torch.Tensor:
        """F2 over arbitrary (row, index) tokens: prefill of new rows or a decode step."""
        L = self._require()
        n = x.shape[0]
        ws = self.workspace(n)
        out = torch.empty(n, self.dims.hidden, dtype=torch.bfloat16, device=self.device)
        _lib.check(L.t3b_forward_tokens(C.byref(self.c), C.byref(kv.c), x.data_ptr(), n, _lib.ptr(tok_row), tok_idx.data_ptr(),
                                        rope_pos.data_ptr(), block_tables.data_ptr(), block_tables.shape[1], out.data_ptr(),
                                        ws.data_ptr(), ws.numel(), _lib.stream_ptr()), "t3b_forward_tokens")
        return out

    def decode_step(self, kv: KVPool, state: DecodeState, rows: int, hidden_out: torch.Tensor,
                    x_in: Optional[torch.Tensor] = None) -> torch.Tensor:
        """F1+F2 for one token per CFG row; embeddings gathered from ``state`` unless ``x_in`` is given."""
        L = self._require()
        ws = self.workspace(rows)
        _lib.check(L.t3b_decode_step(C.byref(self.c), C.byref(kv.c), C.byref(state.c), _lib.ptr(x_in), rows, hidden_out.data_ptr(),
                                     ws.data_ptr(), ws.numel(), _lib.stream_ptr()), "t3b_decode_step")
        return hidden_out

    def mega_logits(self, rows: int) -> torch.Tensor:
        """Persistent bf16 [rows, speech_vocab] head-logit buffer (CUDA graphs hold its address)."""
        if getattr(self, "_mega_lg", None) is None or self._mega_lg.shape[0] < rows:
            self._mega_lg = torch.empty(max(rows, 2), self.dims.speech_vocab, dtype=torch.bfloat16, device=self.device)
        return self._mega_lg

    def mega_ok(self, rows: int) -> bool:
        return self.mega is not None and 0 < rows <= _lib.lib().t3b_mega_max_rows()

    def mega_workspace(self, rows: int) -> torch.Tensor:
        L = self._require()
        need = L.t3b_mega_workspace_bytes(C.byref(self.c), rows)
        if self._mega_ws is None or self._mega_ws.numel() < need:
            self._mega_ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._mega_ws

    def decode_step_mega(self, kv: KVPool, state_c, rows: int, hidden_out: Optional[torch.Tensor], logits_out: torch.Tensor,
                         x_in: Optional[torch.Tensor] = None) -> torch.Tensor:
        """F1+F2+head in ONE persistent launch (csrc/decode_mega.cu): logits_out bf16 [rows, speech_vocab]."""
        L = self._require()
        if not self.mega_ok(rows):
            raise T3BLibraryError(f"decode_step_mega: rows {rows} not supported (bf16 weights, rows <= {L.t3b_mega_max_rows()})")
        ws = self.mega_workspace(rows)
        _lib.check(L.t3b_decode_step_mega(C.byref(self.c), C.byref(self.mega), C.byref(kv.c), C.byref(state_c), _lib.ptr(x_in), rows,
                                          _lib.ptr(hidden_out), logits_out.data_ptr(), ws.data_ptr(), ws.numel(), _lib.stream_ptr()),
                   "t3b_decode_step_mega")
        return logits_out

    def fused_ok(self, rows: int) -> bool:
        return self.fused is not None and 0 < rows <= _lib.lib().t3b_fused_max_rows()

    def fused_workspace(self, rows: int) -> torch.Tensor:
        L = self._require()
        need = L.t3b_fused_workspace_bytes(C.byref(self.c), rows)
        if self._fused_ws is None or self._fused_ws.numel() < need:
            self._fused_ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._fused_ws

    def decode_step_fused(self, kv: KVPool, state_c, rows: int, hidden_out: Optional[torch.Tensor], logits_out: torch.Tensor,
                          x_in: Optional[torch.Tensor] = None) -> torch.Tensor:
        """F1+F2+head in ONE persistent launch with two global exchanges per layer (csrc/decode_fused.cu):
        logits_out bf16 [rows, speech_vocab]."""
        L = self._require()
        if not self.fused_ok(rows):
            raise T3BLibraryError(f"decode_step_fused: rows {rows} not supported (bf16 T3 shape, rows <= {L.t3b_fused_max_rows()})")
        ws = self.fused_workspace(rows)
        _lib.check(L.t3b_decode_step_fused(C.byref(self.c), self.fused.data_ptr(), C.byref(kv.c), C.byref(state_c), _lib.ptr(x_in), rows,
                                           _lib.ptr(hidden_out), logits_out.data_ptr(), ws.data_ptr(), ws.numel(), _lib.stream_ptr()),
                   "t3b_decode_step_fused")
        return logits_out

    def cfg_sample_head_logits(self, head_logits: torch.Tensor, rows: int, sampling_c, state_c, out_tokens_ptr: int, advance: bool = True,
                               logits_out: Optional[torch.Tensor] = None) -> None:
        L = self._require()
        _lib.check(L.t3b_cfg_sample_head_logits(head_logits.data_ptr(), rows, self.dims.speech_vocab, C.byref(sampling_c), C.byref(state_c),
                                                1 if advance else 0, out_tokens_ptr, _lib.ptr(logits_out), _lib.stream_ptr()),
                   "t3b_cfg_sample_head_logits")

    def lane_workspace(self, lane: int, rows: int) -> torch.Tensor:
        """Scratch of decode lane ``lane`` (lanes run concurrently on their own streams, so each owns one)."""
        L = self._require()
        need = L.t3b_workspace_bytes(C.byref(self.c), rows)
        ws = self._lane_ws.get(lane)
        if ws is None or ws.numel() < need:
            ws = self._lane_ws[lane] = torch.empty(need, dtype=torch.uint8, device=self.device)
        return ws

    def decode_lane(self, kv: KVPool, state: DecodeState, sampling: "SamplingTensors", slot0: int, n_slots: int, lane: int,
                    hidden: torch.Tensor, out_tokens: torch.Tensor) -> None:
        """One whole decode step (F1+F2, C1, S1..S4, state advance) for slots [slot0, slot0+n_slots) on the current stream."""
        L = self._require()
        rows = 2 * n_slots
        ws = self.lane_workspace(lane, rows)
        st, sp = state.view(slot0), sampling.view(slot0)
        # The persistent kernel is opt-in: measured on B200 it is slower than the PDL-chained kernels below
        # (profiles/r01c_bench_mega_default.json vs r01c_bench_multikernel.json; DESIGN.md "Persistent step").
        if lane == 0 and slot0 == 0 and self.fused_ok(rows) and os.environ.get("T3B_FUSED", "0") == "1":
            lg = self.mega_logits(rows)
            self.decode_step_fused(kv, st, rows, None, lg)
            self.cfg_sample_head_logits(lg, rows, sp, st, out_tokens.data_ptr())
            return
        if lane == 0 and slot0 == 0 and self.mega_ok(rows) and os.environ.get("T3B_MEGA", "0") == "1":
            lg = self.mega_logits(rows)
            self.decode_step_mega(kv, st, rows, None, lg)
            self.cfg_sample_head_logits(lg, rows, sp, st, out_tokens.data_ptr())
            return
        h = hidden.data_ptr() + 2 * slot0 * self.dims.hidden * 2
        s = _lib.stream_ptr()
        _lib.check(L.t3b_decode_step(C.byref(self.c), C.byref(kv.c), C.byref(st), None, rows, h, ws.data_ptr(), ws.numel(), s),
                   "t3b_decode_step")
        _lib.check(L.t3b_head_cfg_sample(C.byref(self.c), h, rows, C.byref(sp), C.byref(st), 1, out_tokens.data_ptr() + 4 * slot0, None,
                                         ws.data_ptr(), ws.numel(), s), "t3b_head_cfg_sample")

    def head_cfg_sample(self, hidden: torch.Tensor, rows: int, sampling: SamplingTensors, state: Optional[DecodeState],
                        out_tokens: Optional[torch.Tensor], advance: bool = False,
                        logits_out: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
        """C1 + S1..S4 fused: speech head on both rows, CFG combine, penalties, sampling."""
        L = self._require()
        ws = self.workspace(rows)
        _lib.check(L.t3b_head_cfg_sample(C.byref(self.c), hidden.data_ptr(), rows, C.byref(sampling.c),
                                         C.byref(state.c) if state is not None else None, 1 if advance else 0,
                                         _lib.ptr(out_tokens), _lib.ptr(logits_out), ws.data_ptr(), ws.numel(), _lib.stream_ptr()),
                   "t3b_head_cfg_sample")
        return out_tokens

    def prefill_embeds(self, cond_emb: torch.Tensor, text_ids: Sequence[int]) -> Tuple[torch.Tensor, torch.Tensor]:
        """t3.py:542-561: cond = [cond_emb(34) | text_emb+text_pos | BOS], uncond = [cond_emb | zeros | BOS]."""
        dev = self.device
        ids = torch.tensor(list(text_ids), dtype=torch.int32, device=dev)
        pos = torch.arange(len(text_ids), dtype=torch.int32, device=dev)
        text = self.embed("text_emb", "text_pos_emb", ids, pos)
        bos = self.embed("speech_emb", "speech_pos_emb", torch.tensor([START_SPEECH_TOKEN], dtype=torch.int32, device=dev),
                         torch.zeros(1, dtype=torch.int32, device=dev))
        ce = cond_emb.to(device=dev, dtype=torch.bfloat16)
        assert ce.shape == (CONDITIONING_SIZE, self.dims.hidden)
        return torch.cat([ce, text, bos], 0), torch.cat([ce, torch.zeros_like(text), bos], 0)

    def prefill_embeds_batch(self, conds: Sequence[torch.Tensor], text_ids: Sequence[Sequence[int]]) -> List[Tuple[torch.Tensor, torch.Tensor]]:
        """``prefill_embeds`` for many requests with a constant number of device operations: one host->device copy of all text
        ids / positions, one of all conditionals, one embedding gather, three indexed stores.  Row layout per request
        (t3.py:542-561): cond = [cond_emb(34) | text_emb + text_pos | BOS], uncond = [cond_emb | zeros | BOS]."""
        dev, D, C34 = self.device, self.dims.hidden, CONDITIONING_SIZE
        n = len(conds)
        lens = [len(t) for t in text_ids]
        P = [C34 + l + 1 for l in lens]
        big = torch.zeros(2 * sum(P), D, dtype=torch.bfloat16, device=dev)
        ids_all = torch.tensor([t for ids in text_ids for t in ids], dtype=torch.int32)
        pos_all = torch.cat([torch.arange(l, dtype=torch.int32) for l in lens]) if n else torch.zeros(0, dtype=torch.int32)
        ce_all = torch.stack([c[:C34].to(torch.bfloat16) for c in conds]).reshape(n * C34, D)
        if ce_all.device.type == "cpu":
            ce_all = ce_all.pin_memory()
        idx_ce, idx_text, idx_bos, spans = [], [], [], []
        off = 0
        for i in range(n):
            c0, u0 = off, off + P[i]
            idx_ce.append((torch.arange(C34), c0, u0))
            idx_text.append(torch.arange(lens[i]) + c0 + C34)
            idx_bos += [c0 + P[i] - 1, u0 + P[i] - 1]
            spans.append((c0, u0, u0 + P[i]))
            off += 2 * P[i]
        ce_dev = ce_all.to(dev, non_blocking=True)
        i_ce_c = torch.cat([a + c0 for a, c0, _ in idx_ce]).to(dev, non_blocking=True)
        i_ce_u = torch.cat([a + u0 for a, _, u0 in idx_ce]).to(dev, non_blocking=True)
        big.index_copy_(0, i_ce_c, ce_dev)
        big.index_copy_(0, i_ce_u, ce_dev)
        if ids_all.numel():
            text = self.embed("text_emb", "text_pos_emb", ids_all.to(dev), pos_all.to(dev))
            big.index_copy_(0, torch.cat(idx_text).to(dev), text)
        bos = self.embed("speech_emb", "speech_pos_emb", torch.tensor([START_SPEECH_TOKEN], dtype=torch.int32, device=dev),
                         torch.zeros(1, dtype=torch.int32, device=dev))
        big[torch.tensor(idx_bos, device=dev)] = bos
        return [(big[a:b], big[b:c]) for a, b, c in spans]

    # ------------------------------------------------------------------ the reference's operator surface
    def get_language_model(self):                                 # t3.py:715
        return self

    def get_multimodal_embeddings(self, **kwargs) -> List[torch.Tensor]:      # t3.py:335-337
        cond = kwargs.get("conditionals")
        return [] if cond is None else [c for c in cond]

    def get_input_embeddings(self, input_ids: torch.Tensor, multimodal_embeddings=None, *,
                             speech_pos: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Decode branch of t3.py:429-486: ids are offset speech ids (>= 2500).  Returns bf16 [N, 2048]
        (cond‖uncond carry the same decode embedding, t3.py:480).  ``speech_pos`` is the explicit
        positional index the reference fails to track; None = the reference's
        literal fallback, index 0 for every token."""
        if multimodal_embeddings is not None and len(multimodal_embeddings) > 0:
            raise NotImplementedError("prefill batches go through T3B200Engine (prefill_embeds + forward_tokens)")
        ids = (input_ids.to(self.device) - SPEECH_TOKEN_OFFSET).to(torch.int32)
        if bool((ids < 0).any()):
            raise ValueError("get_input_embeddings: decode ids must be speech ids (>= SPEECH_TOKEN_OFFSET)")
        pos = torch.zeros_like(ids) if speech_pos is None else speech_pos.to(device=self.device, dtype=torch.int32)
        return self.embed("speech_emb", "speech_pos_emb", ids, pos, dup=2).view(ids.numel(), 2 * self.dims.hidden)

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
