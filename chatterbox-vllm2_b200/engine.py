"""T3B200Engine: the slice of ``vllm.LLM`` that ``ChatterboxTTS`` drives (tts.py:150-171 construction,
tts.py:445-465 ``self.t3.generate(prompts, sampling_params)``, tts.py:474-492 result shape), for one
B200.

Continuous batching over fixed request slots; paged KV (block 16) whose block tables GROW with the request
(a few blocks ahead of the decode position, so a 100-token and a 1000-token request cost what they use);
the 32 leading positions of a request's cond and uncond rows - bit-identical conditioning prefix,
t3.py:554-556 - live in ONE pair of physical blocks, shared across requests of the same voice (tts.py:256-288);
prefill of new requests as ONE multi-token pass (t3.py:542-561 layout, assembled on the device);
the decode step (embedding gather -> 30 layers -> head -> CFG -> sampler -> state advance) captured
in a CUDA graph per slot-count bucket and replayed with no host writes between steps.  The host only
extends block tables and reads back ``hist_len`` / ``active`` every few steps to retire finished requests.

Multi-GPU: replicas only - one engine per GPU, requests assigned by ``dispatch.assign_requests``;
no collective on the decode path.
"""
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import ctypes as C

import torch

from . import _lib
from .constants import CONDITIONING_SIZE, KV_BLOCK, SPEECH_TOKEN_OFFSET, T3Dims
from .model import (START_SPEECH_TOKEN, STOP_SPEECH_TOKEN, DecodeState, KVPool, SamplingTensors, T3B200Model)

GEMV_MMA_MAX_SEQS = 8            # engines of up to this many requests (<= 16 CFG rows) run the tensor-core GEMV chain by default
SHARED_PREFIX_BLOCKS = (CONDITIONING_SIZE // KV_BLOCK)          # 2 blocks = positions 0..31 of the 34-row conditioning prefix


@dataclass
class SamplingParams:
    """The vLLM SamplingParams fields the reference sets (tts.py:455-464) plus min_p / seed."""
    temperature: float = 0.8
    top_p: float = 1.0
    min_p: float = 0.0
    repetition_penalty: float = 1.0
    max_tokens: int = 1000
    stop_token_ids: Sequence[int] = (STOP_SPEECH_TOKEN + SPEECH_TOKEN_OFFSET,)
    seed: Optional[int] = None
    cfg_scale: Optional[float] = None      # per-request CFG (the reference has one global, t3.py:296)
    ignore_eos: bool = False               # fixed-work benchmarks: never stop early
    analyzer_text_tokens: int = 0          # > 0: run the token heuristics of alignment_stream_analyzer.py:143-195 in the loop with
                                           # text_tokens_count = this (tts.py:496 passes len(prompt.split()) * 2); 0 = off

    def validate(self) -> None:
        """vLLM rejects the same values (sampling_params.py _verify_args)."""
        if not (0.0 < self.top_p <= 1.0):
            raise ValueError(f"top_p must be in (0, 1], got {self.top_p}")
        if not (0.0 <= self.min_p <= 1.0):
            raise ValueError(f"min_p must be in [0, 1], got {self.min_p}")
        if self.temperature < 0.0:
            raise ValueError(f"temperature must be non-negative, got {self.temperature}")
        if self.repetition_penalty <= 0.0:
            raise ValueError(f"repetition_penalty must be positive, got {self.repetition_penalty}")
        if self.max_tokens < 1:
            raise ValueError(f"max_tokens must be at least 1, got {self.max_tokens}")
        if len(self.stop_token_ids) > 1:
            raise ValueError("one stop token per request is supported (tts.py:457 passes exactly one)")

    def stop_token(self) -> int:
        """Un-offset stop id, -1 = never stops."""
        if self.ignore_eos or len(self.stop_token_ids) == 0:
            return -1
        return int(self.stop_token_ids[0]) - SPEECH_TOKEN_OFFSET


@dataclass
class CompletionOutput:
    token_ids: List[int]                   # offset ids (>= 2500), what tts.py:474-492 consumes
    finish_reason: str = "length"


@dataclass
class RequestOutput:
    request_id: int
    outputs: List[CompletionOutput] = field(default_factory=list)


@dataclass
class _Req:
    rid: int
    cond_emb: Optional[torch.Tensor]       # [34, D] conditioning prefix (None when ready-made embeddings were given)
    text_ids: Optional[List[int]]
    embeds: Optional[Tuple[torch.Tensor, torch.Tensor]]   # ready (cond, uncond) prefill rows
    sp: SamplingParams
    n_prompt: int
    slot: int = -1
    blocks: List[List[int]] = field(default_factory=lambda: [[], []])     # physical blocks of the cond / uncond row, in order
    n_gen: int = 0                         # tokens generated so far (host's view)
    voice: Optional[int] = None            # key of the shared conditioning prefix


def shard_requests(n_requests: int, world_size: int, rank: int) -> List[int]:
    """Replica sharding: request i runs on rank i % world_size (independent units, no exchange)."""
    return list(range(rank, n_requests, world_size))


class T3B200Engine:
    def __init__(self, model: T3B200Model, max_num_seqs: int = 32, max_model_len: int = 1200,
                 kv_blocks: Optional[int] = None, check_every: int = 8, prefill_chunk_tokens: int = 8192,
                 use_graphs: bool = True, kv_dtype: str = "bf16", share_prefix: bool = True, grow_steps: int = 32,
                 admit_batch: Optional[int] = None, admit_wait_steps: int = 16, norm_fusion: Optional[bool] = None,
                 gemv_mma: Optional[bool] = None):
        self.m = model
        self.dims: T3Dims = model.dims
        self.dev = model.device
        self.max_req = max_num_seqs
        self.max_len = max_model_len
        self.max_blocks = (max_model_len + KV_BLOCK - 1) // KV_BLOCK
        if kv_blocks is None:
            kv_blocks = 2 * max_num_seqs * self.max_blocks + 1
        self.kv = KVPool(self.dims, kv_blocks, self.dev, kv_dtype)
        self.state = DecodeState(max_num_seqs, self.max_blocks, max_model_len, max_model_len, self.dev)
        self.sampling = SamplingTensors(max_num_seqs, self.dev, model.combine)
        self.hidden = torch.empty(2 * max_num_seqs, self.dims.hidden, dtype=torch.bfloat16, device=self.dev)
        self.out_tokens = torch.empty(max_num_seqs, dtype=torch.int32, device=self.dev)
        self.check_every = max(1, check_every)
        self.grow_steps = max(1, grow_steps)
        # Admission batching: a prefill pass costs milliseconds whether it carries 1 prompt or 16, so with a queue of waiting
        # requests freed slots are refilled in groups - when ``admit_batch`` slots are free (default max_num_seqs / 32) or
        # ``admit_wait_steps`` steps have passed since a slot freed up - instead of one prefill per finished request.
        self.admit_batch = max(1, max_num_seqs // 32) if admit_batch is None else max(1, admit_batch)
        self.admit_wait_steps = max(0, admit_wait_steps)
        self.prefill_chunk = max(prefill_chunk_tokens, 2 * max_model_len)
        self.use_graphs = use_graphs
        self.share_prefix = share_prefix
        self._graphs: Dict[int, torch.cuda.CUDAGraph] = {}
        self.slots: List[Optional[_Req]] = [None] * max_num_seqs
        self.steps_run = 0
        self.preemptions = 0
        self._voices: Dict[int, List] = {}           # voice key -> [blocks (2), live requests]
        # Launch shapes that change the fp32 summation order are fixed per ENGINE, not per step: a batch engine runs the tcgen05
        # projections at every live-row count (also when only one request is left), a single-request engine the GEMV - so a
        # request's tokens do not depend on how many other requests are running beside it.
        if model.options is not None:
            self.opt = model.options
        else:
            items = 2 * max_num_seqs * self.dims.n_heads        # (row, head) items of a full step
            team = 8
            while team > 1 and items * team > 148 * 8:
                team //= 2
            # Same rule for the norms folded into the projections (t3b_options_t.norm_fusion; pays at 16..32 rows): an engine whose
            # largest step is in that range uses them at every step, any other engine never - ``norm_fusion=True / False`` overrides
            # (engines of different sizes produce the same tokens only if they agree on it).
            if norm_fusion is None:
                norm_fusion = 16 <= 2 * max_num_seqs <= 32
            # Engines of up to 8 requests (<= 16 CFG rows, bf16): the tensor-core GEMV chain (t3b_options_t.gemv_mma, 5 launches per
            # layer) at every live-row count, on fragment-ordered copies of the weights made here once (+1 GB), every launch pulling
            # the next layer's copy of its projection into L2 (1 request 0.57 -> 0.44 ms/step, 2 requests 0.87 -> 0.48, 4 requests
            # 0.91 -> 0.56, 8 requests 0.94 -> 0.80; profiles/r2_notes.md r5).  ``gemv_mma=False`` keeps the SIMT GEMV (1 request) / tcgen05 (2..8) chain.
            if gemv_mma is None:
                gemv_mma = max_num_seqs <= GEMV_MMA_MAX_SEQS
            gemv_mma = bool(gemv_mma) and 2 * max_num_seqs <= 16 and model.enable_gemv_mma()
            self.opt = _lib.make_options(gemv_max_rows=16 if gemv_mma else (0 if max_num_seqs == 1 else -1), attn_team_warps=team,
                                         norm_fusion=1 if norm_fusion else -1, gemv_mma=1 if gemv_mma else -1,
                                         weight_prefetch=2 if gemv_mma else 0)
        model.bind(self.kv, self.state)
        # One workspace for the engine's lifetime, sized for its largest launch: captured graphs hold its address, so it must
        # never move (the model's own growable scratch is not used here).
        self.ws = torch.empty(model.workspace_bytes(max(2 * max_num_seqs, self.prefill_chunk)), dtype=torch.uint8, device=self.dev)

    # ------------------------------------------------------------------ one decode step
    def _step_eager(self, n_slots: int) -> None:
        self.m.decode_and_sample(self.kv, self.state, self.sampling, n_slots, self.hidden, self.out_tokens, self.ws, opt=self.opt)

    def launches_per_step(self, n_slots: int) -> int:
        """Kernels one decode step launches: embed + per layer [norm, qkv, attention(+RoPE+KV write), o, norm,
        gate_up(+SiLU*mul), down] + final norm + head + sampler (the two norms ride inside the GEMVs at <= 2 rows and inside the
        tcgen05 projections at 3..32 rows)."""
        rows = 2 * self._bucket(n_slots)
        gmax = 2 if self.opt.gemv_max_rows == 0 else max(0, min(16 if self.opt.gemv_mma > 0 else 4, self.opt.gemv_max_rows))
        gemv = rows <= gmax and self.m.quantization is None and not self.opt.gemv_no_norm
        nf = self.opt.norm_fusion
        if not gemv and rows <= 32 and (nf > 0 or (nf == 0 and rows >= 9)) and rows > gmax and self.dims.hidden == 1024:
            # the norms ride inside the tcgen05 projections (csrc/api.cu forward_impl): layer 0's first norm is the only one left
            return 1 + 1 + self.dims.n_layers * 5 + 3
        return 1 + self.dims.n_layers * (5 if gemv else 7) + 3

    def _bucket(self, n_slots: int) -> int:
        b = 1
        while b < n_slots:
            b *= 2
        return min(b, self.max_req)

    def step(self, n_slots: int) -> None:
        """Advance every active request in slots [0, n_slots) by one token (device state only)."""
        nb = self._bucket(n_slots)
        if not self.use_graphs:
            self._step_eager(nb)
        else:
            g = self._graphs.get(nb)
            if g is None:
                g = self._capture(nb)
            g.replay()
        self.steps_run += 1

    def _capture(self, nb: int) -> torch.cuda.CUDAGraph:
        # Warm-up + capture run the step for real; park every slot as inactive so no state advances
        # (idle rows write their K/V into scratch block 0).
        st = self.state
        saved = (st.active.clone(), st.ctx_lens.clone(), st.rope_pos.clone(), st.block_tables.clone())
        st.active.zero_(); st.ctx_lens.zero_(); st.rope_pos.zero_(); st.block_tables.zero_()
        s = torch.cuda.Stream(device=self.dev)
        s.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(s):
            self._step_eager(nb)
        torch.cuda.current_stream(self.dev).wait_stream(s)
        torch.cuda.synchronize(self.dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            self._step_eager(nb)
        torch.cuda.synchronize(self.dev)
        st.active.copy_(saved[0]); st.ctx_lens.copy_(saved[1]); st.rope_pos.copy_(saved[2]); st.block_tables.copy_(saved[3])
        self._graphs[nb] = g
        return g

    # ------------------------------------------------------------------ KV blocks
    @staticmethod
    def _nblk(tokens: int) -> int:
        return (tokens + KV_BLOCK - 1) // KV_BLOCK

    def _voice_key(self, r: _Req) -> Optional[int]:
        if not self.share_prefix or r.cond_emb is None or r.n_prompt < SHARED_PREFIX_BLOCKS * KV_BLOCK:
            return None
        c = r.cond_emb[:SHARED_PREFIX_BLOCKS * KV_BLOCK].detach().to("cpu").contiguous()
        return hash((str(c.dtype), c.view(torch.uint8).numpy().tobytes()))

    def _alloc_prompt(self, r: _Req) -> bool:
        """Blocks for the prompt plus the growth look-ahead; False (nothing allocated) if the pool cannot hold them."""
        nb = self._nblk(min(self.max_len, r.n_prompt + self.grow_steps))
        key = self._voice_key(r)
        if key is None:
            if 2 * nb > len(self.kv.free):
                return False
            r.blocks = [self.kv.alloc(nb), self.kv.alloc(nb)]
            return True
        v = self._voices.get(key)
        own = 2 * (nb - SHARED_PREFIX_BLOCKS)
        if own + (0 if v else SHARED_PREFIX_BLOCKS) > len(self.kv.free):
            return False
        if v is None:
            shared = self.kv.alloc(SHARED_PREFIX_BLOCKS)
            self.kv.share(shared)                              # the cond and the uncond row both name them
            v = self._voices[key] = [shared, 0]
        else:
            shared = v[0]
            self.kv.share(shared); self.kv.share(shared)
        v[1] += 1
        r.voice = key
        r.blocks = [list(shared) + self.kv.alloc(nb - SHARED_PREFIX_BLOCKS), list(shared) + self.kv.alloc(nb - SHARED_PREFIX_BLOCKS)]
        return True

    def _release(self, r: _Req) -> None:
        self.kv.release(r.blocks[0]); self.kv.release(r.blocks[1])
        r.blocks = [[], []]
        if r.voice is not None:
            v = self._voices[r.voice]
            v[1] -= 1
            if v[1] == 0:
                del self._voices[r.voice]
            r.voice = None

    def _grow(self, live: List[_Req], steps: int) -> bool:
        """Every live row gets the blocks its next ``steps`` tokens need; one small host->device copy.  False if the pool is short."""
        want = []
        total = 0
        for r in live:
            nb = self._nblk(min(self.max_len, r.n_prompt + r.n_gen + steps))
            add = nb - len(r.blocks[0])
            if add > 0:
                want.append((r, add))
                total += 2 * add
        if total > len(self.kv.free):
            return False
        if not want:
            return True
        idx, val = [], []
        for r, add in want:
            for half in (0, 1):
                new = self.kv.alloc(add)
                base = (2 * r.slot + half) * self.max_blocks + len(r.blocks[half])
                idx += list(range(base, base + add))
                val += new
                r.blocks[half] += new
        upd = torch.tensor([idx, val], dtype=torch.int64).pin_memory().to(self.dev, non_blocking=True)
        self.state.block_tables.view(-1)[upd[0]] = upd[1].to(torch.int32)
        return True

    def _reorder(self) -> None:
        """Deal order of the attention (t3b_state_t.row_order): rows of the current slot bucket by context, longest first, idle rows
        last.  Every live row grows by one token per step, so the order only changes when requests come or go - recomputed on the
        host then, one small asynchronous copy."""
        live = [r for r in self.slots if r is not None]
        if not live:
            return
        nb = self._bucket(max(r.slot for r in live) + 1)
        by_ctx = sorted(live, key=lambda r: -(r.n_prompt + r.n_gen))
        rows = [2 * r.slot + h for r in by_ctx for h in (0, 1)]
        used = {r.slot for r in live}
        rows += [2 * sl + h for sl in range(nb) if sl not in used for h in (0, 1)]
        self.state.row_order[:2 * nb].copy_(torch.tensor(rows, dtype=torch.int32).pin_memory(), non_blocking=True)

    def kv_tokens_outstanding(self) -> int:
        """Token-rows of KV currently reserved (the dispatcher's load measure)."""
        return (self.kv.num_blocks - 1 - len(self.kv.free)) * KV_BLOCK

    # ------------------------------------------------------------------ admission + prefill
    def _make_req(self, rid: int, prompt: dict, sp: SamplingParams) -> _Req:
        sp.validate()
        if "prefill_embeds" in prompt:
            cond, uncond = prompt["prefill_embeds"]
            cond = cond.to(device=self.dev, dtype=torch.bfloat16)
            uncond = uncond.to(device=self.dev, dtype=torch.bfloat16)
            r = _Req(rid, None, None, (cond, uncond), sp, cond.shape[0])
        else:
            conds = prompt["multi_modal_data"]["conditionals"]
            cond_emb = conds[0] if isinstance(conds, (list, tuple)) else conds
            ids = prompt.get("prompt_token_ids")
            if ids is None:
                raise ValueError("prompt needs 'prompt_token_ids' (text tokenisation is outside the T3 decode path)")
            ids = list(ids)
            if ids and ids[0] == 695 and ids[-1] == 697 and len(ids) > CONDITIONING_SIZE:     # packed like t3.py:189-200
                from .packing import unpack_prompt
                ids = unpack_prompt(ids)
            if cond_emb.shape[0] < CONDITIONING_SIZE:
                raise ValueError(f"conditionals must have {CONDITIONING_SIZE} rows (t3.py:210), got {cond_emb.shape[0]}")
            r = _Req(rid, cond_emb[:CONDITIONING_SIZE], ids, None, sp, CONDITIONING_SIZE + len(ids) + 1)
        if r.n_prompt + 1 > self.max_len:
            raise ValueError(f"prompt of {r.n_prompt} tokens does not fit max_model_len {self.max_len}")
        return r

    def _limit(self, r: _Req) -> int:
        return min(r.sp.max_tokens, self.max_len - r.n_prompt)

    def _admit(self, reqs: List[_Req]) -> None:
        """Run the prompts of ``reqs`` (blocks already allocated, slots assigned) through the backbone in one multi-token pass
        per chunk, sample each request's first speech token and install its slot state."""
        st, dev, m = self.state, self.dev, self.m
        bt_rows, bt_host = [], []
        for r in reqs:
            for half in (0, 1):
                bt_host.append(r.blocks[half] + [0] * (self.max_blocks - len(r.blocks[half])))
            bt_rows += [2 * r.slot, 2 * r.slot + 1]
        st.block_tables[torch.tensor(bt_rows, device=dev)] = torch.tensor(bt_host, dtype=torch.int32).to(dev)
        chunks: List[List[_Req]] = [[]]
        ntok = 0
        for r in reqs:
            if ntok + 2 * r.n_prompt > self.prefill_chunk and chunks[-1]:
                chunks.append([]); ntok = 0
            chunks[-1].append(r); ntok += 2 * r.n_prompt
        for ch in chunks:
            if all(r.embeds is None for r in ch):
                x, tok_row, tok_idx, starts = m.prefill_assemble([r.cond_emb for r in ch], [r.text_ids for r in ch], [r.slot for r in ch])
                seq_start, seq_row = m.last_seq
            else:                                            # ready-made embeddings: rows built on the host side
                xs, rows, idx, starts, srow = [], [], [], [0], []
                for r in ch:
                    pair = r.embeds if r.embeds is not None else m.prefill_embeds(r.cond_emb, r.text_ids)
                    for half, xx in enumerate(pair):
                        xs.append(xx)
                        rows.append(torch.full((r.n_prompt,), 2 * r.slot + half, dtype=torch.int32))
                        idx.append(torch.arange(r.n_prompt, dtype=torch.int32))
                        starts.append(starts[-1] + r.n_prompt)
                        srow.append(2 * r.slot + half)
                x = torch.cat(xs, 0).contiguous()
                tok_row, tok_idx = torch.cat(rows).to(dev), torch.cat(idx).to(dev)
                seq_start = torch.tensor(starts, dtype=torch.int32, device=dev)
                seq_row = torch.tensor(srow, dtype=torch.int32, device=dev)
            max_seq = max(r.n_prompt for r in ch)
            h = m.forward_tokens(self.kv, x, tok_row, tok_idx, tok_idx, st.block_tables, seq_start, seq_row, max_seq, ws=self.ws, opt=self.opt)
            h_last = h[torch.tensor([s - 1 for s in starts[1:]], device=dev)].contiguous()      # hidden of each row's BOS token
            nreq = len(ch)
            tmp_state = DecodeState(nreq, self.max_blocks, self.max_len, self.max_len, dev)
            tmp_sp = SamplingTensors(nreq, dev, m.combine)
            tmp_sp.set_many([dict(cfg_scale=m.cfg_scale if r.sp.cfg_scale is None else r.sp.cfg_scale,
                                  repetition_penalty=r.sp.repetition_penalty, temperature=r.sp.temperature, min_p=r.sp.min_p,
                                  top_p=r.sp.top_p, seed=(r.rid if r.sp.seed is None else r.sp.seed),
                                  analyzer_text_tokens=r.sp.analyzer_text_tokens) for r in ch])
            tmp_state.stop_tokens.copy_(torch.tensor([r.sp.stop_token() for r in ch], dtype=torch.int32))
            tmp_state.active.fill_(1)
            toks = torch.empty(nreq, dtype=torch.int32, device=dev)
            m.head_cfg_sample(h_last, 2 * nreq, tmp_sp, tmp_state, toks, advance=True, ws=self.ws, opt=self.opt)
            slots = torch.tensor([r.slot for r in ch], device=dev)
            rows2 = torch.stack([2 * slots, 2 * slots + 1], 1).reshape(-1)
            npr = torch.tensor([r.n_prompt for r in ch], dtype=torch.int32, device=dev).repeat_interleave(2)
            st.tokens[slots] = tmp_state.tokens
            st.speech_pos[slots] = tmp_state.speech_pos              # 1: BOS sat at 0 (t3.py:551)
            st.history[slots] = tmp_state.history
            st.hist_len[slots] = tmp_state.hist_len
            st.step[slots] = tmp_state.step
            st.active[slots] = tmp_state.active
            st.stop_tokens[slots] = tmp_state.stop_tokens
            st.ctx_lens[rows2] = npr
            st.rope_pos[rows2] = npr
            for name in SamplingTensors.FIELDS:
                getattr(self.sampling, name)[slots] = getattr(tmp_sp, name)
            for r in ch:
                r.n_gen = 1

    def install(self, prompts: Sequence[dict], sampling_params, reserve_steps: int = 0) -> List[_Req]:
        """Admit ``prompts`` into free slots (prefill + first token) WITHOUT running decode steps, with blocks for ``reserve_steps``
        more tokens per row: the entry point of benchmarks and teacher-forced tests that drive ``step()`` themselves."""
        sps = list(sampling_params) if isinstance(sampling_params, (list, tuple)) else [sampling_params] * len(prompts)
        reqs = [self._make_req(i, p, sp) for i, (p, sp) in enumerate(zip(prompts, sps))]
        free = [i for i, s in enumerate(self.slots) if s is None]
        if len(free) < len(reqs):
            raise ValueError(f"{len(reqs)} requests but {len(free)} free slots")
        for r, slot in zip(reqs, free):
            r.slot = slot
            if not self._alloc_prompt(r):
                raise MemoryError("KV pool exhausted")
            self.slots[slot] = r
        self._admit(reqs)
        if reserve_steps > 0 and not self._grow(reqs, reserve_steps):
            raise MemoryError("KV pool exhausted")
        self._reorder()
        return reqs

    def uninstall(self, reqs: Sequence[_Req]) -> None:
        for r in reqs:
            self._release(r)
            self.slots[r.slot] = None
        self._park([r.slot for r in reqs])

    # ------------------------------------------------------------------ the caller's entry point
    def _retire(self, done: List[_Req], pending: List[Tuple[List[Tuple[int, int, int]], torch.Tensor]]) -> None:
        """Finished requests leave their slots at once; their history rows start an ASYNCHRONOUS copy into pinned host memory
        (stream-ordered in front of whatever reuses the slots) and are turned into results by ``_collect`` when generate()
        returns - the host never waits for the GPU inside the decode loop, so the next steps are queued while this one runs."""
        rows = torch.tensor([r.slot for r in done], device=self.dev)
        host = torch.empty((len(done), self.state.hist_stride), dtype=torch.int32, pin_memory=True)
        host.copy_(self.state.history[rows], non_blocking=True)
        pending.append(([(r.rid, r.n_gen, r.sp.stop_token()) for r in done], host))
        for r in done:
            self._release(r)
            self.slots[r.slot] = None
        self._park([r.slot for r in done])

    @staticmethod
    def _collect(pending, results: Dict[int, RequestOutput]) -> None:
        for metas, host in pending:
            for i, (rid, n_gen, stop) in enumerate(metas):
                toks = host[i, :n_gen].tolist()
                stopped = stop >= 0 and stop in toks
                if stopped:
                    toks = toks[:toks.index(stop) + 1]
                results[rid] = RequestOutput(rid, [CompletionOutput([t + SPEECH_TOKEN_OFFSET for t in toks], "stop" if stopped else "length")])

    def _park(self, slots: List[int]) -> None:
        """Freed slots become idle rows (they write into scratch block 0): four indexed stores in all."""
        if not slots:
            return
        st = self.state
        sl = torch.tensor(slots, device=self.dev)
        rw = torch.stack([2 * sl, 2 * sl + 1], 1).reshape(-1)
        st.active[sl] = 0
        st.ctx_lens[rw] = 0
        st.rope_pos[rw] = 0
        st.block_tables[rw] = 0

    def generate(self, prompts: Sequence[dict], sampling_params=None) -> List[RequestOutput]:
        """tts.py:445-465.  ``prompts``: dicts with ``prompt_token_ids`` (plain text ids, or packed like t3.py:189-200) +
        ``multi_modal_data.conditionals`` ([34, 1024] tensor, tts.py:452) or ready ``prefill_embeds`` (cond, uncond).
        ``sampling_params``: one SamplingParams or one per prompt (stop token, ignore_eos and every knob are per request).
        Returns results in prompt order; ``outputs[0].token_ids`` are offset by SPEECH_TOKEN_OFFSET like the reference's
        (tts.py:474-492)."""
        if sampling_params is None:
            sampling_params = SamplingParams()
        sps = list(sampling_params) if isinstance(sampling_params, (list, tuple)) else [sampling_params] * len(prompts)
        if len(sps) != len(prompts):
            raise ValueError(f"{len(prompts)} prompts but {len(sps)} sampling params")
        waiting = [self._make_req(i, p, sp) for i, (p, sp) in enumerate(zip(prompts, sps))]
        waiting.reverse()
        results: Dict[int, RequestOutput] = {}
        pending: List[Tuple[List[Tuple[int, int, int]], torch.Tensor]] = []
        st = self.state
        waited = 0                                              # steps run since a slot became free with requests still waiting
        changed = True                                          # the live set changed: the attention's deal order is stale
        while waiting or any(s is not None for s in self.slots):
            new: List[_Req] = []
            n_free = sum(s is None for s in self.slots)
            admit_now = n_free == self.max_req or n_free >= min(self.admit_batch, len(waiting)) or waited >= self.admit_wait_steps
            for slot in range(self.max_req):
                if not waiting or not admit_now:
                    break
                if self.slots[slot] is None:
                    r = waiting[-1]
                    r.slot = slot
                    if not self._alloc_prompt(r):
                        r.slot = -1
                        break                                   # pool is full: wait for a running request to finish
                    waiting.pop()
                    self.slots[slot] = r
                    new.append(r)
            if new:
                self._admit(new)
                waited = 0
                changed = True
            live = [r for r in self.slots if r is not None]
            if not live:
                raise MemoryError(f"KV pool ({self.kv.num_blocks} blocks) cannot hold a single prompt of {waiting[-1].n_prompt} tokens")
            n_slots = max(r.slot for r in live) + 1
            may_stop = any(r.sp.stop_token() >= 0 or r.sp.analyzer_text_tokens > 0 for r in live)
            # steps until the host has to look again: the shortest remaining length, the block look-ahead, and - only when a
            # stop token can end a request early - at most check_every
            budget = min(min(self._limit(r) - r.n_gen for r in live), self.grow_steps)
            if may_stop:
                budget = min(budget, self.check_every)
            if waiting and len(live) < self.max_req:
                budget = min(budget, max(1, self.admit_wait_steps - waited))     # come back in time to refill the free slots
            budget = max(0, budget)
            while not self._grow(live, budget):
                # pool exhausted: preempt the youngest request (its tokens are regenerated later - sampling is keyed by
                # (seed, index, step), so the stream is the same)
                if len(live) == 1:
                    raise MemoryError(f"KV pool ({self.kv.num_blocks} blocks) cannot hold one request of {live[0].n_prompt + live[0].n_gen} tokens")
                victim = max(live, key=lambda r: r.rid)
                live.remove(victim)
                self._release(victim)
                self.slots[victim.slot] = None
                self._park([victim.slot])
                victim.slot, victim.n_gen = -1, 0
                waiting.append(victim)
                self.preemptions += 1
                changed = True
                n_slots = max(r.slot for r in live) + 1
            if changed:
                self._reorder()
                changed = False
            for _ in range(budget):
                self.step(n_slots)
            if waiting and len(live) < self.max_req:
                waited += budget
            if may_stop:
                poll = torch.stack([st.hist_len[:n_slots], st.active[:n_slots].to(st.hist_len.dtype)]).cpu()    # one D2H per poll
                hist_len, active = poll[0], poll[1]
                for r in live:
                    r.n_gen = int(hist_len[r.slot])
                done = [r for r in live if not bool(active[r.slot]) or r.n_gen >= self._limit(r)]
            else:
                for r in live:                                  # nothing can end early: the host knows every length
                    r.n_gen += budget
                done = [r for r in live if r.n_gen >= self._limit(r)]
            if done:
                self._retire(done, pending)
                changed = True
        torch.cuda.current_stream(self.dev).synchronize()      # the one wait of the call: every queued step and result copy
        self._collect(pending, results)
        return [results[i] for i in range(len(prompts))]
