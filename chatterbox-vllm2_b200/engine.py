"""T3B200Engine: the slice of ``vllm.LLM`` that ``ChatterboxTTS`` drives (tts.py:150-171 construction,
tts.py:445-465 ``self.t3.generate(prompts, sampling_params)``, tts.py:474-492 result shape), for one
B200.

Continuous batching over fixed request slots; paged KV (block 16) with all blocks of a request
reserved at admission; prefill of new requests as ONE multi-token pass (t3.py:542-561 layout);
the decode step (embedding gather -> 30 layers -> head -> CFG -> sampler -> state advance) captured
in a CUDA graph per slot-count bucket and replayed with no host writes between steps.  The host
only reads back ``hist_len`` / ``active`` every few steps to retire finished requests.

Multi-GPU: replicas only - one engine per GPU, requests sharded by the caller
(``shard_requests``); no collective on the decode path.
"""
import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import torch

from . import _lib
from .constants import CONDITIONING_SIZE, KV_BLOCK, SPEECH_TOKEN_OFFSET, T3Dims
from .model import (START_SPEECH_TOKEN, STOP_SPEECH_TOKEN, DecodeState, KVPool, SamplingTensors, T3B200Model)


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
    cond: torch.Tensor                     # [P, D] bf16 prefill embeddings, cond row
    uncond: torch.Tensor
    sp: SamplingParams
    slot: int = -1
    blocks: List[int] = field(default_factory=list)
    n_prompt: int = 0
    n_gen: int = 0


def shard_requests(n_requests: int, world_size: int, rank: int) -> List[int]:
    """Replica sharding: request i runs on rank i % world_size (independent units, no exchange)."""
    return list(range(rank, n_requests, world_size))


class T3B200Engine:
    def __init__(self, model: T3B200Model, max_num_seqs: int = 32, max_model_len: int = 1200,
                 kv_blocks: Optional[int] = None, check_every: int = 8, prefill_chunk_tokens: int = 8192,
                 use_graphs: bool = True, decode_lanes: Optional[int] = None):
        self.m = model
        self.dims: T3Dims = model.dims
        self.dev = model.device
        self.max_req = max_num_seqs
        self.max_len = max_model_len
        self.max_blocks = (max_model_len + KV_BLOCK - 1) // KV_BLOCK
        if kv_blocks is None:
            kv_blocks = 2 * max_num_seqs * self.max_blocks + 1
        self.kv = KVPool(self.dims, kv_blocks, self.dev)
        self.state = DecodeState(max_num_seqs, self.max_blocks, max_model_len, max_model_len, self.dev)
        self.sampling = SamplingTensors(max_num_seqs, self.dev, model.combine)
        self.hidden = torch.empty(2 * max_num_seqs, self.dims.hidden, dtype=torch.bfloat16, device=self.dev)
        self.out_tokens = torch.empty(max_num_seqs, dtype=torch.int32, device=self.dev)
        self.check_every = max(1, check_every)
        self.prefill_chunk = prefill_chunk_tokens
        self.use_graphs = use_graphs
        # Decode lanes: the step of a slot bucket is cut into `lanes` independent slices of requests, each a
        # chain of ~214 small dependent kernels on its own stream (forked / joined inside the CUDA graph).
        # One chain is latency-bound (a few us per kernel, far below HBM speed); several chains in flight
        # fill the machine.  Rows never interact, so every request's arithmetic is unchanged.
        self.decode_lanes = decode_lanes if decode_lanes is not None else (int(os.environ.get("T3B_DECODE_LANES", "0")) or None)
        self._lane_streams: List[torch.cuda.Stream] = []
        self._graphs: Dict[int, torch.cuda.CUDAGraph] = {}
        self.slots: List[Optional[_Req]] = [None] * max_num_seqs
        self.steps_run = 0
        model.bind(self.kv, self.state)
        model.workspace(max(2 * max_num_seqs, prefill_chunk_tokens))     # sized once: graphs hold its address
        if model.mega is not None:
            r = min(2 * max_num_seqs, _lib.lib().t3b_mega_max_rows())
            model.mega_workspace(r)
            model.mega_logits(r)
        if model.fused is not None:
            r = min(2 * max_num_seqs, _lib.lib().t3b_fused_max_rows())
            model.fused_workspace(r)
            model.mega_logits(r)

    # ------------------------------------------------------------------ one decode step
    def lane_plan(self, n_slots: int) -> List[tuple]:
        """[(first slot, slots)] per lane.  Default one lane: measured on B200 (cfg2, 32 requests) 1/2/4/8 lanes give
        19.0k / 20.0k / 19.5k / 13.3k tokens/s - the lanes contend for SMs more than they hide latency."""
        lanes = self.decode_lanes if self.decode_lanes else 1
        lanes = max(1, min(lanes, n_slots))
        base, extra = divmod(n_slots, lanes)
        plan, s0 = [], 0
        for g in range(lanes):
            n = base + (1 if g < extra else 0)
            plan.append((s0, n))
            s0 += n
        return plan

    def _step_eager(self, n_slots: int) -> None:
        plan = self.lane_plan(n_slots)
        if len(plan) == 1:
            self.m.decode_lane(self.kv, self.state, self.sampling, 0, n_slots, 0, self.hidden, self.out_tokens)
            return
        while len(self._lane_streams) < len(plan):
            self._lane_streams.append(torch.cuda.Stream(device=self.dev))
        cur = torch.cuda.current_stream(self.dev)
        for g, (s0, n) in enumerate(plan):
            st = self._lane_streams[g]
            st.wait_stream(cur)                                   # fork
            with torch.cuda.stream(st):
                self.m.decode_lane(self.kv, self.state, self.sampling, s0, n, g, self.hidden, self.out_tokens)
        for g in range(len(plan)):
            cur.wait_stream(self._lane_streams[g])                # join

    def launches_per_step(self, n_slots: int) -> int:
        """Kernels one decode step launches: embed + per layer [norm, qkv, attention(+RoPE+KV write), o, norm,
        gate_up(+SiLU*mul), down] + final norm + head + sampler."""
        return len(self.lane_plan(self._bucket(n_slots))) * (1 + self.dims.n_layers * 7 + 3)

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

    # ------------------------------------------------------------------ admission + prefill
    def _make_req(self, rid: int, prompt: dict, sp: SamplingParams) -> _Req:
        if "prefill_embeds" in prompt:
            cond, uncond = prompt["prefill_embeds"]
            cond = cond.to(device=self.dev, dtype=torch.bfloat16)
            uncond = uncond.to(device=self.dev, dtype=torch.bfloat16)
        else:
            conds = prompt["multi_modal_data"]["conditionals"]
            cond_emb = conds[0] if isinstance(conds, (list, tuple)) else conds
            ids = prompt.get("prompt_token_ids")
            if ids is None:
                raise ValueError("prompt needs 'prompt_token_ids' (text tokenisation is outside the T3 decode path)")
            cond, uncond = self.m.prefill_embeds(cond_emb[:CONDITIONING_SIZE], ids)
        r = _Req(rid, cond, uncond, sp)
        r.n_prompt = cond.shape[0]
        if r.n_prompt + 1 > self.max_len:
            raise ValueError(f"prompt of {r.n_prompt} tokens does not fit max_model_len {self.max_len}")
        return r

    def _embed_prompts(self, prompts: List[dict]) -> List[dict]:
        """Prompts that come as (conditionals, prompt_token_ids) get their prefill embeddings in ONE batched pass
        (T3B200Model.prefill_embeds_batch) instead of a dozen small device operations each."""
        todo = [i for i, p in enumerate(prompts) if "prefill_embeds" not in p and p.get("prompt_token_ids") is not None
                and "multi_modal_data" in p]
        if len(todo) < 2:
            return prompts
        conds = []
        for i in todo:
            c = prompts[i]["multi_modal_data"]["conditionals"]
            conds.append(c[0] if isinstance(c, (list, tuple)) else c)
        if any(c.shape[0] < CONDITIONING_SIZE for c in conds) or len({c.device for c in conds}) != 1:
            return prompts
        pairs = self.m.prefill_embeds_batch(conds, [prompts[i]["prompt_token_ids"] for i in todo])
        out = list(prompts)
        for i, pr in zip(todo, pairs):
            out[i] = {"prefill_embeds": pr}
        return out

    def _admit(self, reqs: List[_Req]) -> None:
        """Reserve KV, run the prompts of ``reqs`` through the backbone in one multi-token pass per chunk,
        sample each request's first speech token and install its slot state."""
        st, dev = self.state, self.dev
        # block tables of all admitted requests: built on the host, ONE copy + ONE indexed store on the device
        bt = torch.zeros(2 * len(reqs), self.max_blocks, dtype=torch.int32)
        bt_rows = []
        for i, r in enumerate(reqs):
            total = min(self.max_len, r.n_prompt + r.sp.max_tokens)
            nblk = (total + KV_BLOCK - 1) // KV_BLOCK
            r.blocks = self.kv.alloc(2 * nblk)
            bt[2 * i, :nblk] = torch.tensor(r.blocks[:nblk], dtype=torch.int32)
            bt[2 * i + 1, :nblk] = torch.tensor(r.blocks[nblk:], dtype=torch.int32)
            bt_rows += [2 * r.slot, 2 * r.slot + 1]
        st.block_tables[torch.tensor(bt_rows, device=dev)] = bt.to(dev)
        # token arrays for the multi-token pass: rows in slot order, tokens of a row contiguous
        chunks: List[List[_Req]] = [[]]
        ntok = 0
        for r in reqs:
            if ntok + 2 * r.n_prompt > self.prefill_chunk and chunks[-1]:
                chunks.append([]); ntok = 0
            chunks[-1].append(r); ntok += 2 * r.n_prompt
        for ch in chunks:
            xs, rows, idx, last = [], [], [], []
            n = 0
            for r in ch:
                for half, x in enumerate((r.cond, r.uncond)):
                    xs.append(x)
                    rows.append(torch.full((r.n_prompt,), 2 * r.slot + half, dtype=torch.int32))
                    idx.append(torch.arange(r.n_prompt, dtype=torch.int32))
                    n += r.n_prompt
                    last.append(n - 1)
            x = torch.cat(xs, 0).contiguous()
            if x.shape[0] > self.prefill_chunk:
                self.m.workspace(x.shape[0]); self._graphs.clear()
            tok_row, tok_idx = torch.cat(rows).to(dev), torch.cat(idx).to(dev)
            h = self.m.forward_tokens(self.kv, x, tok_row, tok_idx, tok_idx, st.block_tables)
            h_last = h[torch.tensor(last, device=dev)].contiguous()          # hidden of each row's BOS token
            nreq = len(ch)
            tmp_state = DecodeState(nreq, self.max_blocks, self.max_len, self.max_len, dev)
            tmp_sp = SamplingTensors(nreq, dev, self.m.combine)
            for j, r in enumerate(ch):
                sp = r.sp
                tmp_sp.set(j, cfg_scale=self.m.cfg_scale if sp.cfg_scale is None else sp.cfg_scale,
                           repetition_penalty=sp.repetition_penalty, temperature=sp.temperature, min_p=sp.min_p,
                           top_p=sp.top_p, seed=(r.rid if sp.seed is None else sp.seed))
            tmp_state.ctx_lens.fill_(0)
            tmp_state.active.fill_(1)
            tmp_state.c.stop_token = st.c.stop_token
            toks = torch.empty(nreq, dtype=torch.int32, device=dev)
            self.m.head_cfg_sample(h_last, 2 * nreq, tmp_sp, tmp_state, toks, advance=True)
            slots = torch.tensor([r.slot for r in ch], device=dev)
            rows2 = torch.stack([2 * slots, 2 * slots + 1], 1).reshape(-1)
            npr = torch.tensor([r.n_prompt for r in ch], dtype=torch.int32, device=dev).repeat_interleave(2)
            st.tokens[slots] = tmp_state.tokens
            st.speech_pos[slots] = tmp_state.speech_pos              # 1: BOS sat at 0 (t3.py:551)
            st.history[slots] = tmp_state.history
            st.hist_len[slots] = tmp_state.hist_len
            st.step[slots] = tmp_state.step
            st.active[slots] = tmp_state.active
            st.ctx_lens[rows2] = npr
            st.rope_pos[rows2] = npr
            for name in ("cfg_scale", "repetition_penalty", "temperature", "min_p", "top_p", "seed"):
                getattr(self.sampling, name)[slots] = getattr(tmp_sp, name)
            for r in ch:
                r.n_gen = 1

    # ------------------------------------------------------------------ the caller's entry point
    def generate(self, prompts: Sequence[dict], sampling_params=None) -> List[RequestOutput]:
        """tts.py:445-465.  ``prompts``: dicts with ``prompt_token_ids`` + ``multi_modal_data.conditionals``
        ([34, 1024] tensor, tts.py:452) or ready ``prefill_embeds`` (cond, uncond).  ``sampling_params``:
        one SamplingParams or one per prompt.  Returns results in prompt order; ``outputs[0].token_ids``
        are offset by SPEECH_TOKEN_OFFSET like the reference's (tts.py:474-492)."""
        if sampling_params is None:
            sampling_params = SamplingParams()
        sps = list(sampling_params) if isinstance(sampling_params, (list, tuple)) else [sampling_params] * len(prompts)
        ignore_eos = any(sp.ignore_eos for sp in sps)
        stop = -1 if ignore_eos else int(sps[0].stop_token_ids[0]) - SPEECH_TOKEN_OFFSET     # tts.py:457
        if stop != self.state.c.stop_token:
            self.state.c.stop_token = stop
            self._graphs.clear()                       # the stop token is baked into captured kernel arguments
        prompts = self._embed_prompts(list(prompts))
        waiting = [self._make_req(i, p, sp) for i, (p, sp) in enumerate(zip(prompts, sps))]
        waiting.reverse()
        results: Dict[int, RequestOutput] = {}
        st = self.state
        while waiting or any(s is not None for s in self.slots):
            new: List[_Req] = []
            for slot in range(self.max_req):
                if not waiting:
                    break
                if self.slots[slot] is None:
                    r = waiting.pop()
                    r.slot = slot
                    self.slots[slot] = r
                    new.append(r)
            if new:
                self._admit(new)
            live = [r for r in self.slots if r is not None]
            n_slots = max(r.slot for r in live) + 1
            # steps until the host has to look again: the shortest remaining length, and - only when a stop token can end a
            # request early - at most check_every
            budget = min(min(r.sp.max_tokens, self.max_len - r.n_prompt) - r.n_gen for r in live)
            if stop >= 0:
                budget = min(budget, self.check_every)
            for _ in range(max(0, budget)):
                self.step(n_slots)
            poll = torch.stack([st.hist_len[:n_slots], st.active[:n_slots].to(st.hist_len.dtype)]).cpu()    # one D2H per poll
            hist_len, active = poll[0], poll[1]
            done = []
            for r in live:
                r.n_gen = int(hist_len[r.slot])
                limit = min(r.sp.max_tokens, self.max_len - r.n_prompt)
                if not bool(active[r.slot]) or r.n_gen >= limit:
                    done.append(r)
            hist_host = st.history[:n_slots].cpu() if done else None          # one D2H for every request that finished
            for r in done:
                toks = hist_host[r.slot, :r.n_gen].tolist()
                stopped = stop >= 0 and stop in toks
                if stopped:
                    toks = toks[:toks.index(stop) + 1]
                results[r.rid] = RequestOutput(r.rid, [CompletionOutput([t + SPEECH_TOKEN_OFFSET for t in toks],
                                                                        "stop" if stopped else "length")])
                self.kv.release(r.blocks)
                self.slots[r.slot] = None
            if done:                                                           # park the freed slots: four indexed stores in all
                sl = torch.tensor([r.slot for r in done], device=self.dev)
                rw = torch.stack([2 * sl, 2 * sl + 1], 1).reshape(-1)
                st.active[sl] = 0
                st.ctx_lens[rw] = 0
                st.rope_pos[rw] = 0
                st.block_tables[rw] = 0
        return [results[i] for i in range(len(prompts))]
