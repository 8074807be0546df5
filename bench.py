#!/usr/bin/env python
"""bench.py - T3 CFG decode throughput on B200 (BASELINE.json metric: T3 speech tokens/s per B200).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--batch B]

Workload at N=1 = BASELINE.json configs[1]: T3 bf16, batch 32 requests (64 CFG rows), synthetic 34-row
conditioning prefix + 64 text tokens + BOS (P = 99), 300 speech tokens per request, greedy, CFG w = 0.5,
repetition penalty 2.0, paged KV (block 16).  Weights are the seeded random-init checkpoint of the exact
architecture (no network for the real file).  N > 1: one replica per GPU, each with its own 32 requests
(weak scaling, no collective on the data path; NCCL is used for the barrier / max-over-ranks only).

A "step" is one decode step of the whole batch: embedding gather -> 30 layers -> speech head -> CFG combine
-> penalty -> argmax -> state advance.  The job (300 steps from the prefilled state) is repeated until K
steps have run; the prefill is outside the timed region and reported separately.

  value   device-resident: CUDA-graph replay of the step, inputs already in HBM, CUDA events.  The K timed steps are
          spread evenly over the 300-step job (decode positions 1..299 -> contexts 100..398, mean 249 at any K), so a
          short run measures the configuration it names.
  e2e     the same job through the reference-facing call a user makes, T3B200Engine.generate()
          (tts.py:445-465): host prompts in pinned memory -> H2D -> prefill -> 300 tokens per request -> token ids
          D2H; everything (including the prefill and the host polling) is inside the timed region; ceil(K/300) jobs.
  roofline  dominant kernel class of the step (per-class CUDA-event timing over all 30 layers'
          weights, L2-cold because 30 layers x weights >> L2), algorithmic bytes / launch / duration
          against MEASURED_PEAKS.json.
  cpu_baseline  the fp32 oracle (the reference has no CPU implementation of this path)
          timed on this box's host cores on a bounded sample of the same workload.

--impl reference times that CPU oracle as the whole arm (rank 0 only).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "T3 speech tokens/s per B200 (CFG, bf16, batch 32 requests = 64 CFG rows)"
UNIT = "tokens/s"
N_TEXT, GEN_LEN, MAX_MODEL_LEN = 64, 300, 1200
KV_BYTES_PER_TOKEN_ROW = 122880          # 30 layers x 2 (K,V) x 16 heads x 64 x 2 B   


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=GEN_LEN)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=32, help="requests per GPU (CFG rows = 2x)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-breakdown", action="store_true")
    ap.add_argument("--sweep", default="1,8,awq256,cfg4", help="extra points for the 'sweep' key: N = N bf16 requests greedy; awqN = configs[2] "
                                                             "(AWQ W4A16, sampling); cfg4 = configs[3] (128 requests at context ~1000); fp8N = N "
                                                             "requests on an FP8 KV pool; '' = none")
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg5"], help="cfg2 = BASELINE.json configs[1] (default); cfg5 = "
                    "configs[4]: 4096 mixed-length requests through the dispatcher + generate()")
    ap.add_argument("--requests", type=int, default=4096, help="cfg5: total requests over all GPUs")
    ap.add_argument("--opt", default="", help="A/B of launch-shape options: comma-separated t3b_options_t fields, e.g. "
                                              "no_weight_prefetch=1,attn_team_warps=2 (applied to every engine this run builds)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                "source": "MEASURED_PEAKS.json"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1400.0, "source": "fallback (B200_PROFILING.md)"}


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""
    BAD = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20}
    NOTE = {"sw_power_cap": 0x4}

    def __init__(self, index: int):
        self.samples, self.reasons, self.ok = [], set(), False
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # noqa: BLE001 - clocks are evidence, not a dependency
            self.err = repr(e)

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(self.nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for name, bit in {**self.BAD, **self.NOTE}.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(0.01)

    def __enter__(self):
        if self.ok:
            self.t = threading.Thread(target=self._run, daemon=True)
            self.t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self.ok:
            self.t.join()

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": getattr(self, "err", "no samples")}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}




WORKLOAD = ("BASELINE.json configs[1]: T3 bf16, 32 requests (64 CFG rows) per GPU, P=99 prompt, 300 speech tokens, greedy, CFG 0.5, "
            "rep. penalty 2.0, paged KV block 16")
P_PROMPT = 34 + N_TEXT + 1


def spread_positions(k_steps: int, gen_len: int = GEN_LEN):
    """Decode positions (tokens already generated, 1..gen_len-1) of the K timed steps, spread evenly over the job: the step at
    position g attends to P + g tokens.  K >= gen_len - 1 wraps around the job."""
    n = gen_len - 1
    if k_steps >= n:
        return [1 + (i % n) for i in range(k_steps)]
    if k_steps == 1:
        return [(n + 1) // 2]
    return [1 + round(i * (n - 1) / (k_steps - 1)) for i in range(k_steps)]


def mean_context(k_steps: int) -> float:
    pos = spread_positions(k_steps)
    return P_PROMPT + sum(pos) / len(pos)


# ------------------------------------------------------------------------------------------ CPU oracle arm
def cpu_oracle_rate(batch: int, steps: int, warmup: int, ctx: int, sd=None, budget_s: float = None):
    """Decode steps of the fp32 CPU oracle (oracle/t3_ref.py) for ``batch`` requests whose KV already holds
    ``ctx`` tokens (filled with seeded noise: the cost of a step does not depend on the values)."""
    import torch
    from chatterbox_vllm2_b200 import T3Dims
    from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint
    from oracle.t3_ref import T3Oracle
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    dims = T3Dims()
    if sd is None:
        sd = synthetic_checkpoint(dims)
    orc = T3Oracle(sd, dims, mode="fp32", max_len=ctx + steps + warmup + 2)
    rows = orc.new_rows(2 * batch)
    g = torch.Generator().manual_seed(5)
    for r in rows:
        for li in range(dims.n_layers):
            orc.kc[r][li][:ctx].uniform_(-1, 1, generator=g)
            orc.vc[r][li][:ctx].uniform_(-1, 1, generator=g)
        orc.len[r] = ctx
    toks = [(17 * b) % 6561 for b in range(batch)]
    done = 0
    for _ in range(warmup):
        orc.decode_step(rows, toks, [1 + done] * batch)
        done += 1
    t0 = time.perf_counter()
    n = 0
    for _ in range(steps):
        h = orc.decode_step(rows, toks, [1 + done] * batch)
        lg = orc.compute_logits(h)
        toks = lg.argmax(-1).tolist()
        done += 1
        n += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return {"value": batch * n / dt, "steps": n, "seconds": dt, "cores": cores, "ms_per_step": 1e3 * dt / n}


def bench_config(batch: int, world: int, k_steps: int):
    """The `config` object both arms print (the driver compares them)."""
    sb = 1023539200 + KV_BYTES_PER_TOKEN_ROW * 2 * batch * mean_context(k_steps)
    return {"workload": WORKLOAD, "batch_requests_per_gpu": batch, "cfg_rows_per_gpu": 2 * batch, "prompt_tokens": P_PROMPT,
            "gen_len": GEN_LEN, "mean_context": round(mean_context(k_steps), 1),
            "weights": "seeded random-init, exact T3 Llama-520M shapes",
            "l2": "inputs larger than L2: every step streams %.2f GB of weights + KV (L2 = 126 MB)" % (sb / 1e9),
            "parallelism": f"replicas x{world} (no collective on the data path)"}


def run_reference(args):
    """The reference's CPU arm.  The reference has no CPU (or any own) implementation of this path - its arithmetic is vLLM's CUDA
    kernels - so this times the fp32 oracle port on the host cores, on the same workload shape as the B200 arm: the same batch, decode
    steps at the job's mean context (every step of the port costs the same at a given context, so K steps at the mean context are
    the bounded sample of the 300-step job)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    batch = args.batch
    ctx = int(round(mean_context(args.steps))) - 1          # cached tokens in front of the step's own token
    r = cpu_oracle_rate(batch, args.steps, args.warmup, ctx)
    sample = (f"{r['steps']} decode steps of {batch} requests ({2 * batch} CFG rows) at context {ctx + 1} (the job's mean), fp32 oracle "
              f"(oracle/t3_ref.py), torch {torch.__version__} CPU, {r['cores']} threads")
    cfg = bench_config(batch, args.gpus, args.steps)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": r["steps"],
            "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": cfg,
            "notes": {"arm": "CPU restatement of the path (the reference has no CPU implementation; SURVEY.md 8(d)); rank 0 only"},
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": sample},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ B200 arm
def build_engine(batch, dev, sd=None, max_model_len=MAX_MODEL_LEN, quant=None, kv_dtype="bf16", kv_tokens_per_row=None):
    import torch
    from chatterbox_vllm2_b200 import T3Dims
    from chatterbox_vllm2_b200.engine import T3B200Engine
    from chatterbox_vllm2_b200.model import T3B200Model
    from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint
    dims = T3Dims()
    if sd is None:
        sd = synthetic_checkpoint(dims)
    m = T3B200Model(dims=dims, device=dev, cfg_scale=0.5, quantization=quant)
    m.load_weights(sd.items())
    kv_blocks = None
    if kv_tokens_per_row is not None:
        kv_blocks = 2 * batch * ((kv_tokens_per_row + 15) // 16 + 1) + 8
    eng = T3B200Engine(m, max_num_seqs=batch, max_model_len=max_model_len, kv_blocks=kv_blocks, kv_dtype=kv_dtype)
    for kv in [t for t in OPT_OVERRIDES.split(",") if t]:
        k, v = kv.split("=")
        setattr(eng.opt, k, int(v))
    if eng.opt.gemv_mma > 0:
        m.enable_gemv_mma()
    return m, eng, sd


def make_prompts(batch, rank, pin=True):
    """Host-side prompts: conditionals [34,1024] bf16 in pinned memory + 64 text ids (seed 1000+i)."""
    from chatterbox_vllm2_b200 import T3Dims
    from chatterbox_vllm2_b200.synthetic import synthetic_request
    out = []
    for i in range(batch):
        cond, ids = synthetic_request(rank * 100000 + i, T3Dims(), n_text=N_TEXT)
        out.append({"prompt_token_ids": ids, "multi_modal_data": {"conditionals": [cond.pin_memory() if pin else cond]}})
    return out


def device_resident(eng, prompts, steps, warmup, sampler, sampled=False, gen_len=GEN_LEN):
    """K graph-replayed decode steps at positions spread evenly over the ``gen_len``-token job (spread_positions): before every
    step ONE device-to-device copy sets the counters of that position (tokens generated, context length, RoPE / speech positions,
    sampler step); K/V beyond what the warm-up really generated is whatever the pool holds - a step's cost does not depend on the
    values.  The prefill (admission of all requests) is timed separately."""
    import torch
    from chatterbox_vllm2_b200.engine import SamplingParams
    B = len(prompts)
    st = eng.state
    sp = SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=gen_len, ignore_eos=True)
    if sampled:      # configs[2]: min-p / top-p sampling with repetition penalty (tts.py:455-464 defaults)
        sp = SamplingParams(temperature=0.8, top_p=0.8, min_p=0.05, repetition_penalty=2.0, max_tokens=gen_len, ignore_eos=True, seed=1234)
    eng.uninstall(eng.install(prompts, sp, reserve_steps=1))        # warm the prefill path (tensor maps, kernel attributes, allocator)
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0.record()
    reqs = eng.install(prompts, sp, reserve_steps=gen_len + 1)
    t1.record()
    torch.cuda.synchronize()
    prefill_ms = t0.elapsed_time(t1)
    P = reqs[0].n_prompt
    M = st.max_req
    base = st.small.clone()
    pos = spread_positions(steps, gen_len)

    def counters(g):
        s = base.clone()
        s[1 * M:1 * M + B] = g            # speech_pos
        s[2 * M:2 * M + B] = g            # hist_len
        s[3 * M:3 * M + B] = g            # sampler step
        s[5 * M:5 * M + 2 * B] = P + g - 1    # rope_pos
        s[7 * M:7 * M + 2 * B] = P + g - 1    # ctx_lens (cached tokens in front of the step's own)
        return s

    table = torch.stack([counters(g) for g in pos])
    # history of a position g must hold g valid ids: fill the rows once with a seeded id stream (the first one is the real token)
    gen = torch.Generator(device="cpu").manual_seed(11)
    st.history[:B, 1:gen_len] = torch.randint(0, 6561, (B, gen_len - 1), generator=gen, dtype=torch.int32).to(st.history.device)
    for i in range(max(warmup, 3)):
        st.small.copy_(table[i % len(pos)])
        eng.step(B)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    with sampler:
        e0.record()
        for i in range(steps):
            st.small.copy_(table[i])
            eng.step(B)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    toks = st.history[:B, :4].cpu().tolist()
    eng.uninstall(reqs)
    return {"ms": ms, "prefill_ms": prefill_ms, "mean_ctx": P + sum(pos) / len(pos), "P": P, "first_tokens": toks[0],
            "positions": [pos[0], pos[-1]]}


def end_to_end(eng, prompts, steps):
    """The user-facing call: generate() from host prompts to host token ids, ceil(K/300) full 300-token jobs."""
    import torch
    from chatterbox_vllm2_b200.engine import SamplingParams
    B = len(prompts)
    jobs = max(1, -(-steps // GEN_LEN))
    h2d = d2h = 0
    t_tokens = 0
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(jobs):
        outs = eng.generate(prompts, SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=GEN_LEN, ignore_eos=True))
        t_tokens += sum(len(o.outputs[0].token_ids) for o in outs)
        # conditionals (bf16) + one int32 buffer (text ids, offsets, slots, sequence rows) + block tables (int32 rows) and the
        # int64 (index, block) pairs of every block-table extension
        blocks = 2 * ((P_PROMPT + GEN_LEN + 15) // 16 + 2)
        h2d += sum(34 * 1024 * 2 + (len(p["prompt_token_ids"]) + 6) * 4 for p in prompts) + B * 2 * eng.max_blocks * 4 + B * blocks * 16
        # the finished requests' history rows come back once per job
        d2h += B * int(eng.state.hist_stride) * 4
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return {"tokens": t_tokens, "seconds": dt, "h2d": h2d / (jobs * GEN_LEN), "d2h": d2h / (jobs * GEN_LEN), "jobs": jobs}


def kernel_breakdown(m, eng, B, ctx):
    """Per kernel class: its 30 per-layer launches (distinct weights / KV per layer, so every launch streams from HBM) captured
    into ONE CUDA graph and replayed between two CUDA events; 5 repetitions, best of the last 4.  Returns ms per launch and the
    algorithmic bytes per launch."""
    import ctypes as C
    import torch
    from chatterbox_vllm2_b200 import _lib
    L = _lib.lib()
    d, dev = m.dims, m.device
    R = 2 * B
    bf = lambda *shape: (0.5 * torch.randn(*shape, device=dev)).to(torch.bfloat16)
    x, x4 = bf(R, d.hidden), bf(R, d.ffn)
    y_qkv, y_d = bf(R, 3 * d.hidden), bf(R, d.hidden)
    act, attn, res = bf(R, d.ffn), bf(R, d.hidden), bf(R, d.hidden)
    kv = eng.kv
    nblk = (ctx + 16) // 16
    blocks = kv.alloc(R * nblk)
    bt = torch.zeros(R, eng.max_blocks, dtype=torch.int32)
    bt[:, :nblk] = torch.tensor(blocks, dtype=torch.int32).reshape(R, nblk)
    bt = bt.to(dev)
    idx = torch.full((R,), ctx - 1, dtype=torch.int32, device=dev)
    ns = C.c_int32(0)
    auto = m.c.layers
    cur = {"s": torch.cuda.current_stream().cuda_stream}
    awq = m.quantization == "awq"
    lin_fn = L.t3b_linear_awq if awq else L.t3b_linear_bf16

    def lin(name, xin, yout, epi=0):
        def f(li):
            w = getattr(auto[li], name)
            _lib.check(lin_fn(xin.data_ptr(), C.byref(w), R, yout.data_ptr(), None, 0, C.byref(ns), 0, epi, C.byref(eng.opt), cur["s"]), name)
        return f

    def attn_f(li):
        off = li * kv.c.layer_stride
        _lib.check(L.t3b_attn_decode_fused(y_qkv.data_ptr(), R, idx.data_ptr(), idx.data_ptr(), m.w["rope"].data_ptr(), d.max_pos,
                                           bt.data_ptr(), eng.max_blocks, kv.k.data_ptr() + off, kv.v.data_ptr() + off, d.n_heads,
                                           kv.kv_dtype, attn.data_ptr(), C.byref(eng.opt), cur["s"]), "attn")

    def norm_f(li):
        _lib.check(L.t3b_add_rmsnorm(y_d.data_ptr(), None, 1, res.data_ptr(), m.w[f"{li}.ln1"].data_ptr(), x.data_ptr(), R, d.hidden,
                                     d.rms_eps, 1, cur["s"]), "norm")

    wb = (lambda n, k: n * k // 2 + (k // 128) * n * 3) if awq else (lambda n, k: n * k * 2)
    kv_tok = 2 * d.n_heads * kv.slice_bytes // 16            # K + V bytes per token-row per layer
    classes = {
        "qkv_proj": (lin("qkv", x, y_qkv), wb(3 * d.hidden, d.hidden), 1),
        "attention_rope_kvwrite": (attn_f, R * ctx * kv_tok, 1),
        "o_proj": (lin("o", attn, y_d), wb(d.hidden, d.hidden), 1),
        "gate_up_silu_proj": (lin("gate_up", x, act, 1), wb(2 * d.ffn, d.hidden), 1),
        "down_proj": (lin("down", x4, y_d), wb(d.hidden, d.ffn), 1),
        "add_rmsnorm": (norm_f, R * d.hidden * 2 * 4, 2),
    }
    out = {}
    side = torch.cuda.Stream(device=dev)
    for name, (fn, bytes_per_launch, per_layer) in classes.items():
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            cur["s"] = side.cuda_stream
            for li in range(d.n_layers):
                fn(li)                                   # warm-up (tensor maps, attributes) outside the capture
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                cur["s"] = side.cuda_stream
                for li in range(d.n_layers):
                    fn(li)
            best = None
            for rep in range(5):
                e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                e0.record(side)
                g.replay()
                e1.record(side)
                torch.cuda.synchronize()
                t = e0.elapsed_time(e1) / d.n_layers
                if rep > 0:
                    best = t if best is None else min(best, t)
            del g
        torch.cuda.current_stream().wait_stream(side)
        out[name] = {"ms_per_launch": best, "bytes_per_launch": bytes_per_launch, "launches_per_step": per_layer * d.n_layers,
                     "gbs": bytes_per_launch / best / 1e6}
    kv.release(blocks)
    return out


def step_bytes(m, eng, B, mean_ctx):
    """Algorithmic bytes of one decode step: weights once + K/V of every row's context."""
    kv_tok = 2 * m.dims.n_layers * m.dims.n_heads * eng.kv.slice_bytes // 16
    return m.weight_bytes() + kv_tok * 2 * B * mean_ctx


def sweep_point(tok, dev, sd, rank, K, W, pk, local):
    """One extra measurement of the metric's other batch points / BASELINE.json configs:
    "N" = N bf16 requests, greedy (300-token job); "awqN" = configs[2] (AWQ W4A16, N requests, min-p / top-p sampling, rep. penalty);
    "cfg4" = configs[3] (bf16, 128 requests, 1000-token job: contexts 100..1098); "fp8N" = N bf16-weight requests on an FP8 KV pool;
    "awqfp8N" = configs[2] on an FP8 KV pool (the KV-traffic reducer of SURVEY.md 8(f) rank 2 where K/V is 98 % of the bytes)."""
    import torch
    import re
    awq, fp8, cfg4 = "awq" in tok, "fp8" in tok, tok == "cfg4"
    b = 128 if cfg4 else int(re.sub(r"[a-z]", "", tok.replace("fp8", "")))
    gen_len = 1000 if cfg4 else GEN_LEN
    m2, eng, _ = build_engine(b, dev, sd, quant="awq" if awq else None, kv_dtype="fp8" if fp8 else "bf16",
                              kv_tokens_per_row=P_PROMPT + gen_len + 40)
    rr = device_resident(eng, make_prompts(b, rank, pin=False), K, W, ClockSampler(local), sampled=awq, gen_len=gen_len)
    bytes_ = step_bytes(m2, eng, b, rr["mean_ctx"])
    ms_step = rr["ms"] / K
    flops = 2 * b * 2 * 511707136 + 122880 * 2 * b * rr["mean_ctx"]
    out = {"tokens_per_s": b * K / (rr["ms"] * 1e-3), "ms_per_step": ms_step, "steps": K, "mean_context": round(rr["mean_ctx"], 1),
           "bytes_per_step": bytes_, "hbm_frac": bytes_ / ms_step / 1e6 / pk["hbm_gbs"], "weight_bytes": m2.weight_bytes(),
           "tensor_frac_whole_step": flops / (ms_step * 1e-3) / 1e12 / pk["bf16_tflops"],
           "weights": "awq_w4a16" if awq else "bf16", "kv": eng.kv.dtype, "gen_len": gen_len, "prefill_ms": rr["prefill_ms"],
           "sampling": "temperature 0.8, top_p 0.8, min_p 0.05, rep 2.0" if awq else "greedy, rep 2.0"}
    del eng, m2
    torch.cuda.empty_cache()
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (the T3 decode path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(dev))
    from chatterbox_vllm2_b200 import _lib
    _lib.check(_lib.lib().t3b_check_device(), "t3b_check_device")
    pk = peaks()
    B, K, W = args.batch, args.steps, args.warmup

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if args.workload == "cfg5":
        run_cfg5(args, rank, world, local, dev, pk, barrier, max_over_ranks)
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    m, eng, sd = build_engine(B, dev, kv_tokens_per_row=2 * (P_PROMPT + GEN_LEN + 40))
    prompts = make_prompts(B, rank)

    # ---- device-resident
    sampler = ClockSampler(local)
    barrier()
    r = device_resident(eng, prompts, K, W, sampler)
    barrier()
    ms = max_over_ranks(r["ms"])
    value = world * B * K / (ms * 1e-3)
    launches = eng.launches_per_step(B) * K

    # ---- end to end through generate(): the full 300-token job
    end_to_end(eng, prompts, 1)          # warm the graphs / allocator for this path
    barrier()
    e = end_to_end(eng, prompts, K)
    barrier()
    e_s = max_over_ranks(e["seconds"])
    e2e_value = world * e["tokens"] / e_s

    sb = step_bytes(m, eng, B, r["mean_ctx"])
    cfg = bench_config(B, world, K)
    notes = {"options": args.opt or "defaults", "decode_positions": r["positions"], "measured_mean_context": r["mean_ctx"],
             "timed_region": "K CUDA-graph replays of the step + one 9*B int32 device copy per step that selects the decode position"}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(W, 3), "ms_per_step": ms / K,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": cfg, "notes": notes,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": e["h2d"], "d2h_bytes_per_step": e["d2h"],
                "api": "T3B200Engine.generate(prompts, SamplingParams): %d job(s) of %d requests x %d tokens incl. prefill, block-table "
                       "growth and result read-back" % (e["jobs"], B, GEN_LEN), "seconds": e_s},
        "gpu_launches": launches,
        "prefill_ms": r["prefill_ms"],
        "step_roofline": {"bound": "hbm", "bytes_per_step": sb, "achieved": sb / (ms / K) / 1e6, "peak": pk["hbm_gbs"],
                          "unit": "GB/s", "frac": sb / (ms / K) / 1e6 / pk["hbm_gbs"], "peak_source": pk["source"]},
        "clocks": sampler.summary(),
    }
    if rank == 0 and not args.no_breakdown:
        ctx = int(round(r["mean_ctx"]))
        bd = kernel_breakdown(m, eng, B, ctx)
        tot = sum(v["ms_per_launch"] * v["launches_per_step"] for v in bd.values())
        top = max(bd, key=lambda k: bd[k]["ms_per_launch"] * bd[k]["launches_per_step"])
        t = bd[top]
        line["kernels"] = {k: {"ms_per_launch": round(v["ms_per_launch"], 5), "share": round(v["ms_per_launch"] * v["launches_per_step"] / tot, 4),
                               "GBps": round(v["gbs"], 1)} for k, v in bd.items()}
        traffic, traffic_note = None, None
        tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(tp):
            tj = json.load(open(tp)).get(top)
            if tj:
                traffic = tj["dram_bytes_per_launch"]
                traffic_note = {k: tj[k] for k in ("context", "algorithmic_bytes_at_that_context", "source") if k in tj}
        line["roofline"] = {"bound": "hbm", "kernel": top, "achieved": t["gbs"], "peak": pk["hbm_gbs"], "unit": "GB/s",
                            "frac": t["gbs"] / pk["hbm_gbs"], "traffic": traffic, "traffic_capture": traffic_note,
                            "bytes_per_launch": t["bytes_per_launch"],
                            "ms_per_launch": t["ms_per_launch"], "context": ctx, "peak_source": pk["source"]}
    pp = os.path.join(ROOT, "profiles", "parity_latest.json")
    if os.path.exists(pp):
        line["parity"] = json.load(open(pp))
    if rank == 0 and world == 1 and args.sweep:
        del eng
        torch.cuda.empty_cache()
        sw = {}
        for tok in [v for v in args.sweep.split(",") if v]:
            try:
                sw[tok] = sweep_point(tok, dev, sd, rank, min(K, 100), W, pk, local)
            except Exception as ex:  # noqa: BLE001 - a sweep point must not take the headline line down
                sw[tok] = {"error": repr(ex)[:300]}
        line["sweep"] = sw
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ctx = int(round(mean_context(K))) - 1
        c = cpu_oracle_rate(B, 12, 1, ctx, sd=sd, budget_s=20.0)
        c0 = cpu_oracle_rate(1, 40, 1, ctx, sd=sd, budget_s=10.0)      # BASELINE.json configs[0]: batch 1, CFG 0.5, greedy, fp32 CPU
        line["cpu_baseline"] = {"value": c["value"], "unit": UNIT, "cores": c["cores"], "kind": "port",
                                "sample": f"{c['steps']} decode steps of {B} requests ({2 * B} CFG rows) at context {ctx + 1}, fp32 oracle "
                                          f"(oracle/t3_ref.py; the reference has no CPU path), torch CPU, {c['cores']} threads, "
                                          f"{c['seconds']:.1f} s",
                                "configs0_batch1": {"value": c0["value"], "unit": UNIT, "ms_per_step": c0["ms_per_step"],
                                                    "sample": f"{c0['steps']} decode steps of 1 request (2 CFG rows) at context {ctx + 1}"}}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def cfg5_lengths(n_requests: int):
    """BASELINE.json configs[4]: lengths U{100..1000}, seed 7."""
    import random
    rng = random.Random(7)
    return [rng.randint(100, 1000) for _ in range(n_requests)]


def run_cfg5(args, rank, world, local, dev, pk, barrier, max_over_ranks):
    """BASELINE.json configs[4]: `--requests` synthetic requests with mixed 100-1000-token lengths (EOS masked, so the work is
    fixed), assigned to the replicas by the least-outstanding-KV-tokens dispatcher and run through T3B200Engine.generate() -
    continuous batching, growing block tables, prefix sharing.  Total work is fixed: strong scaling."""
    import torch
    from chatterbox_vllm2_b200.dispatch import assign_requests, kv_traffic_cost
    from chatterbox_vllm2_b200.engine import SamplingParams
    lens = cfg5_lengths(args.requests)
    shares = assign_requests([kv_traffic_cost(P_PROMPT, n) for n in lens], world)
    mine = shares[rank]
    slots = min(512, max(1, len(mine)))
    free_b = torch.cuda.mem_get_info()[0]
    budget = int(free_b * 0.80) - 6 * (1 << 30)
    kv_blocks = max(1024, budget // (16 * KV_BYTES_PER_TOKEN_ROW))
    from chatterbox_vllm2_b200 import T3Dims
    from chatterbox_vllm2_b200.engine import T3B200Engine
    from chatterbox_vllm2_b200.model import T3B200Model
    from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint, synthetic_request
    dims = T3Dims()
    m = T3B200Model(dims=dims, device=dev, cfg_scale=0.5)
    m.load_weights(synthetic_checkpoint(dims).items())
    eng = T3B200Engine(m, max_num_seqs=slots, max_model_len=MAX_MODEL_LEN, kv_blocks=kv_blocks, prefill_chunk_tokens=16384)
    prompts, sps = [], []
    for i in mine:
        cond, ids = synthetic_request(i % 64, dims, n_text=N_TEXT)            # 64 voices / texts, reused across requests
        prompts.append({"prompt_token_ids": ids, "multi_modal_data": {"conditionals": [cond]}})
        sps.append(SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=lens[i], ignore_eos=True))
    # warm-up: capture the graphs of every slot bucket on a small job
    eng.generate(prompts[:min(len(prompts), slots)], [SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=4, ignore_eos=True)] *
                 min(len(prompts), slots))
    sampler = ClockSampler(local)
    barrier()
    t0 = time.perf_counter()
    with sampler:
        outs = eng.generate(prompts, sps)
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    barrier()
    dt = max_over_ranks(dt)
    toks = sum(len(o.outputs[0].token_ids) for o in outs)
    assert toks == sum(lens[i] for i in mine)
    total = sum(lens)
    steps = eng.steps_run
    line = {"metric": "T3 speech tokens/s, BASELINE.json configs[4] (4096 mixed-length requests, data-parallel replicas)", "value": total / dt,
            "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": 1, "ms_per_step": 1e3 * dt / max(steps, 1), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "BASELINE.json configs[4]: T3 bf16, %d synthetic requests, lengths U{100..1000} seed 7, EOS masked, "
                                   "P=99, greedy, dispatcher = least outstanding KV traffic" % args.requests,
                       "requests": args.requests, "slots_per_gpu": slots, "kv_blocks_per_gpu": kv_blocks, "total_tokens": total,
                       "requests_rank0": len(mine), "preemptions_rank0": eng.preemptions,
                       "parallelism": f"replicas x{world} (no collective on the data path)"},
            "e2e": {"value": total / dt, "unit": UNIT, "h2d_bytes_per_step": sum(34 * 1024 * 2 + 70 * 4 for _ in mine) / max(steps, 1),
                    "d2h_bytes_per_step": len(mine) * MAX_MODEL_LEN * 4 / max(steps, 1),
                    "api": "T3B200Engine.generate(prompts, per-request SamplingParams)", "seconds": dt},
            "gpu_launches": steps * eng.launches_per_step(slots), "clocks": sampler.summary()}
    if rank == 0:
        print(json.dumps(line), flush=True)


OPT_OVERRIDES = ""


def main():
    global OPT_OVERRIDES
    args = parse()
    OPT_OVERRIDES = args.opt
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
