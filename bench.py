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

  value   device-resident: CUDA-graph replay of the step, inputs already in HBM, CUDA events.
  e2e     the same K steps through the reference-facing call a user makes, T3B200Engine.generate()
          (tts.py:445-465): host prompts in pinned memory -> H2D -> prefill -> decode -> token ids D2H;
          everything (including the prefill and the host polling) is inside the timed region.
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
    ap.add_argument("--sweep", default="", help="extra batch sizes for the 'sweep' key, e.g. 1,4,128")
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


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    # calibrate the per-step sample so K steps end within a few minutes
    batch = args.batch
    probe = cpu_oracle_rate(batch, 2, 1, 250)
    while batch > 1 and probe["ms_per_step"] * 1e-3 * (args.steps + args.warmup) > 150.0:
        batch //= 2
        probe = cpu_oracle_rate(batch, 2, 1, 250)
    r = cpu_oracle_rate(batch, args.steps, args.warmup, 250)
    sample = (f"{r['steps']} decode steps of {batch} requests ({2 * batch} CFG rows) at context 250, fp32 oracle "
              f"(oracle/t3_ref.py), torch {torch.__version__} CPU, {r['cores']} threads")
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": r["steps"],
            "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[1] decode step, CPU restatement (the reference has no CPU path; SURVEY.md 8(d))",
                       "batch_requests": batch, "cfg_rows": 2 * batch, "context": 250},
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": sample},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ B200 arm
def build_engine(batch, dev, sd=None, max_model_len=MAX_MODEL_LEN, quant=None):
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
    eng = T3B200Engine(m, max_num_seqs=batch, max_model_len=max_model_len)
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


def device_resident(eng, prompts, steps, warmup, sampler, sampled=False):
    """K graph-replayed steps from the prefilled state; the 300-step job restarts from a state snapshot."""
    import torch
    from chatterbox_vllm2_b200.engine import SamplingParams
    B = len(prompts)
    st = eng.state
    sp = SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=GEN_LEN + warmup + 2, ignore_eos=True)
    if sampled:      # configs[2]: min-p / top-p sampling with repetition penalty (tts.py:455-464 defaults)
        sp = SamplingParams(temperature=0.8, top_p=0.8, min_p=0.05, repetition_penalty=2.0, max_tokens=GEN_LEN + warmup + 2,
                            ignore_eos=True, seed=1234)
    eng.state.c.stop_token = -1
    reqs = [eng._make_req(i, p, sp) for i, p in enumerate(prompts)]
    for i, r in enumerate(reqs):
        r.slot = i
        eng.slots[i] = r
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0.record()
    eng._admit(reqs)
    t1.record()
    torch.cuda.synchronize()
    prefill_ms = t0.elapsed_time(t1)
    names = ("tokens", "speech_pos", "rope_pos", "ctx_lens", "hist_len", "step", "active")
    snap = {n: getattr(st, n).clone() for n in names}

    def restore():
        for n in names:
            getattr(st, n).copy_(snap[n])

    for _ in range(max(warmup, 3)):
        eng.step(B)
    restore()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    ctx_sum = 0
    P = reqs[0].n_prompt
    with sampler:
        e0.record()
        done = 0
        while done < steps:
            n = min(GEN_LEN - 1, steps - done)      # token 1 of the job came from the prefill
            for j in range(n):
                eng.step(B)
                ctx_sum += P + 1 + j                 # tokens each row attends to, incl. its own
            done += n
            if done < steps:
                restore()
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    toks = st.history[:B, :4].cpu().tolist()
    for r in reqs:
        eng.kv.release(r.blocks)
        eng.slots[r.slot] = None
    st.active.zero_(); st.ctx_lens.zero_(); st.rope_pos.zero_(); st.block_tables.zero_()
    return {"ms": ms, "prefill_ms": prefill_ms, "mean_ctx": ctx_sum / steps, "P": P, "first_tokens": toks[0]}


def end_to_end(eng, prompts, steps, sampler=None):
    """The user-facing call: generate() from host prompts to host token ids, ceil(K/300) jobs."""
    import torch
    from chatterbox_vllm2_b200.engine import SamplingParams
    B = len(prompts)
    h2d = d2h = 0
    t_tokens = 0
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    done = 0
    while done < steps:
        n = min(GEN_LEN, steps - done)
        outs = eng.generate(prompts, SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=n, ignore_eos=True))
        t_tokens += sum(len(o.outputs[0].token_ids) for o in outs)
        # conditionals (bf16) + text ids and positions (int32) + the int64 row indices of the batched embedding assembly
        h2d += sum(p["multi_modal_data"]["conditionals"][0].numel() * 2 + len(p["prompt_token_ids"]) * 8
                   + (2 * 34 + len(p["prompt_token_ids"]) + 2) * 8 for p in prompts)
        # generate(ignore_eos=True) polls once per job (2 int32 per slot) and reads the slots' history rows back once
        d2h += B * 8 + B * int(eng.state.c.hist_stride) * 4
        done += n
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return {"tokens": t_tokens, "seconds": dt, "h2d": h2d / steps, "d2h": d2h / steps}


def kernel_breakdown(m, eng, B, ctx):
    """Per kernel class: launch it for all 30 layers back to back (distinct weights / KV per layer, so every
    launch streams from HBM) between two CUDA events; 3 repetitions, best.  Returns ms per launch and the
    algorithmic bytes per launch."""
    import ctypes as C
    import torch
    from chatterbox_vllm2_b200 import _lib
    L = _lib.lib()
    d, dev = m.dims, m.device
    R = 2 * B
    s = torch.cuda.current_stream().cuda_stream
    bf = lambda *shape: (0.5 * torch.randn(*shape, device=dev)).to(torch.bfloat16)
    x, x4 = bf(R, d.hidden), bf(R, d.ffn)
    y_qkv, y_gu, y_d = bf(R, 3 * d.hidden), bf(R, 2 * d.ffn), bf(R, d.hidden)
    act, q, attn, res = bf(R, d.ffn), bf(R, d.hidden), bf(R, d.hidden), bf(R, d.hidden)
    st, kv = eng.state, eng.kv
    # context: every row holds `ctx` tokens in its own blocks
    nblk = (ctx + 16) // 16
    blocks = kv.alloc(R * nblk)
    bt = torch.zeros(R, eng.max_blocks, dtype=torch.int32)
    bt[:, :nblk] = torch.tensor(blocks, dtype=torch.int32).reshape(R, nblk)
    bt = bt.to(dev)
    idx = torch.full((R,), ctx - 1, dtype=torch.int32, device=dev)
    ns = C.c_int32(0)
    auto = m.c.layers

    def lin(name, xin, yout, epi=0):
        def f(li):
            w = getattr(auto[li], name)
            _lib.check(L.t3b_linear_bf16(xin.data_ptr(), C.byref(w), R, yout.data_ptr(), None, 0, C.byref(ns), 0, epi, s), name)
        return f

    def attn_f(li):
        off = li * kv.c.layer_stride * 2
        _lib.check(L.t3b_attn_decode_fused(y_qkv.data_ptr(), R, idx.data_ptr(), idx.data_ptr(), m.w["rope"].data_ptr(), d.max_pos,
                                           bt.data_ptr(), eng.max_blocks, kv.k.data_ptr() + off, kv.v.data_ptr() + off, d.n_heads,
                                           attn.data_ptr(), s), "attn")

    def norm_f(li):
        _lib.check(L.t3b_add_rmsnorm(y_d.data_ptr(), None, 1, res.data_ptr(), m.w[f"{li}.ln1"].data_ptr(), x.data_ptr(), R, d.hidden,
                                     d.rms_eps, 1, s), "norm")

    classes = {
        "qkv_proj": (lin("qkv", x, y_qkv), 3 * d.hidden * d.hidden * 2, 1),
        "attention_rope_kvwrite": (attn_f, R * ctx * d.n_heads * d.head_dim * 2 * 2, 1),
        "o_proj": (lin("o", attn, y_d), d.hidden * d.hidden * 2, 1),
        "gate_up_silu_proj": (lin("gate_up", x, act, 1), 2 * d.ffn * d.hidden * 2, 1),
        "down_proj": (lin("down", x4, y_d), d.hidden * d.ffn * 2, 1),
        "add_rmsnorm": (norm_f, R * d.hidden * 2 * 4, 2),
    }
    out = {}
    side = torch.cuda.Stream(device=dev)
    for name, (fn, bytes_per_launch, per_layer) in classes.items():
        # the 30 launches of a class are captured into ONE CUDA graph (like the real step): a Python loop of ctypes calls
        # is bound by the host's ~5 us per launch and hides every kernel faster than that
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            s = side.cuda_stream
            for li in range(d.n_layers):
                fn(li)                                   # warm-up (tensor maps, attributes) outside the capture
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                s = side.cuda_stream
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
    m, eng, sd = build_engine(B, dev)
    prompts = make_prompts(B, rank)

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

    # ---- device-resident
    sampler = ClockSampler(local)
    barrier()
    r = device_resident(eng, prompts, K, W, sampler)
    barrier()
    ms = max_over_ranks(r["ms"])
    value = world * B * K / (ms * 1e-3)
    launches = eng.launches_per_step(B) * K

    # ---- end to end through generate()
    end_to_end(eng, prompts, min(K, 16))          # warm the graphs / allocator for this path
    barrier()
    e = end_to_end(eng, prompts, K)
    barrier()
    e_s = max_over_ranks(e["seconds"])
    e2e_value = world * e["tokens"] / e_s

    W_bytes = m.weight_bytes()
    step_bytes = W_bytes + KV_BYTES_PER_TOKEN_ROW * 2 * B * r["mean_ctx"]
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(W, 3), "ms_per_step": ms / K,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "BASELINE.json configs[1]: T3 bf16, 32 requests (64 CFG rows) per GPU, P=99 prompt, 300 speech tokens, "
                               "greedy, CFG 0.5, rep. penalty 2.0, paged KV block 16",
                   "batch_requests_per_gpu": B, "cfg_rows_per_gpu": 2 * B, "prompt_tokens": r["P"], "gen_len": GEN_LEN,
                   "mean_context": r["mean_ctx"], "weights": "seeded random-init, exact T3 Llama-520M shapes",
                   "l2": "inputs larger than L2: every step streams %.2f GB of weights + KV (L2 = 126 MB)" % (step_bytes / 1e9),
                   "parallelism": f"replicas x{world} (no collective on the data path)"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": e["h2d"], "d2h_bytes_per_step": e["d2h"],
                "api": "T3B200Engine.generate(prompts, SamplingParams) incl. prefill and host polling", "seconds": e_s},
        "gpu_launches": launches,
        "prefill_ms": r["prefill_ms"],
        "step_roofline": {"bound": "hbm", "bytes_per_step": step_bytes, "achieved": step_bytes / (ms / K) / 1e6, "peak": pk["hbm_gbs"],
                          "unit": "GB/s", "frac": step_bytes / (ms / K) / 1e6 / pk["hbm_gbs"], "peak_source": pk["source"]},
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
    if rank == 0 and args.sweep:
        sw = {}
        for tok in [v for v in args.sweep.split(",") if v]:
            # "N" = N bf16 requests, greedy; "awqN" = N requests on AWQ W4A16 weights with min-p / top-p sampling (configs[2])
            awq = tok.startswith("awq")
            b = int(tok[3:] if awq else tok)
            del eng
            torch.cuda.empty_cache()
            m2, eng, _ = build_engine(b, dev, sd, quant="awq" if awq else None)
            rr = device_resident(eng, make_prompts(b, rank, pin=False), min(K, GEN_LEN - 1), W, ClockSampler(local), sampled=awq)
            kk = min(K, GEN_LEN - 1)
            wb = m2.weight_bytes()
            bytes_ = wb + KV_BYTES_PER_TOKEN_ROW * 2 * b * rr["mean_ctx"]
            sw[tok] = {"tokens_per_s": b * kk / (rr["ms"] * 1e-3), "ms_per_step": rr["ms"] / kk,
                       "hbm_frac": bytes_ / (rr["ms"] / kk) / 1e6 / pk["hbm_gbs"], "weight_bytes": wb,
                       "weights": "awq_w4a16" if awq else "bf16", "sampling": "temperature 0.8, top_p 0.8, min_p 0.05, rep 2.0" if awq else "greedy, rep 2.0"}
            del m2
        line["sweep"] = sw
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        c = cpu_oracle_rate(B, 12, 1, 250, sd=sd, budget_s=25.0)
        line["cpu_baseline"] = {"value": c["value"], "unit": UNIT, "cores": c["cores"], "kind": "port",
                                "sample": f"{c['steps']} decode steps of {B} requests ({2 * B} CFG rows) at context 250, fp32 oracle "
                                          f"(oracle/t3_ref.py; the reference has no CPU path), torch CPU, {c['cores']} threads, "
                                          f"{c['seconds']:.1f} s"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
