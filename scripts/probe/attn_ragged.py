#!/usr/bin/env python
"""Decode attention with ragged vs uniform contexts of the same total size (load balance of the static item -> warp deal).
    python scripts/probe/attn_ragged.py [rows]"""
import ctypes as C
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from chatterbox_vllm2_b200 import T3Dims, _lib
from chatterbox_vllm2_b200.model import llama3_rope_table

L = _lib.lib()
dev = "cuda:0"
R = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
d = T3Dims()
rng = random.Random(7)
ragged = [rng.randint(100, 1100) for _ in range(R // 2) for _ in (0, 1)]
mean = sum(ragged) // R
cases = {"uniform": [mean] * R, "ragged": ragged, "ragged_sorted_desc": sorted(ragged, reverse=True)}
mb = 75
nblk_total = sum((c + 16) // 16 for c in ragged) + R + 8
kp = torch.randn(nblk_total + 1, 16, 1024, device=dev).to(torch.bfloat16)
vp = torch.randn(nblk_total + 1, 16, 1024, device=dev).to(torch.bfloat16)
rope = llama3_rope_table(d).to(device=dev, dtype=torch.bfloat16).contiguous()
qkv = torch.randn(R, 3072, device=dev).to(torch.bfloat16)
out = torch.empty(R, 1024, dtype=torch.bfloat16, device=dev)
for name, ctxs in cases.items():
    bt = torch.zeros(R, mb, dtype=torch.int32)
    nb = 1
    for r, c in enumerate(ctxs):
        n = (c + 16) // 16
        bt[r, :n] = torch.arange(nb, nb + n, dtype=torch.int32)
        nb += n
    bt = bt.to(dev)
    idx = torch.tensor([c - 1 for c in ctxs], dtype=torch.int32, device=dev)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        def run():
            for _ in range(10):
                _lib.check(L.t3b_attn_decode_fused(qkv.data_ptr(), R, idx.data_ptr(), idx.data_ptr(), rope.data_ptr(), d.max_pos, bt.data_ptr(), mb,
                                                   kp.data_ptr(), vp.data_ptr(), 16, 0, out.data_ptr(), None, side.cuda_stream))
        run(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            run()
        t = 1e9
        for rep in range(5):
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(side); g.replay(); e1.record(side); torch.cuda.synchronize()
            if rep: t = min(t, e0.elapsed_time(e1) / 10 * 1e3)
    torch.cuda.current_stream().wait_stream(side)
    byt = sum(c - 1 for c in ctxs) * 4096
    print(f"{name:20s} rows {R} mean ctx {sum(ctxs) / R:.0f}: {t:8.1f} us  {byt / t / 1e3:7.1f} GB/s")
