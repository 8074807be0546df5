#!/usr/bin/env python
"""Time the SIMT GEMV of the four decode projections in a CUDA graph of 30 PDL-chained launches (bench.py's per-class breakdown):
SIMT kernel (<= 4 rows) vs the tensor-core kernel (t3b_options_t.gemv_mma, <= 8 rows), 30 distinct weight sets (HBM) vs one weight set (L2-resident: what is left is the
dependent latency of a launch).      python scripts/probe/gemv_probe.py [rows ...]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from chatterbox_vllm2_b200 import _lib

L = _lib.lib()
dev = "cuda:0"
rows_list = [int(a) for a in sys.argv[1:] if a.isdigit()] or [2, 4, 8]
SHAPES = [("qkv", 3072, 1024, 0), ("o", 1024, 1024, 0), ("gate_up", 8192, 1024, 1), ("down", 1024, 4096, 0)]
NL = 30
print("rows shape weights kernel us_per_launch GB/s")
for name, N, K, epi in SHAPES:
    ws = [(torch.randn(N, K, device=dev) * K ** -0.5).to(torch.bfloat16) for _ in range(NL)]
    frags = []
    for w in ws:
        f = torch.empty(L.t3b_gemv_frag_bytes(N, K, epi), dtype=torch.uint8, device=dev)
        _lib.check(L.t3b_repack_weights_gemv(w.data_ptr(), N, K, epi, f.data_ptr(), torch.cuda.current_stream().cuda_stream))
        frags.append(f)
    lins = [_lib.Linear(_lib.T3B_W_BF16, N, K, w.data_ptr(), f.data_ptr(), None, 0) for w, f in zip(ws, frags)]
    for rows in rows_list:
        x = torch.randn(rows, K, device=dev).to(torch.bfloat16)
        y = torch.empty(rows, N // 2 if epi else N, dtype=torch.bfloat16, device=dev)
        for same in (0, 1):
            for kern in (("simt", "mma") if rows <= 4 else ("mma",)):
                opt = _lib.make_options(gemv_max_rows=8, gemv_mma=1) if kern == "mma" else _lib.make_options(gemv_max_rows=4, gemv_mma=-1)
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    def run():
                        for li in range(NL):
                            _lib.check(L.t3b_linear_bf16(x.data_ptr(), C.byref(lins[0 if same else li]), rows, y.data_ptr(), None, 1, None,
                                                         _lib.T3B_IMPL_GEMV, epi, C.byref(opt), side.cuda_stream))
                    run(); torch.cuda.synchronize()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=side):
                        run()
                    t = 1e9
                    for rep in range(6):
                        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                        e0.record(side); g.replay(); e1.record(side); torch.cuda.synchronize()
                        if rep: t = min(t, e0.elapsed_time(e1) / NL * 1e3)
                    del g
                torch.cuda.current_stream().wait_stream(side)
                print(f"{rows:3d} {name:8s} {'same' if same else 'distinct':8s} {kern:6s} {t:7.2f} {N * K * 2 / t / 1e3:8.0f}")
