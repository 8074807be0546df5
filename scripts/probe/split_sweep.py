#!/usr/bin/env python
"""Time every cluster split-K factor of the four decode projections at several row counts (CUDA graph of 30 launches over 30
distinct weight sets, like bench.py's per-class breakdown): the table t3b_umma_pick_split encodes.
    python scripts/probe/split_sweep.py [rows ...]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from chatterbox_vllm2_b200 import _lib

L = _lib.lib()
dev = "cuda:0"
rows_list = [int(a) for a in sys.argv[1:] if a.isdigit()] or [64, 128, 256, 512, 1024]
awq = "--awq" in sys.argv
SHAPES = [("qkv", 3072, 1024, 0), ("o", 1024, 1024, 0), ("gate_up", 8192, 1024, 1), ("down", 1024, 4096, 0)]
NL = 30
print("rows shape split chunk us_per_launch")
for name, N, K, epi in SHAPES:
    ws = [(torch.randn(N, K, device=dev) * K ** -0.5).to(torch.bfloat16) for _ in range(NL)]
    lins = [_lib.Linear(_lib.T3B_W_BF16, N, K, w.data_ptr(), None, None, 0) for w in ws]
    for rows in rows_list:
        x = torch.randn(rows, K, device=dev).to(torch.bfloat16)
        y = torch.empty(rows, N // 2 if epi else N, dtype=torch.bfloat16, device=dev)
        best = None
        for split in (1, 2, 4, 8):
            for chunk, stages, big in (((0, 0, -1),) if split > 1 else ((0, 0, -1), (0, 0, 1))):
                opt = _lib.make_options(umma_row_chunk=chunk, umma_stages=stages, umma_big_min_rows=big)
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                try:
                    with torch.cuda.stream(side):
                        def run():
                            for li in range(NL):
                                _lib.check(L.t3b_linear_bf16(x.data_ptr(), C.byref(lins[li]), rows, y.data_ptr(), None, split, None, _lib.T3B_IMPL_UMMA, epi,
                                                             C.byref(opt), side.cuda_stream))
                        run(); torch.cuda.synchronize()
                        g = torch.cuda.CUDAGraph()
                        with torch.cuda.graph(g, stream=side):
                            run()
                        t = 1e9
                        for rep in range(5):
                            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                            e0.record(side); g.replay(); e1.record(side); torch.cuda.synchronize()
                            if rep: t = min(t, e0.elapsed_time(e1) / NL * 1e3)
                        del g
                except Exception as e:  # noqa: BLE001
                    print(f"{rows:5d} {name:8s} {split} {chunk:3d} st{stages} big{big} failed: {str(e)[:80]}")
                    continue
                torch.cuda.current_stream().wait_stream(side)
                print(f"{rows:5d} {name:8s} {split} {chunk:3d} st{stages} big{big} {t:8.2f}")
                if best is None or t < best[0]:
                    best = (t, split, chunk)
        print(f"# best rows {rows} {name}: split {best[1]} chunk {best[2]} {best[0]:.2f} us")
