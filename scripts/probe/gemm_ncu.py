#!/usr/bin/env python
"""The decode / prefill projections at large row counts, a few launches each, for an ncu --set full capture (tensor-pipe evidence):
    ncu --set full --clock-control none -k regex:"linear_umma|linear_big" -o out python scripts/probe/gemm_ncu.py"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from chatterbox_vllm2_b200 import _lib

L = _lib.lib()
dev = "cuda:0"
s = torch.cuda.current_stream().cuda_stream
for rows in (256, 512, 1024, 6336):
    for name, N, K, epi, split in (("gate_up", 8192, 1024, 1, 1), ("qkv", 3072, 1024, 0, 1), ("down", 1024, 4096, 0, 0 if rows <= 1024 else 1)):
        w = (torch.randn(N, K, device=dev) * K ** -0.5).to(torch.bfloat16)
        lin = _lib.Linear(_lib.T3B_W_BF16, N, K, w.data_ptr(), None, None, 0)
        x = torch.randn(rows, K, device=dev).to(torch.bfloat16)
        y = torch.empty(rows, N // 2 if epi else N, dtype=torch.bfloat16, device=dev)
        for _ in range(2):
            _lib.check(L.t3b_linear_bf16(x.data_ptr(), C.byref(lin), rows, y.data_ptr(), None, split, None, _lib.T3B_IMPL_UMMA, epi, None, s))
        torch.cuda.synchronize()
        print(rows, name, "ok")
