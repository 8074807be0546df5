#!/usr/bin/env python
"""Timeline of the decode attention kernel inside a PDL chain (profiling build of the library, -DT3B_ATTN_TRACE):
    python scripts/probe/attn_trace.py [rows] [context]
builds scripts/probe/bin/libt3b200_trace.so on first use (nvcc), launches qkv-GEMM -> attention pairs back to back like a decode step
and prints, for the CTAs of one attention launch, the %globaltimer stamps relative to the earliest CTA start."""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
CS = os.path.join(ROOT, "chatterbox-vllm2_b200", "csrc")
OUT = os.path.join(ROOT, "scripts", "probe", "bin", "libt3b200_trace.so")


def build():
    srcs = ["api.cu", "elementwise.cu", "attention.cu", "linear_gemv.cu", "linear_umma.cu", "linear_big.cu", "sampler.cu"]
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    subprocess.check_call(["nvcc", "-shared", "-o", OUT, "-DT3B_ATTN_TRACE", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17",
                           "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-cudart", "static"] + [os.path.join(CS, s) for s in srcs])


def main():
    if "--build" in sys.argv:
        build()
        return
    import torch
    from chatterbox_vllm2_b200 import T3Dims, _lib
    _lib.LIB_PATH = OUT
    L = _lib.lib()
    L.t3b_debug_set_attn_trace.argtypes = [C.c_void_p]
    R = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    ctx = int(sys.argv[2]) if len(sys.argv) > 2 else 249
    d, dev = T3Dims(), "cuda:0"
    nblk = (ctx + 16) // 16
    mb = 75
    kp = torch.randn(R * nblk + 1, 16, 1024, device=dev).to(torch.bfloat16)
    vp = torch.randn_like(kp.float()).to(torch.bfloat16)
    bt = torch.zeros(R, mb, dtype=torch.int32)
    bt[:, :nblk] = (torch.randperm(R * nblk) + 1).reshape(R, nblk).to(torch.int32)
    bt = bt.to(dev)
    idx = torch.full((R,), ctx - 1, dtype=torch.int32, device=dev)
    x = torch.randn(R, 1024, device=dev).to(torch.bfloat16)
    w = (torch.randn(3072, 1024, device=dev) / 32).to(torch.bfloat16)
    lin = _lib.Linear(_lib.T3B_W_BF16, 3072, 1024, w.data_ptr(), None, None)
    qkv = torch.empty(R, 3072, dtype=torch.bfloat16, device=dev)
    out = torch.empty(R, 1024, dtype=torch.bfloat16, device=dev)
    from chatterbox_vllm2_b200.model import llama3_rope_table
    rope = llama3_rope_table(d).to(device=dev, dtype=torch.bfloat16).contiguous()
    trace = torch.zeros(4, 148, 16, dtype=torch.int64, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    only_attn = "--attn-only" in sys.argv
    for rep in range(4):
        torch.cuda.synchronize()
        L.t3b_debug_set_attn_trace(trace[rep].data_ptr())
        for _ in range(3):
            if not only_attn:
                _lib.check(L.t3b_linear_bf16(x.data_ptr(), C.byref(lin), R, qkv.data_ptr(), None, 0, None, 0, 0, None, s))
            _lib.check(L.t3b_attn_decode_fused(qkv.data_ptr(), R, idx.data_ptr(), idx.data_ptr(), rope.data_ptr(), d.max_pos, bt.data_ptr(), mb,
                                               kp.data_ptr(), vp.data_ptr(), 16, 0, out.data_ptr(), None, s))
        torch.cuda.synchronize()
    t = trace[3].cpu().double()
    t = t[t[:, 0] > 0]
    t0 = t[:, 0].min()
    names = ["start", "meta_done", "head_done", "wait_done", "first_pair", "q_arrived", "item0_done", "end", "q_rotated", "k_rotated", "hblk_arrived", "new_token_dot"]
    print(f"rows {R} context {ctx}: {t.shape[0]} CTAs of the last launch, ns after the earliest CTA start (min / median / max)")
    for i, n in enumerate(names):
        c = t[:, i] - t0
        print(f"  {n:12s} {c.min():9.0f} {c.median():9.0f} {c.max():9.0f}")


if __name__ == "__main__":
    main()
