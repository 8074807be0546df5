"""GPU debug: clock64 timeline inside linear_umma_kernel for the decode shapes, plain and with F2.1 fused
(t3b_linear_fused: norm operand build / residual epilogue).  Cycles relative to the CTA's first stamp."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm2_b200 import _lib
L = _lib.lib()
DEV = "cuda:0"
L.t3b_debug_set_umma_trace.argtypes = [C.c_void_p]
names = ["start", "setup", "pdlwait", "mma0", "mmaLast", "tmemfull", "parked", "csync1", "reduced", "end", "w4", "w5", "w6", "w7", "t0", "t32"]
rows = int(os.environ.get("ROWS", "64"))
junk = torch.ones(512 << 20, dtype=torch.uint8, device=DEV)
for (N, K, epi, mode) in [(8192, 1024, 1, "plain"), (8192, 1024, 1, "normx"), (3072, 1024, 0, "plain"), (3072, 1024, 0, "normx"),
                          (1024, 4096, 0, "plain"), (1024, 4096, 2, "resid"), (1024, 1024, 0, "plain"), (1024, 1024, 2, "resid")]:
    x = torch.randn(rows, K, device=DEV).bfloat16()
    w = (torch.randn(N, K, device=DEV) * K ** -0.5).bfloat16()
    y = torch.empty(rows, N, dtype=torch.bfloat16, device=DEV)
    z = torch.randn(rows, max(N, K), device=DEV)
    zk = torch.randn(rows, K, device=DEV)
    ss = torch.rand(32, rows, device=DEV) * 128
    nw = torch.ones(K, device=DEV).bfloat16()
    lin = _lib.Linear(0, N, K, w.data_ptr(), None, None)
    tr = torch.zeros(4096 * 16, dtype=torch.int64, device=DEV)
    s = torch.cuda.current_stream().cuda_stream
    ns = L.t3b_umma_pick_split(rows, N, K) if hasattr(L, "t3b_umma_pick_split") else 0
    for it in range(3):
        L.t3b_debug_set_umma_trace(tr.data_ptr() if it == 2 else None)
        junk.sum()                      # read pass: L2 ends up full of CLEAN lines (a write pass leaves it dirty)
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        if mode == "plain":
            _lib.check(L.t3b_linear_bf16(x.data_ptr(), C.byref(lin), rows, y.data_ptr(), None, 0, None, 2, epi, s))
        elif mode == "normx":
            f = _lib.LinearFuse(zk.data_ptr(), ss.data_ptr(), K // 128, nw.data_ptr(), 1e-5, None, None, None)
            _lib.check(L.t3b_linear_fused(None, C.byref(lin), rows, y.data_ptr(), C.byref(f), 0, epi, s))
        else:
            f = _lib.LinearFuse(None, None, 0, None, 0.0, z.data_ptr(), z.data_ptr(), ss.data_ptr())
            _lib.check(L.t3b_linear_fused(x.data_ptr(), C.byref(lin), rows, None, C.byref(f), 0, epi, s))
        e1.record()
        torch.cuda.synchronize()
    L.t3b_debug_set_umma_trace(None)
    t = tr.reshape(4096, 16).cpu()
    t = t[t[:, 0] != 0]
    rel = (t[:, :16] - t[:, :1]).float()
    print(f"rows {rows} N {N} K {K} epi {epi} {mode} ctas {t.shape[0]}  event {e0.elapsed_time(e1)*1e3:.1f} us")
    print("   " + " ".join(f"{n:>9s}" for n in names))
    print("med" + " ".join(f"{v:9.0f}" for v in rel.median(0).values.tolist()))
    print("max" + " ".join(f"{v:9.0f}" for v in rel.max(0).values.tolist()))
