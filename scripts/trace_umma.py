"""GPU debug: clock64 timeline inside linear_umma_kernel for the decode shapes."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm2_b200 import _lib
L = _lib.lib()
DEV = "cuda:0"
L.t3b_debug_set_umma_trace.argtypes = [C.c_void_p]
names = ["start", "setup", "pdlwait", "mma0", "mmaLast", "tmemfull", "parked", "csync1", "reduced", "end", "w4", "w5", "w6", "w7", "t0", "t32"]
for (rows, N, K, epi) in [(64, 8192, 1024, 1), (64, 3072, 1024, 0), (64, 1024, 4096, 0), (64, 1024, 1024, 0)]:
    x = torch.randn(rows, K, device=DEV).bfloat16()
    w = (torch.randn(N, K, device=DEV) * K ** -0.5).bfloat16()
    y = torch.empty(rows, N, dtype=torch.bfloat16, device=DEV)
    lin = _lib.Linear(0, N, K, w.data_ptr(), None, None)
    tr = torch.zeros(4096 * 16, dtype=torch.int64, device=DEV)
    junk = torch.ones(512 << 20, dtype=torch.uint8, device=DEV)
    ns = C.c_int32(0)
    s = torch.cuda.current_stream().cuda_stream
    for it in range(3):
        L.t3b_debug_set_umma_trace(tr.data_ptr() if it == 2 else None)
        # flush L2 so the weights come from HBM
        junk.sum()                      # read pass: L2 ends up full of CLEAN lines (a write pass leaves it dirty)
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(L.t3b_linear_bf16(x.data_ptr(), C.byref(lin), rows, y.data_ptr(), None, 0, C.byref(ns), 2, epi, s))
        e1.record()
        torch.cuda.synchronize()
    L.t3b_debug_set_umma_trace(None)
    ncta = ((N // 128) if epi else (N + 127) // 128) * ns.value
    t = tr[: ncta * 16].reshape(ncta, 16).cpu()
    rel = (t[:, :16] - t[:, :1]).float()
    print(f"rows {rows} N {N} K {K} epi {epi} split {ns.value} ctas {ncta}  event {e0.elapsed_time(e1)*1e3:.1f} us")
    print("   " + " ".join(f"{n:>9s}" for n in names))
    print("med" + " ".join(f"{v:9.0f}" for v in rel.median(0).values.tolist()))
    print("max" + " ".join(f"{v:9.0f}" for v in rel.max(0).values.tolist()))
    print("min" + " ".join(f"{v:9.0f}" for v in rel.min(0).values.tolist()))
