"""GPU debug: clock64 timeline inside linear_umma_kernel when it runs WARM inside a PDL chain: 30 launches of one decode shape
(distinct weights per launch, like the 30 layers) captured in a CUDA graph; the trace buffer keeps the last launch's stamps."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm2_b200 import _lib
L = _lib.lib()
DEV = "cuda:0"
L.t3b_debug_set_umma_trace.argtypes = [C.c_void_p]
names = ["start", "setup", "pdlwait", "mma0", "mmaLast", "tmemfull", "parked", "csync1", "reduced", "end", "w4", "w5", "w6", "w7", "t0", "t32"]
rows = int(os.environ.get("ROWS", "64"))
NL = 30
for (N, K, epi, nm) in [(3072, 1024, 0, "qkv"), (1024, 1024, 0, "o"), (8192, 1024, 1, "gate_up"), (1024, 4096, 0, "down")]:
    x = torch.randn(rows, K, device=DEV).bfloat16()
    ws = [(torch.randn(N, K, device=DEV) * K ** -0.5).bfloat16() for _ in range(NL)]
    y = torch.empty(rows, N, dtype=torch.bfloat16, device=DEV)
    lins = [_lib.Linear(0, N, K, w.data_ptr(), None, None) for w in ws]
    tr = torch.zeros(4096 * 16, dtype=torch.int64, device=DEV)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        s = side.cuda_stream
        def run():
            for l in lins:
                _lib.check(L.t3b_linear_bf16(x.data_ptr(), C.byref(l), rows, y.data_ptr(), None, 0, None, 2, epi, s))
        run(); torch.cuda.synchronize()
        L.t3b_debug_set_umma_trace(tr.data_ptr())
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            run()
        L.t3b_debug_set_umma_trace(None)
        for _ in range(3):
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(side); g.replay(); e1.record(side)
            torch.cuda.synchronize()
    t = tr.reshape(4096, 16).cpu()
    t = t[t[:, 0] != 0]
    rel = (t[:, :16] - t[:, :1]).float()
    print(f"{nm}: rows {rows} N {N} K {K} ctas {t.shape[0]}  {e0.elapsed_time(e1) * 1e3 / NL:.2f} us per launch in the chain (trace on)")
    print("   " + " ".join(f"{n:>8s}" for n in names[:10]))
    print("med" + " ".join(f"{v:8.0f}" for v in rel.median(0).values.tolist()[:10]))
    print("max" + " ".join(f"{v:8.0f}" for v in rel.max(0).values.tolist()[:10]))
    # spread of CTA start times relative to the earliest CTA of the launch
    st = (t[:, 0] - t[:, 0].min()).float()
    print(f"   CTA start spread: median {st.median():.0f} max {st.max():.0f} cycles (clock64 is per SM: indicative only)")
