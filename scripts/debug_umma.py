"""GPU debug: where does the tcgen05 linear differ from fp32/fp64 references?"""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm2_b200 import _lib
L = _lib.lib()
DEV = "cuda:0"
torch.backends.cuda.matmul.allow_tf32 = False
for (rows, N, K) in [(64, 3072, 1024), (128, 3072, 1024), (50, 8192, 1024), (256, 1024, 4096), (512, 8192, 1024)]:
    g = torch.Generator().manual_seed(N + K + rows)
    x = torch.randn(rows, K, generator=g).bfloat16()
    w = (torch.randn(N, K, generator=g) * K ** -0.5).bfloat16()
    xd, wd = x.to(DEV), w.to(DEV)
    torch.cuda.synchronize()
    y = torch.full((rows, N), float("nan"), dtype=torch.bfloat16, device=DEV)
    lin = _lib.Linear(0, N, K, wd.data_ptr(), None, None)
    _lib.check(L.t3b_linear_bf16(xd.data_ptr(), C.byref(lin), rows, y.data_ptr(), None, 1, None, 2, 0, torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    yc = y.float().cpu()
    r32 = (x.float() @ w.float().t())
    r64 = (x.double() @ w.double().t()).float()
    rg = (xd.float() @ wd.float().t()).cpu()
    def stat(a, b):
        d = (a - b).abs()
        return f"max {d.max():.4g} mean {d.mean():.4g} cos {torch.nn.functional.cosine_similarity(a.flatten().double(), b.flatten().double(), dim=0):.8f}"
    print(f"rows {rows} N {N} K {K}: nan {int(torch.isnan(yc).sum())}")
    print("   ours vs cpu32 :", stat(yc, r32.bfloat16().float()))
    print("   ours vs cpu64 :", stat(yc, r64.bfloat16().float()))
    print("   ours vs gpu32 :", stat(yc, rg.bfloat16().float()))
    print("   cpu32 vs cpu64:", stat(r32, r64))
    d = (yc - r64.bfloat16().float()).abs()
    bad = (d > 0.05).nonzero()
    print("   bad count", len(bad), "first", bad[:8].tolist(), "rows set", sorted(set(bad[:, 0].tolist()))[:20], "cols%128", sorted(set((bad[:, 1] % 128).tolist()))[:20])
