"""debug: NaN rows of the tensor-core GEMV chain at 9..16 rows, with and without the fused normalisation"""
import ctypes as C, sys, os, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from chatterbox_vllm2_b200 import _lib, T3Dims
from chatterbox_vllm2_b200.model import T3B200Model
from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint
from helpers import random_state
L = _lib.lib(); DEV = "cuda:0"
SMALL = T3Dims().small(n_layers=2, max_pos=512)
sd = synthetic_checkpoint(SMALL)
m = T3B200Model(dims=SMALL, device=DEV, cfg_scale=0.5); m.load_weights(sd.items()); assert m.enable_gemv_mma()
sp = lambda: torch.cuda.current_stream().cuda_stream
for rows in (8, 9, 10, 12, 16):
    for no_norm in (0, 1):
        g = torch.Generator().manual_seed(rows)
        lens = torch.randint(5, 300, ((rows + 1) // 2,), generator=g).repeat_interleave(2).tolist()[:rows]
        kv, st = random_state(SMALL, rows + (rows & 1), lens + lens[-1:] * (rows & 1), seed=5)
        opt = _lib.make_options(gemv_max_rows=16, gemv_mma=1, gemv_no_norm=no_norm)
        ws = m.workspace(16)
        hid = torch.zeros(rows, 1024, dtype=torch.bfloat16, device=DEV)
        rc = L.t3b_decode_step(C.byref(m.c), C.byref(kv.c), C.byref(st.c), None, rows, hid.data_ptr(), ws.data_ptr(), ws.numel(), C.byref(opt), sp())
        torch.cuda.synchronize()
        bad = torch.isnan(hid.float()).any(dim=1).nonzero().flatten().tolist()
        print("rows", rows, "no_norm", no_norm, "rc", rc, "nan rows", bad, "absmax", float(hid.float().nan_to_num().abs().max()))
