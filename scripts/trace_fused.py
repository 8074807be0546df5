"""GPU debug: globaltimer timeline of decode_fused_kernel.

    python scripts/trace_fused.py [rows ...]

Stamps per CTA (worker thread 0):  exchange: [wait-done (e>0)] arrive wait-done | phase A: after cluster sync 1, 2, 3,
grid arrive | exchange ... | phase B: after cluster sync 4, grid arrive | ... | final exchange | end."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from chatterbox_vllm2_b200 import T3Dims, _lib
from chatterbox_vllm2_b200.model import T3B200Model
from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint
from test_gpu_mega import random_state

L = _lib.lib()
DEV = "cuda:0"
L.t3b_debug_set_fused_trace.argtypes = [C.c_void_p]
L.t3b_fused_trace_stride.restype = C.c_int32
dims = T3Dims()
m = T3B200Model(dims=dims, device=DEV, cfg_scale=0.5)
m.load_weights(synthetic_checkpoint(dims).items())
G, S = 128, L.t3b_fused_trace_stride()
ctx = int(os.environ.get("CTX", "250"))
L.t3b_debug_set_fused_flags.argtypes = [C.c_int32]
L.t3b_debug_set_fused_flags(int(os.environ.get("FZ_DBG", "0")))
for rows in [int(a) for a in sys.argv[1:]] or [2, 64]:
    kv, st = random_state(dims, rows, [ctx] * rows, seed=1)
    lg = torch.zeros(rows, dims.speech_vocab, dtype=torch.bfloat16, device=DEV)
    tr = torch.zeros(2 * G * S, dtype=torch.int64, device=DEV)
    best = 1e9
    for it in range(6):
        L.t3b_debug_set_fused_trace(tr.data_ptr() if it == 5 else None)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        m.decode_step_fused(kv, st.c, rows, None, lg)
        e1.record()
        torch.cuda.synchronize()
        if it < 5:
            best = min(best, e0.elapsed_time(e1))
    L.t3b_debug_set_fused_trace(None)
    tm2 = tr[G * S:].reshape(G, S).cpu()
    t = tr[:G * S].reshape(G, S).cpu()
    ids, tm = t[:, 0::2], t[:, 1::2].double()
    # sequence of one layer (>= 1): 2 1 2 | 10 11 12 13 14 15 16 17 3 1 | 2 1 2 | 20 21 22 23 3 1
    seq = [2, 1, 2, 10, 11, 12, 13, 14, 15, 16, 17, 3, 1, 2, 1, 2, 20, 21, 22, 23, 3, 1]
    # interval ENDING at each stamp
    names = ["grid barrier (end B)", "xA reduce", "xA barrier", "build_x", "qkv mma wait", "qkv drain+bcast", "cluster sync1",
             "attention", "cluster sync2", "combine", "cluster sync3", "o mma wait", "o drain+store", "grid barrier (end A)", "xB reduce",
             "xB barrier", "build_x", "gu mma wait", "gu drain+act", "cluster sync4", "down mma wait", "down drain+store"]
    per = len(seq)
    first = per - 1                       # layer 0 lacks the leading wait stamp
    assert ids[0, first:first + per].tolist() == seq, ids[0, :60].tolist()
    seg = torch.zeros(per, G)
    nl = 0
    for layer in range(2, dims.n_layers):
        o = first + (layer - 1) * per
        st_ = tm[:, o - 1:o + per]
        seg += (st_[:, 1:] - st_[:, :-1]).T
        nl += 1
    seg /= nl
    print(f"rows {rows} ctx {ctx}: best event {best * 1e3:.1f} us; per layer (us): median over CTAs / max over CTAs")
    for i, n in enumerate(names):
        print(f"  {n:22s} {seg[i].median() / 1e3:7.2f} {seg[i].max() / 1e3:7.2f}")
    print(f"  {'layer total':18s} {seg.sum(0).median() / 1e3:7.2f}")
    # MMA-issuer side: per full-K GEMM stamps 40 (begin) 41 (first weights) 42 (first X) 44 (last weights) 43 (all issued)
    mi, mt = tm2[:, 0::2], tm2[:, 1::2].double()
    n = int((mi[0] != 0).sum())
    assert n % 5 == 0
    d = (mt[:, 1:n] - mt[:, :n - 1])
    lab = ["wait first weights", "wait first X", "wait last weights", "issue rest"]
    for gi, gname in ((0, "qkv"), (1, "gate_up")):
        sel = [5 * (2 * l + gi) for l in range(2, dims.n_layers)]
        print(f"  MMA {gname}: " + ", ".join(f"{lab[k]} {torch.stack([d[:, o + k] for o in sel]).mean(0).median() / 1e3:.2f}" for k in range(4)))
