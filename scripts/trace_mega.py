"""GPU debug: globaltimer timeline of decode_mega_kernel (stamps at every grid_arrive / after every grid_wait).

    python scripts/trace_mega.py [rows ...]

Prints, per phase kind, the median time a CTA spends computing (wait-done -> arrive) and waiting at the grid
barrier (arrive -> wait-done), and the wall time of the launch."""
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
L.t3b_debug_set_mega_trace.argtypes = [C.c_void_p]
L.t3b_mega_trace_stride.restype = C.c_int32
dims = T3Dims()
m = T3B200Model(dims=dims, device=DEV, cfg_scale=0.5)
m.load_weights(synthetic_checkpoint(dims).items())
G = torch.cuda.get_device_properties(0).multi_processor_count
S = L.t3b_mega_trace_stride()
KIND = ["qkv", "att", "o", "gate_up", "down"]
for rows in [int(a) for a in sys.argv[1:]] or [2, 64]:
    kv, st = random_state(dims, rows, [250] * rows, seed=1)
    lg = torch.zeros(rows, dims.speech_vocab, dtype=torch.bfloat16, device=DEV)
    tr = torch.zeros(G * S, dtype=torch.int64, device=DEV)
    for it in range(4):
        L.t3b_debug_set_mega_trace(tr.data_ptr() if it == 3 else None)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        m.decode_step_mega(kv, st.c, rows, None, lg)
        e1.record()
        torch.cuda.synchronize()
    L.t3b_debug_set_mega_trace(None)
    t = tr.reshape(G, S).cpu().double()
    # stamps: arrive(E1), then per phase: wait-done, arrive.  151 phases -> 1 + 302 stamps
    nph = dims.n_layers * 5 + 1
    arrive = torch.cat([t[:, 0:1], t[:, 2:2 * nph + 1:2]], 1)      # [G, nph+1]  arrive before phase i (i=0: E1)
    done = t[:, 1:2 * nph:2]                                        # [G, nph]    wait-done of phase i
    comp = arrive[:, 1:] - done                                     # compute of phase i
    wait = done - arrive[:, :-1]                                    # barrier wait before phase i
    span = (arrive[:, 1:].max(0).values - done.min(0).values)      # phase wall: first CTA through -> last arrive
    release = done.min(0).values - arrive[:, :-1].max(0).values    # barrier latency: last arrive -> first release
    print(f"rows {rows}: event {e0.elapsed_time(e1) * 1e3:.1f} us, trace span {(t[:, :2 * nph + 1].max() - t[:, 0].min()) / 1e3:.1f} us")
    print(f"  {'phase':8s} {'comp med':>9s} {'comp max':>9s} {'wait med':>9s} {'span':>9s} {'barrier':>9s}   (us, mean over layers 2..29)")
    for k in range(5):
        idx = torch.arange(2, dims.n_layers) * 5 + k
        print(f"  {KIND[k]:8s} {comp[:, idx].median(0).values.mean() / 1e3:9.2f} {comp[:, idx].max(0).values.mean() / 1e3:9.2f} "
              f"{wait[:, idx].median(0).values.mean() / 1e3:9.2f} {span[idx].mean() / 1e3:9.2f} {release[idx].mean() / 1e3:9.2f}")
    print(f"  {'head':8s} {comp[:, -1].median() / 1e3:9.2f} {comp[:, -1].max() / 1e3:9.2f} {wait[:, -1].median() / 1e3:9.2f} "
          f"{span[-1] / 1e3:9.2f} {release[-1] / 1e3:9.2f}")
