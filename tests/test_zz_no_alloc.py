"""t3b_head_cfg_sample_static (include/t3b200.h): the allocation-free variant of t3b_head_cfg_sample - head logits scratch in a
file-scope ``static __device__`` array inside the library instead of a caller workspace.  No counterpart in the reference; added
for callers that cannot call cudaMalloc.

Named test_zz_* so that pytest collects it LAST (a failure here must not hide other tests).
Passes alone in a fresh process and after the other test files (it builds its own model and touches no shared state).

Checked, through the C ABI, on two calls in a row (64 rows, then 6 rows with other inputs - the array still holds the first call's
logits): tokens and combined fp32 logits equal those of the existing entry point bit for bit (same kernels, same launch shapes,
deterministic; the tolerance of the other parity tests of this pair of kernels is 0 as well), sampled and greedy; the second call
leaves the device's free memory untouched; sizes beyond the array are rejected with a message naming the limit.
"""
import ctypes as C

import pytest
import torch

from chatterbox_vllm2_b200 import T3Dims, _lib
from chatterbox_vllm2_b200.model import DecodeState, SamplingTensors, T3B200Model
from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SMALL = T3Dims().small(n_layers=2, max_pos=512)
STATIC_MAX_ROWS = 64          # T3B_STATIC_MAX_ROWS


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _inputs(m, rows, seed):
    """hidden bf16 [rows, D], a decode state with some history (so that the repetition penalty bites) and sampling knobs."""
    g = torch.Generator().manual_seed(seed)
    n = rows // 2
    hidden = torch.randn(rows, SMALL.hidden, generator=g).bfloat16().to(DEV)
    st = DecodeState(n, 4, 64, 64, DEV)
    hist = torch.randint(0, SMALL.speech_vocab, (n, 10), generator=g).to(torch.int32)
    st.history[:n, :10] = hist.to(DEV)
    st.hist_len[:n] = 10
    st.step[:n] = 10
    st.active[:n] = 1
    sp = SamplingTensors(n, DEV, m.combine)
    sp.set_many([dict(cfg_scale=0.5, repetition_penalty=2.0, temperature=0.8 if i % 2 else 0.0, min_p=0.05, top_p=0.8, seed=100 + i)
                 for i in range(n)])
    return hidden, st, sp


def _both(L, m, rows, seed):
    hidden, st, sp = _inputs(m, rows, seed)
    n, V = rows // 2, SMALL.speech_vocab
    ws = m.workspace(rows)
    tok_a = torch.full((n,), -1, dtype=torch.int32, device=DEV)
    tok_b = torch.full((n,), -2, dtype=torch.int32, device=DEV)
    log_a = torch.zeros(n, V, dtype=torch.float32, device=DEV)
    log_b = torch.ones(n, V, dtype=torch.float32, device=DEV)
    _lib.check(L.t3b_head_cfg_sample(C.byref(m.c), hidden.data_ptr(), rows, C.byref(sp.c), C.byref(st.c), 0, tok_a.data_ptr(),
                                     log_a.data_ptr(), ws.data_ptr(), ws.numel(), None, _stream()), "t3b_head_cfg_sample")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    _lib.check(L.t3b_head_cfg_sample_static(C.byref(m.c), hidden.data_ptr(), rows, C.byref(sp.c), C.byref(st.c), 0, tok_b.data_ptr(),
                                            log_b.data_ptr(), None, _stream()), "t3b_head_cfg_sample_static")
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return (tok_a, log_a), (tok_b, log_b), free0 - free1


def test_static_scratch_variant_equals_workspace_variant_on_two_calls_in_a_row():
    L = _lib.lib()
    _lib.check(L.t3b_check_device(), "t3b_check_device")
    m = T3B200Model(dims=SMALL, device=DEV, cfg_scale=0.5)
    m.load_weights(synthetic_checkpoint(SMALL).items())
    used = []
    for call, (rows, seed) in enumerate([(STATIC_MAX_ROWS, 1), (6, 2)]):
        (tok_a, log_a), (tok_b, log_b), grew = _both(L, m, rows, seed)
        assert torch.equal(tok_a, tok_b), (call, tok_a.tolist(), tok_b.tolist())
        assert torch.equal(log_a, log_b), (call, (log_a - log_b).abs().max().item())
        assert (tok_a >= 0).all() and (tok_a < SMALL.speech_vocab).all()
        used.append(grew)
    # the first call on a device may load the module's data lazily; the second must not take a byte from the device
    assert used[1] <= 0, used

    # beyond the array: a clear error, nothing launched
    rows = STATIC_MAX_ROWS + 2
    hidden, st, sp = _inputs(m, rows, 3)
    tok = torch.zeros(rows // 2, dtype=torch.int32, device=DEV)
    rc = L.t3b_head_cfg_sample_static(C.byref(m.c), hidden.data_ptr(), rows, C.byref(sp.c), C.byref(st.c), 0, tok.data_ptr(), None, None,
                                      _stream())
    assert rc == -1                      # T3B_ERR_ARG
    msg = L.t3b_last_error().decode()
    assert "T3B_STATIC_MAX_ROWS" in msg and str(rows) in msg, msg
    torch.cuda.synchronize()
