"""AWQ W4A16 path on the GPU: format pin against vLLM's own dequantiser, the
dequant-fused tcgen05 linear against the oracle, and a short end-to-end decode with a quantised backbone."""
import ctypes as C

import numpy as np
import pytest
import torch

from chatterbox_vllm2_b200 import SPEECH_TOKEN_OFFSET, T3Dims, _lib, awq
from chatterbox_vllm2_b200.engine import SamplingParams, T3B200Engine
from chatterbox_vllm2_b200.model import T3B200Model
from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint, synthetic_request
from helpers import first_divergence, oracle_generate, record
from oracle import awq_ref
from oracle.t3_ref import T3Oracle

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SMALL = T3Dims().small(n_layers=2, max_pos=512)


def sp():
    return torch.cuda.current_stream().cuda_stream


def test_oracle_dequantize_matches_vllm_triton_kernel():
    """Pins the restated AutoAWQ layout (nibble order, zero/scale broadcasting) on the engine the reference runs."""
    vq = pytest.importorskip("vllm.model_executor.layers.quantization.awq_triton")
    g = torch.Generator().manual_seed(0)
    w = torch.randn(256, 512, generator=g) * 0.05
    q, z, s = awq_ref.quantize(w)
    qw, qz = awq_ref.pack(q), awq_ref.pack(z)
    ref = vq.awq_dequantize_triton(torch.from_numpy(qw).to(DEV), torch.from_numpy(s).to(DEV), torch.from_numpy(qz).to(DEV))
    mine = awq_ref.dequantize_fp16(qw, qz, s)
    assert ref.shape == mine.shape == (512, 256)
    assert torch.equal(ref.cpu(), mine)


def test_package_quantizer_equals_oracle_quantizer():
    g = torch.Generator().manual_seed(1)
    w = (torch.randn(384, 256, generator=g) * 0.04).bfloat16()
    t = awq.quantize_linear(w)
    q, z, s = awq_ref.quantize(w)
    assert np.array_equal(t["qweight"].numpy(), awq_ref.pack(q)) and np.array_equal(t["qzeros"].numpy(), awq_ref.pack(z))
    assert np.array_equal(t["scales"].numpy(), s)


def _awq_linear(L, x, w, rows, gate_up=False, split=0):
    q = awq.quantize_and_repack(w.to(DEV), gate_up)
    N, K = w.shape
    lin = _lib.Linear(_lib.T3B_W_AWQ, N, K, q["w"].data_ptr(), q["scales"].data_ptr(), q["zeros"].data_ptr(),
                      _lib.T3B_LIN_GATE_UP_INTERLEAVED if gate_up else 0)
    xd = x.to(DEV)
    y = torch.full((rows, N // 2 if gate_up else N), float("nan"), dtype=torch.bfloat16, device=DEV)
    _lib.check(L.t3b_linear_awq(xd.data_ptr(), C.byref(lin), rows, y.data_ptr(), None, split, None, _lib.T3B_IMPL_AUTO,
                                _lib.T3B_EPI_SILU_MUL if gate_up else _lib.T3B_EPI_NONE, None, sp()), "t3b_linear_awq")
    torch.cuda.synchronize()
    return y.float().cpu()


@pytest.mark.parametrize("rows", [1, 2, 16, 64, 200, 512])
@pytest.mark.parametrize("N,K,split", [(3072, 1024, 0), (1024, 4096, 0), (1024, 4096, 1), (1024, 1024, 8)])
def test_awq_linear_matches_oracle(rows, N, K, split):
    L = _lib.lib()
    g = torch.Generator().manual_seed(N + K + rows)
    x = torch.randn(rows, K, generator=g).bfloat16()
    w = (torch.randn(N, K, generator=g) * K ** -0.5).bfloat16()
    y = _awq_linear(L, x, w, rows, split=split)
    q, z, s = awq_ref.quantize(w)
    wd = awq_ref.dequantize_build(awq_ref.pack(q), awq_ref.pack(z), s)            # [N, K] fp32 holding bf16 values
    ref = (x.float() @ wd.t()).bfloat16().float()
    mag = torch.maximum(y.abs(), ref.abs()).clamp_min(1e-30)
    ulp = (y - ref).abs() / torch.exp2(torch.floor(torch.log2(mag)) - 7)
    ok = (ulp <= 1.0) | ((ref.abs() < 1e-2) & ((y - ref).abs() < 2e-4))
    assert torch.isfinite(y).all() and ok.float().mean() > 0.999, (float(ulp.max()), float((~ok).float().mean()))
    assert torch.nn.functional.cosine_similarity(y.double().flatten(), ref.double().flatten(), dim=0) > 0.99999
    # and the quantisation itself is sane: close to the unquantised product
    full = x.float() @ w.float().t()
    assert torch.nn.functional.cosine_similarity(y.double().flatten(), full.double().flatten(), dim=0) > 0.99


@pytest.mark.parametrize("rows", [2, 64, 300])
def test_awq_gate_up_silu(rows):
    L = _lib.lib()
    g = torch.Generator().manual_seed(rows)
    K, F = 1024, 4096
    x = torch.randn(rows, K, generator=g).bfloat16()
    w = (torch.randn(2 * F, K, generator=g) * K ** -0.5).bfloat16()
    act = _awq_linear(L, x, w, rows, gate_up=True)
    q, z, s = awq_ref.quantize(w)
    wd = awq_ref.dequantize_build(awq_ref.pack(q), awq_ref.pack(z), s)
    gu = (x.float() @ wd.t()).bfloat16()
    ref = (torch.nn.functional.silu(gu[:, :F]) * gu[:, F:]).float()
    assert torch.nn.functional.cosine_similarity(act.double().flatten(), ref.double().flatten(), dim=0) > 0.99999
    assert (act - ref).abs().max() <= 0.0625


def test_awq_decode_matches_oracle_with_dequantised_weights():
    """quantization="awq" end to end: greedy decode of a 2-layer T3 whose projections are W4A16 vs the bf16-mode oracle
    run on the dequantised weights (the definition of the AWQ linear: y = x W_deq^T)."""
    sd = synthetic_checkpoint(SMALL)
    sd_deq, _ = awq_ref.quantize_state_dict(sd)
    m = T3B200Model(dims=SMALL, device=DEV, cfg_scale=0.5, quantization="awq")
    m.load_weights(sd.items())
    assert m.weight_bytes() < 0.45 * sum(v.numel() * 2 for k, v in sd.items() if k.startswith("tfmr.") or k == "speech_head.weight")
    prompts = [synthetic_request(i, SMALL, n_text=12) for i in range(3)]
    T = 12
    ref = oracle_generate(T3Oracle(sd_deq, SMALL, mode="bf16"), prompts, T, rho=2.0)
    eng = T3B200Engine(m, max_num_seqs=4, max_model_len=256)
    outs = eng.generate([{"prompt_token_ids": ids, "multi_modal_data": {"conditionals": [c]}} for c, ids in prompts],
                        SamplingParams(temperature=0.0, repetition_penalty=2.0, max_tokens=T, ignore_eos=True))
    for b, o in enumerate(outs):
        ids = [t - SPEECH_TOKEN_OFFSET for t in o.outputs[0].token_ids]
        k = first_divergence(ids, ref["tokens"][b])
        record("awq_engine_greedy_stream", {"req": b, "first_divergence": k, "steps": T,
                                            "margin_at_divergence": float(ref["margins"][k, b]) if k < T else None})
        assert k == T or ref["margins"][k, b] <= 0.125, (b, k, float(ref["margins"][k, b]))


def test_awq_large_batch_dequantise_once_equals_fused_kernel():
    """From 256 rows on t3b_decode_step dequantises a layer's AWQ projections once into a bf16 scratch and runs the bf16 kernels
    (upstream switches to dequantise-then-GEMM at 256 tokens too, vllm awq.py:276-281).  The same 256 rows as two 128-row steps go
    through the dequant-fused kernel: same definition bf16_rn((q - z) * s), so the hidden states agree to bf16 rounding (the
    split-K factor differs between the two row counts) and the KV pool bit for bit where both wrote it."""
    import ctypes as C
    from helpers import random_state
    m = T3B200Model(dims=SMALL, device=DEV, cfg_scale=0.5, quantization="awq")
    m.load_weights(synthetic_checkpoint(SMALL).items())
    L = _lib.lib()
    rows = 256
    lens = [20 + (7 * r) % 90 for r in range(rows)]
    kv, st = random_state(SMALL, rows, lens, seed=11)
    ws = m.workspace(rows)
    big = torch.empty(rows, 1024, dtype=torch.bfloat16, device=DEV)
    s = torch.cuda.current_stream().cuda_stream
    _lib.check(L.t3b_decode_step(C.byref(m.c), C.byref(kv.c), C.byref(st.c), None, rows, big.data_ptr(), ws.data_ptr(), ws.numel(), None, s))
    k_big, v_big = kv.k.clone(), kv.v.clone()
    halves = torch.empty_like(big)
    for h in range(2):
        view = st.view(64 * h)
        _lib.check(L.t3b_decode_step(C.byref(m.c), C.byref(kv.c), C.byref(view), None, 128, halves[128 * h:].data_ptr(), ws.data_ptr(),
                                     ws.numel(), None, s))
    torch.cuda.synchronize()
    a, b = big.float().cpu(), halves.float().cpu()
    assert torch.isfinite(a).all() and (a - b).abs().max() <= 0.0625, float((a - b).abs().max())
    assert torch.nn.functional.cosine_similarity(a.flatten().double(), b.flatten().double(), dim=0) > 0.9999
    # layer 0's new K/V come from the same embeddings through the same weights: the pools agree up to the split-K rounding of the projections
    for now, then in ((kv.k, k_big), (kv.v, v_big)):
        d = (now.view(torch.bfloat16).float() - then.view(torch.bfloat16).float()).abs().max()
        assert float(d) <= 0.0625, float(d)
