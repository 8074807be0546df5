"""Parity of the fused persistent step kernel (csrc/decode_fused.cu, C ABI t3b_decode_step_fused) against the
multi-kernel path (t3b_decode_step + head), which test_gpu_decode.py pins to the CPU oracle; the engine tests
of test_gpu_decode.py run THROUGH this kernel (it is the default decode path for <= 128 rows) and compare
token streams with the oracle directly.

Both CUDA paths round where the oracle's bf16 mode rounds; they differ in fp32 summation order only, so the
tolerances are those of test_gpu_decode.py (2 layers: max-abs 0.0625 on logits, cosine >= 0.9999).
Bit-exact properties checked here: run-to-run determinism, batch invariance (a request's logits do not depend
on which other requests share the step) and the repacked weight layout.
"""
import ctypes as C

import pytest
import torch

from chatterbox_vllm2_b200 import T3Dims, _lib
from chatterbox_vllm2_b200.model import T3B200Model
from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint
from test_gpu_mega import cosine, random_state

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SMALL = T3Dims().small(n_layers=2, max_pos=2048)
_models = {}


def model_for(dims):
    key = dims.n_layers
    if key not in _models:
        m = T3B200Model(dims=dims, device=DEV, cfg_scale=0.5)
        m.load_weights(synthetic_checkpoint(dims).items())
        if m.fused is None:
            m._build_fused()            # opt-in in production (T3B_FUSED=1); the parity tests always build it
        assert m.fused is not None
        _models[key] = m
    return _models[key]


def run_both(m, kv, st, rows):
    d = m.dims
    k0, v0 = kv.k.clone(), kv.v.clone()
    hid_ref = torch.empty(rows, d.hidden, dtype=torch.bfloat16, device=DEV)
    m.decode_step(kv, st, rows, hid_ref)
    torch.cuda.synchronize()
    k_ref, v_ref = kv.k.clone(), kv.v.clone()
    kv.k.copy_(k0); kv.v.copy_(v0)
    hid = torch.zeros(rows, d.hidden, dtype=torch.bfloat16, device=DEV)
    lg = torch.zeros(rows, d.speech_vocab, dtype=torch.bfloat16, device=DEV)
    m.decode_step_fused(kv, st.c, rows, hid, lg)
    torch.cuda.synchronize()
    return hid_ref, k_ref, v_ref, hid, lg


def test_fused_repack_layout_matches_definition():
    """Every 16-byte chunk of the blob holds the weights the kernel comment says it does."""
    m = model_for(SMALL)
    d = m.dims
    blob = m.fused.cpu()
    layer_bytes, a_bytes, b_bytes, sec_a = 32 << 20, 131072, 196608, 8 << 20
    g = torch.Generator().manual_seed(0)
    wq, wo, wgu, wd = (m.w[f"1.{n}"].cpu() for n in ("qkv", "o", "gate_up", "down"))
    layer1 = blob[layer_bytes:2 * layer_bytes].view(torch.bfloat16)
    ri = lambda n: int(torch.randint(0, n, (1,), generator=g))
    for _ in range(400):
        j, sec, pc = ri(4), ri(4), ri(8)
        if sec == 0:
            h, kb, rr = ri(16), ri(16), ri(48)
            off = (h * 4 + j) * a_bytes + kb * 6144 + rr * 128 + pc * 16
            f = 48 * j + rr
            src = wq[(f >> 6) * 1024 + h * 64 + (f & 63), kb * 64 + (pc ^ (rr & 7)) * 8:][:8]
        elif sec == 1:
            h, t, rr = ri(16), ri(2), ri(128)
            off = (h * 4 + j) * a_bytes + 98304 + t * 16384 + rr * 128 + pc * 16
            src = wo[256 * j + 128 * t + rr, h * 64 + (pc ^ (rr & 7)) * 8:][:8]
        elif sec == 2:
            c, kb, rr = ri(32), ri(16), ri(64)
            off = sec_a + (c * 4 + j) * b_bytes + kb * 8192 + rr * 128 + pc * 16
            f = 128 * c + 32 * j + (rr & 31)
            src = wgu[f if rr < 32 else 4096 + f, kb * 64 + (pc ^ (rr & 7)) * 8:][:8]
        else:
            c, t, kb, rr = ri(32), ri(2), ri(2), ri(128)
            off = sec_a + (c * 4 + j) * b_bytes + 131072 + (t * 2 + kb) * 16384 + rr * 128 + pc * 16
            src = wd[256 * j + 128 * t + rr, 128 * c + kb * 64 + (pc ^ (rr & 7)) * 8:][:8]
        got = layer1[off // 2: off // 2 + 8]
        assert torch.equal(got.view(torch.int16), src.contiguous().view(torch.int16)), (sec, j, off)
    # head + norms
    hb = 2 * layer_bytes
    head = blob[hb:hb + 128 * 147456].view(torch.bfloat16).reshape(128, 147456 // 2)
    wh = m.w["head"].cpu()
    for _ in range(100):
        cta, kb, rr, pc = (int(torch.randint(0, n, (1,), generator=g)) for n in (114, 16, 72, 8))
        n = cta * 72 + rr
        got = head[cta, (kb * 9216 + rr * 128 + pc * 16) // 2:][:8]
        ref = wh[n, kb * 64 + (pc ^ (rr & 7)) * 8:][:8] if n < d.speech_vocab else torch.zeros(8, dtype=torch.bfloat16)
        assert torch.equal(got.view(torch.int16), ref.contiguous().view(torch.int16)), (cta, kb, rr, pc)
    norms = blob[hb + 128 * 147456:].view(torch.bfloat16).reshape(2 * d.n_layers + 1, 1024)
    assert torch.equal(norms[1].view(torch.int16), m.w["0.ln2"].cpu().view(torch.int16))
    assert torch.equal(norms[-1].view(torch.int16), m.w["norm"].cpu().view(torch.int16))


@pytest.mark.parametrize("rows,lo,hi", [(2, 1, 40), (2, 99, 100), (2, 0, 0), (6, 0, 130), (16, 60, 260), (64, 90, 200), (128, 1, 150),
                                        (34, 255, 258), (4, 500, 700), (8, 1100, 1180)])
def test_fused_matches_multi_kernel_path(rows, lo, hi):
    m = model_for(SMALL)
    g = torch.Generator().manual_seed(rows * 7 + lo)
    lens = torch.randint(lo, hi + 1, (rows,), generator=g).tolist()
    kv, st = random_state(SMALL, rows, lens, seed=rows)
    hid_ref, k_ref, v_ref, hid, lg = run_both(m, kv, st, rows)
    assert torch.isfinite(hid.float()).all() and torch.isfinite(lg.float()).all()
    dh = (hid.float() - hid_ref.float()).abs().max().item()
    ch = cosine(hid, hid_ref)
    lg_ref = hid_ref.float() @ m.w["head"].float().T
    dl = (lg.float() - lg_ref).abs().max().item()
    cl = cosine(lg, lg_ref)
    assert ch >= 0.9999 and cl >= 0.9999 and dl <= 0.0625, dict(rows=rows, max_abs_hidden=dh, cos_hidden=ch, max_abs_logits=dl, cos_logits=cl)
    # this step's K/V landed in the same cache slots (bf16: <= 1 ulp of rounding noise on K after RoPE)
    assert (kv.k.float() - k_ref.float()).abs().max().item() <= 0.0625
    assert (kv.v.float() - v_ref.float()).abs().max().item() <= 0.0625
    assert int((kv.k != k_ref).sum()) <= rows * SMALL.hidden * SMALL.n_layers      # only this step's slots may differ at all


def test_fused_deterministic_and_batch_invariant():
    m = model_for(SMALL)
    rows = 64
    g = torch.Generator().manual_seed(5)
    lens = torch.randint(70, 220, (rows,), generator=g).tolist()
    kv, st = random_state(SMALL, rows, lens, seed=11)
    k0, v0 = kv.k.clone(), kv.v.clone()
    outs = []
    for _ in range(2):
        kv.k.copy_(k0); kv.v.copy_(v0)
        lg = torch.zeros(rows, SMALL.speech_vocab, dtype=torch.bfloat16, device=DEV)
        m.decode_step_fused(kv, st.c, rows, None, lg)
        torch.cuda.synchronize()
        outs.append(lg)
    assert torch.equal(outs[0].view(torch.int16), outs[1].view(torch.int16))
    # the first request (rows 0, 1) alone in a step: same bits as inside the 64-row step
    kv.k.copy_(k0); kv.v.copy_(v0)
    lg2 = torch.zeros(2, SMALL.speech_vocab, dtype=torch.bfloat16, device=DEV)
    m.decode_step_fused(kv, st.c, 2, None, lg2)
    torch.cuda.synchronize()
    assert torch.equal(lg2.view(torch.int16), outs[0][:2].view(torch.int16))


def test_fused_full_depth():
    dims = T3Dims()
    m = model_for(dims)
    rows = 8
    lens = [99, 99, 130, 130, 257, 257, 64, 64]
    kv, st = random_state(dims, rows, lens, seed=2)
    hid_ref, k_ref, v_ref, hid, lg = run_both(m, kv, st, rows)
    lg_ref = hid_ref.float() @ m.w["head"].float().T
    dl = (lg.float() - lg_ref).abs().max().item()
    cl = cosine(lg, lg_ref)
    assert cl >= 0.999 and dl <= 0.25, dict(max_abs_logits=dl, cos_logits=cl)


def test_fused_rejects_bad_rows():
    m = model_for(SMALL)
    kv, st = random_state(SMALL, 2, [5, 5])
    lg = torch.zeros(130, SMALL.speech_vocab, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(_lib.T3BLibraryError):
        m.decode_step_fused(kv, st.c, 130, None, lg)
    with pytest.raises(_lib.T3BLibraryError):
        m.decode_step_fused(kv, st.c, 3, None, lg)
