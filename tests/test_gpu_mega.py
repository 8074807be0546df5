"""Parity of the persistent whole-step kernel (csrc/decode_mega.cu, C ABI t3b_decode_step_mega) against the
multi-kernel path (t3b_decode_step + head), which test_gpu_decode.py pins to the CPU oracle, and directly
against the oracle through the engine.

Both CUDA paths round where the oracle's bf16 mode rounds; they differ in fp32 summation order only, so the
tolerances are those of test_gpu_decode.py (2 layers: max-abs 0.0625 on logits, cosine >= 0.9999).
Bit-exact properties checked here: run-to-run determinism and batch invariance (a request's logits do not
depend on which other requests share the step), and the repacked weight layout.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from chatterbox_vllm2_b200 import T3Dims, _lib
from chatterbox_vllm2_b200.model import DecodeState, KVPool, T3B200Model
from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SMALL = T3Dims().small(n_layers=2, max_pos=512)
_models = {}


def model_for(dims):
    key = dims.n_layers
    if key not in _models:
        m = T3B200Model(dims=dims, device=DEV, cfg_scale=0.5)
        m.load_weights(synthetic_checkpoint(dims).items())
        if m.mega is None:
            m._build_mega()          # opt-in path (T3B_MEGA=1): build its repacked weights for these tests
        assert m.mega is not None
        _models[key] = m
    return _models[key]


def random_state(dims, rows, lens, seed=0):
    """KV pool filled with random bf16 K/V, a decode state whose row r has lens[r] cached tokens."""
    g = torch.Generator().manual_seed(seed)
    max_blocks = (max(lens) + 16) // 16 + 1
    nblk = rows * max_blocks + 1
    kv = KVPool(dims, nblk, DEV)
    kv.k.copy_((torch.randn(kv.k.shape, generator=g) * 0.5).to(torch.bfloat16))
    kv.v.copy_((torch.randn(kv.v.shape, generator=g) * 0.5).to(torch.bfloat16))
    st = DecodeState(rows // 2, max_blocks, 64, max(lens) + 8, DEV)
    perm = torch.randperm(nblk - 1, generator=g)[: rows * max_blocks] + 1
    st.block_tables.copy_(perm.reshape(rows, max_blocks).to(torch.int32))
    st.ctx_lens.copy_(torch.tensor(lens, dtype=torch.int32))
    st.rope_pos.copy_(torch.tensor(lens, dtype=torch.int32))
    st.tokens.copy_(torch.randint(0, dims.speech_vocab, (rows // 2,), generator=g).to(torch.int32))
    st.speech_pos.copy_(torch.randint(0, 300, (rows // 2,), generator=g).to(torch.int32))
    return kv, st


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
    m.decode_step_mega(kv, st.c, rows, hid, lg)
    torch.cuda.synchronize()
    return hid_ref, k_ref, v_ref, hid, lg


def cosine(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a * b).sum() / (a.norm() * b.norm()))


def test_repack_layout_matches_definition():
    L = _lib.lib()
    g = torch.Generator().manual_seed(3)
    for n, k, gu in ((256, 128, 0), (1024, 4096, 0), (512, 1024, 1), (8194, 1024, 0)):
        w = torch.randn(n, k, generator=g).to(torch.bfloat16).to(DEV)
        out = torch.empty(L.t3b_repacked_bytes(n, k), dtype=torch.uint8, device=DEV)
        _lib.check(L.t3b_repack_weights_bf16(w.data_ptr(), n, k, gu, out.data_ptr(), _lib.stream_ptr()))
        torch.cuda.synchronize()
        nt, nkb = (n + 127) // 128, k // 64
        got = out.view(torch.bfloat16).reshape(nt, nkb, 128, 8, 8).cpu()
        wp = torch.zeros(nt * 128, k, dtype=torch.bfloat16)
        wp[:n] = w.cpu()
        if gu:
            idx = torch.cat([torch.cat([torch.arange(64) + 64 * t, torch.arange(64) + n // 2 + 64 * t]) for t in range(nt)])
            wp = wp[idx]
        ref = wp.reshape(nt, 128, nkb, 8, 8).permute(0, 2, 1, 3, 4)          # [tile][kb][row][chunk][8]
        r = torch.arange(128).reshape(128, 1)
        pc = torch.arange(8).reshape(1, 8)
        src_chunk = (pc ^ (r & 7)).reshape(1, 1, 128, 8, 1).expand(nt, nkb, 128, 8, 8)
        ref_sw = torch.gather(ref, 3, src_chunk)                              # stored chunk pc holds logical chunk pc ^ (r & 7)
        assert torch.equal(got.view(torch.int16), ref_sw.contiguous().view(torch.int16)), (n, k, gu)


@pytest.mark.parametrize("rows,lo,hi", [(2, 1, 40), (2, 99, 100), (6, 0, 130), (16, 60, 260), (64, 90, 200), (128, 1, 150), (34, 255, 258)])
def test_mega_matches_multi_kernel_path(rows, lo, hi):
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


def test_mega_deterministic_and_batch_invariant():
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
        m.decode_step_mega(kv, st.c, rows, None, lg)
        torch.cuda.synchronize()
        outs.append(lg)
    assert torch.equal(outs[0].view(torch.int16), outs[1].view(torch.int16))
    # the first request (rows 0, 1) alone in a step: same bits as inside the 64-row step
    kv.k.copy_(k0); kv.v.copy_(v0)
    lg2 = torch.zeros(2, SMALL.speech_vocab, dtype=torch.bfloat16, device=DEV)
    m.decode_step_mega(kv, st.c, 2, None, lg2)
    torch.cuda.synchronize()
    assert torch.equal(lg2.view(torch.int16), outs[0][:2].view(torch.int16))


def test_mega_full_depth():
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


def test_mega_rejects_bad_rows():
    m = model_for(SMALL)
    kv, st = random_state(SMALL, 2, [5, 5])
    lg = torch.zeros(130, SMALL.speech_vocab, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(_lib.T3BLibraryError):
        m.decode_step_mega(kv, st.c, 130, None, lg)
