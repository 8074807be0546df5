"""Round-2 kernels through the C ABI against the CPU oracle: FP8 paged KV (write + attention), the read-once prefill attention,
the prefill embedding assembly, the S3Gen token filter, the in-loop analyzer heuristics (pinned to the reference's own class through
tests/golden/analyzer_cases.json) and the launch-shape options."""
import ctypes as C
import json
import os

import numpy as np
import pytest
import torch

from chatterbox_vllm2_b200 import T3Dims, _lib
from chatterbox_vllm2_b200.model import DecodeState, KVPool, SamplingTensors, T3B200Model
from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint, synthetic_request
from helpers import GOLD, random_state
from oracle import analyzer_ref, kv_fp8_ref
from oracle.t3_ref import T3Oracle, rope_cos_sin_table

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SMALL = T3Dims().small(n_layers=2, max_pos=512)


@pytest.fixture(scope="module")
def L():
    lib = _lib.lib()
    _lib.check(lib.t3b_check_device(), "t3b_check_device")
    return lib


@pytest.fixture(scope="module")
def small_model():
    sd = synthetic_checkpoint(SMALL)
    m = T3B200Model(dims=SMALL, device=DEV, cfg_scale=0.5)
    m.load_weights(sd.items())
    return m, sd


def sp():
    return torch.cuda.current_stream().cuda_stream


def i32(x):
    return torch.as_tensor(x, dtype=torch.int32).to(DEV)


def ulp_diff(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    mag = torch.maximum(a.abs(), b.abs()).clamp_min(1e-30)
    return (a - b).abs() / torch.exp2(torch.floor(torch.log2(mag)) - 7)


def _pool(n_rows, max_blocks, kv_dtype, seed):
    nb = n_rows * max_blocks + 1
    perm = torch.randperm(nb - 1, generator=torch.Generator().manual_seed(seed)) + 1
    bt = perm[: n_rows * max_blocks].reshape(n_rows, max_blocks).to(torch.int32)
    slice_b = 1088 if kv_dtype == _lib.T3B_KV_FP8 else 2048
    k = torch.zeros(nb, 16, slice_b, dtype=torch.uint8, device=DEV)
    return bt, k, torch.zeros_like(k)


def _softmax_ref(q, K, V):
    """q [H, d], K/V [L, H, d] fp32 -> [H*d] (bf16-rounded output, like the kernel's store)."""
    s = torch.einsum("hd,lhd->hl", q, K) * 0.125
    return torch.einsum("hl,lhd->hd", torch.softmax(s, -1), V).bfloat16().float().reshape(-1)


# ---------------------------------------------------------------------------------------------- FP8 KV
def test_fp8_kv_write_bit_exact_and_attention_matches_dequantised_oracle(L):
    """t3b_rope_kv_write (T3B_KV_FP8) stores exactly oracle.kv_fp8_ref's e4m3 bytes and scales; t3b_paged_attn_decode over that pool
    equals the fp32 softmax over the DEQUANTISED K/V (1 bf16 ulp: fp32 summation order only)."""
    g = torch.Generator().manual_seed(5)
    ctxs = [1, 15, 16, 17, 33, 100, 250, 700]
    n_rows, mb = len(ctxs), (max(ctxs) + 15) // 16
    bt, kp, vp = _pool(n_rows, mb, _lib.T3B_KV_FP8, 1)
    cs = rope_cos_sin_table(SMALL).to(device=DEV, dtype=torch.bfloat16)
    orc = T3Oracle(synthetic_checkpoint(SMALL), SMALL, mode="bf16")
    rows, idx = [], []
    for r, c in enumerate(ctxs):
        rows += [r] * c
        idx += list(range(c))
    n = len(rows)
    qkv = torch.randn(n, 3072, generator=g).bfloat16()
    pos = torch.tensor(idx) % SMALL.max_pos
    q_out = torch.empty(n, 1024, dtype=torch.bfloat16, device=DEV)
    btd = bt.to(DEV)
    # every device tensor is held in a variable until the launch has run: a temporary's block goes back to the caching
    # allocator as soon as data_ptr() returns and the next temporary would reuse it
    qkv_d, rows_d, idx_d, pos_d = qkv.to(DEV), i32(rows), i32(idx), i32(pos)
    _lib.check(L.t3b_rope_kv_write(qkv_d.data_ptr(), None, 1, n, rows_d.data_ptr(), idx_d.data_ptr(), pos_d.data_ptr(),
                                   cs.data_ptr(), SMALL.max_pos, btd.data_ptr(), mb, kp.data_ptr(), vp.data_ptr(), 16, _lib.T3B_KV_FP8,
                                   q_out.data_ptr(), sp()))
    torch.cuda.synchronize()
    k_ref = orc.rope(qkv[:, 1024:2048].float().reshape(n, 16, 64), pos)
    v_ref = qkv[:, 2048:].float().reshape(n, 16, 64)
    kq, ks = kv_fp8_ref.quantize(k_ref)           # uint8 [n,16,64], fp32 [n,16]
    vq, vs = kv_fp8_ref.quantize(v_ref)
    kh, vh = kp.cpu(), vp.cpu()
    for j in range(0, n, 7):
        blk, t = int(bt[rows[j], idx[j] // 16]), idx[j] % 16
        assert torch.equal(kh[blk, :, t * 64:(t + 1) * 64], kq[j]) and torch.equal(vh[blk, :, t * 64:(t + 1) * 64], vq[j])
        assert torch.equal(kh[blk, :, 1024 + 4 * t:1028 + 4 * t].contiguous().view(torch.float32).reshape(-1), ks[j])
        assert torch.equal(vh[blk, :, 1024 + 4 * t:1028 + 4 * t].contiguous().view(torch.float32).reshape(-1), vs[j])
    # attention of each row's last token over the quantised pool
    last = [sum(ctxs[:r + 1]) - 1 for r in range(n_rows)]
    q = q_out[torch.tensor(last, device=DEV)].contiguous()
    out = torch.empty(n_rows, 1024, dtype=torch.bfloat16, device=DEV)
    last_d = i32([c - 1 for c in ctxs])
    _lib.check(L.t3b_paged_attn_decode(q.data_ptr(), n_rows, None, last_d.data_ptr(), btd.data_ptr(), mb, kp.data_ptr(),
                                       vp.data_ptr(), 16, _lib.T3B_KV_FP8, out.data_ptr(), None, sp()))
    torch.cuda.synchronize()
    kd, vd = kv_fp8_ref.dequantize(kq, ks), kv_fp8_ref.dequantize(vq, vs)
    o0 = 0
    for r, c in enumerate(ctxs):
        ref = _softmax_ref(q[r].float().cpu().reshape(16, 64), kd[o0:o0 + c], vd[o0:o0 + c])
        got = out[r].float().cpu()
        assert ((ulp_diff(got, ref) <= 1.0) | ((got - ref).abs() < 2e-3)).all(), (r, c, float((got - ref).abs().max()))
        o0 += c


@pytest.mark.parametrize("kv_dtype", [_lib.T3B_KV_BF16, _lib.T3B_KV_FP8])
def test_fused_decode_attention_equals_write_then_attend(L, kv_dtype):
    """t3b_attn_decode_fused (new token attended from registers, then stored) == t3b_rope_kv_write + t3b_paged_attn_decode on
    the same pool format: pool bytes identical, outputs within 1 bf16 ulp."""
    ctxs = [1, 2, 16, 17, 32, 33, 64, 65, 250, 511, 1200]
    n, mb = len(ctxs), (max(ctxs) + 15) // 16
    bt, kp, vp = _pool(n, mb, kv_dtype, 3)
    g = torch.Generator().manual_seed(9)
    if kv_dtype == _lib.T3B_KV_BF16:
        kp.view(torch.bfloat16).copy_(torch.randn(kp.numel() // 2, generator=g).bfloat16().reshape(kp.view(torch.bfloat16).shape))
        vp.view(torch.bfloat16).copy_(torch.randn(vp.numel() // 2, generator=g).bfloat16().reshape(vp.view(torch.bfloat16).shape))
    else:
        kq, ks = kv_fp8_ref.quantize(torch.randn(kp.shape[0] * 16, 16, 64, generator=g))
        vq, vs = kv_fp8_ref.quantize(torch.randn(kp.shape[0] * 16, 16, 64, generator=g))
        for pool, qv, sc in ((kp, kq, ks), (vp, vq, vs)):
            nb = pool.shape[0]
            pool[:, :, :1024] = qv.reshape(nb, 16, 16, 64).permute(0, 2, 1, 3).reshape(nb, 16, 1024).to(DEV)
            pool[:, :, 1024:] = sc.reshape(nb, 16, 16).permute(0, 2, 1).contiguous().view(torch.uint8).reshape(nb, 16, 64).to(DEV)
    kp2, vp2 = kp.clone(), vp.clone()
    qkv = torch.randn(n, 3072, generator=g).bfloat16().to(DEV)
    idx, pos = i32([c - 1 for c in ctxs]), i32([(c * 7) % SMALL.max_pos for c in ctxs])
    cs = rope_cos_sin_table(SMALL).to(device=DEV, dtype=torch.bfloat16)
    btd = bt.to(DEV)
    q = torch.empty(n, 1024, dtype=torch.bfloat16, device=DEV)
    out_a, out_b = torch.empty_like(q), torch.empty_like(q)
    _lib.check(L.t3b_rope_kv_write(qkv.data_ptr(), None, 1, n, None, idx.data_ptr(), pos.data_ptr(), cs.data_ptr(), SMALL.max_pos,
                                   btd.data_ptr(), mb, kp.data_ptr(), vp.data_ptr(), 16, kv_dtype, q.data_ptr(), sp()))
    _lib.check(L.t3b_paged_attn_decode(q.data_ptr(), n, None, idx.data_ptr(), btd.data_ptr(), mb, kp.data_ptr(), vp.data_ptr(), 16, kv_dtype,
                                       out_a.data_ptr(), None, sp()))
    _lib.check(L.t3b_attn_decode_fused(qkv.data_ptr(), n, idx.data_ptr(), pos.data_ptr(), cs.data_ptr(), SMALL.max_pos, btd.data_ptr(), mb,
                                       kp2.data_ptr(), vp2.data_ptr(), 16, kv_dtype, out_b.data_ptr(), None, sp()))
    assert torch.equal(kp, kp2) and torch.equal(vp, vp2)
    a, b = out_a.float().cpu(), out_b.float().cpu()
    assert ((ulp_diff(a, b) <= 1.0) | ((a - b).abs() < 2e-3)).all(), float((a - b).abs().max())


@pytest.mark.parametrize("team", [1, 2, 4, 8])
def test_attention_team_sizes_match_fp32_softmax_and_each_other(L, team):
    """t3b_options_t.attn_team_warps: 1, 2, 4 or 8 warps per (row, head) item - every split of the online softmax equals the fp32
    softmax over the pool (1 bf16 ulp), for ragged contexts from 1 token to the 1200-token limit, many items per warp (more items
    than the grid has teams) and the fused decode variant (new token from registers)."""
    g = torch.Generator().manual_seed(40 + team)
    ctxs = [1, 15, 16, 17, 31, 32, 33, 64, 100, 250, 700, 1200] * 8            # 96 rows x 16 heads = 1536 items
    n, mb = len(ctxs), (max(ctxs) + 15) // 16
    bt, kp, vp = _pool(n, mb, _lib.T3B_KV_BF16, 6)
    kp.view(torch.bfloat16).copy_(torch.randn(kp.numel() // 2, generator=g).bfloat16().reshape(kp.view(torch.bfloat16).shape))
    vp.view(torch.bfloat16).copy_(torch.randn(vp.numel() // 2, generator=g).bfloat16().reshape(vp.view(torch.bfloat16).shape))
    kp2, vp2 = kp.clone(), vp.clone()
    qkv = torch.randn(n, 3072, generator=g).bfloat16().to(DEV)
    idx, pos = i32([c - 1 for c in ctxs]), i32([(c * 7) % SMALL.max_pos for c in ctxs])
    cs = rope_cos_sin_table(SMALL).to(device=DEV, dtype=torch.bfloat16)
    btd = bt.to(DEV)
    opt = _lib.make_options(attn_team_warps=team)
    q = torch.empty(n, 1024, dtype=torch.bfloat16, device=DEV)
    out_a, out_b = torch.empty_like(q), torch.empty_like(q)
    _lib.check(L.t3b_rope_kv_write(qkv.data_ptr(), None, 1, n, None, idx.data_ptr(), pos.data_ptr(), cs.data_ptr(), SMALL.max_pos,
                                   btd.data_ptr(), mb, kp.data_ptr(), vp.data_ptr(), 16, 0, q.data_ptr(), sp()))
    _lib.check(L.t3b_paged_attn_decode(q.data_ptr(), n, None, idx.data_ptr(), btd.data_ptr(), mb, kp.data_ptr(), vp.data_ptr(), 16, 0,
                                       out_a.data_ptr(), C.byref(opt), sp()))
    _lib.check(L.t3b_attn_decode_fused(qkv.data_ptr(), n, idx.data_ptr(), pos.data_ptr(), cs.data_ptr(), SMALL.max_pos, btd.data_ptr(), mb,
                                       kp2.data_ptr(), vp2.data_ptr(), 16, 0, out_b.data_ptr(), C.byref(opt), sp()))
    torch.cuda.synchronize()
    assert torch.equal(kp, kp2) and torch.equal(vp, vp2)
    kh = kp.view(torch.bfloat16).view(-1, 16, 16, 64).float().cpu()            # [block, head, tok, d]
    vh = vp.view(torch.bfloat16).view(-1, 16, 16, 64).float().cpu()
    qh = q.float().cpu().reshape(n, 16, 64)
    a, b = out_a.float().cpu(), out_b.float().cpu()
    for r in range(0, n, 5):
        c = ctxs[r]
        blocks = bt[r, :(c + 15) // 16].long()
        K = kh[blocks].permute(0, 2, 1, 3).reshape(-1, 16, 64)[:c]
        V = vh[blocks].permute(0, 2, 1, 3).reshape(-1, 16, 64)[:c]
        ref = _softmax_ref(qh[r], K, V)
        for name, o in (("paged", a[r]), ("fused", b[r])):
            assert ((ulp_diff(o, ref) <= 1.0) | ((o - ref).abs() < 2e-3)).all(), (name, team, r, c, float((o - ref).abs().max()))


# ---------------------------------------------------------------------------------------------- prefill
@pytest.mark.parametrize("kv_dtype", [_lib.T3B_KV_BF16, _lib.T3B_KV_FP8])
def test_prefill_attention_matches_causal_softmax(L, kv_dtype):
    """t3b_prefill_attn: every query of every (ragged) prompt against fp32 causal softmax over the pool's K/V."""
    g = torch.Generator().manual_seed(2)
    lens = [1, 5, 16, 17, 99, 100, 250]
    n_seq, mb = len(lens), (max(lens) + 15) // 16
    bt, kp, vp = _pool(n_seq, mb, kv_dtype, 4)
    starts = [0]
    for l in lens:
        starts.append(starts[-1] + l)
    n = starts[-1]
    rows = [r for r, l in enumerate(lens) for _ in range(l)]
    idx = [t for l in lens for t in range(l)]
    qkv = torch.randn(n, 3072, generator=g).bfloat16()
    cs = rope_cos_sin_table(SMALL).to(device=DEV, dtype=torch.bfloat16)
    q = torch.empty(n, 1024, dtype=torch.bfloat16, device=DEV)
    btd = bt.to(DEV)
    qkv_d, rows_d, idx_d, starts_d, seqrow_d = qkv.to(DEV), i32(rows), i32(idx), i32(starts), i32(list(range(n_seq)))   # held, not temporaries
    _lib.check(L.t3b_rope_kv_write(qkv_d.data_ptr(), None, 1, n, rows_d.data_ptr(), idx_d.data_ptr(), idx_d.data_ptr(),
                                   cs.data_ptr(), SMALL.max_pos, btd.data_ptr(), mb, kp.data_ptr(), vp.data_ptr(), 16, kv_dtype, q.data_ptr(), sp()))
    out = torch.empty(n, 1024, dtype=torch.bfloat16, device=DEV)
    assert max(lens) <= L.t3b_prefill_attn_max_len(kv_dtype)
    _lib.check(L.t3b_prefill_attn(q.data_ptr(), n_seq, starts_d.data_ptr(), seqrow_d.data_ptr(), btd.data_ptr(), mb,
                                  kp.data_ptr(), vp.data_ptr(), 16, kv_dtype, max(lens), out.data_ptr(), sp()))
    # the same queries through the per-token paged kernel: another summation order of the same softmax
    out2 = torch.empty_like(out)
    _lib.check(L.t3b_paged_attn_decode(q.data_ptr(), n, rows_d.data_ptr(), idx_d.data_ptr(), btd.data_ptr(), mb, kp.data_ptr(),
                                       vp.data_ptr(), 16, kv_dtype, out2.data_ptr(), None, sp()))
    torch.cuda.synchronize()
    orc = T3Oracle(synthetic_checkpoint(SMALL), SMALL, mode="bf16")
    k_ref = orc.rope(qkv[:, 1024:2048].float().reshape(n, 16, 64), torch.tensor(idx))
    v_ref = qkv[:, 2048:].float().reshape(n, 16, 64)
    if kv_dtype == _lib.T3B_KV_FP8:
        k_ref, v_ref = kv_fp8_ref.roundtrip(k_ref), kv_fp8_ref.roundtrip(v_ref)
    qh = q.float().cpu().reshape(n, 16, 64)
    got, got2 = out.float().cpu(), out2.float().cpu()
    for s in range(n_seq):
        for j in range(lens[s]):
            t = starts[s] + j
            ref = _softmax_ref(qh[t], k_ref[starts[s]:t + 1], v_ref[starts[s]:t + 1])
            for name, o in (("prefill", got[t]), ("paged", got2[t])):
                assert ((ulp_diff(o, ref) <= 1.0) | ((o - ref).abs() < 2e-3)).all(), (name, s, j, float((o - ref).abs().max()))


def test_prefill_embed_bit_exact(small_model):
    """t3b_prefill_embed == the oracle's row layout of t3.py:542-561 for ragged prompts, plus the token index arrays."""
    m, sd = small_model
    orc = T3Oracle(sd, SMALL, mode="bf16")
    reqs = [synthetic_request(i, SMALL, n_text=n) for i, n in enumerate((1, 9, 64, 17, 2))]
    slots = [3, 0, 7, 1, 5]
    x, tok_row, tok_idx, starts = m.prefill_assemble([c for c, _ in reqs], [ids for _, ids in reqs], slots)
    torch.cuda.synchronize()
    xs = x.float().cpu()
    for r, (cond, ids) in enumerate(reqs):
        c_ref, u_ref = orc.prefill_embeds(cond.float(), ids)
        a, b, c = starts[2 * r], starts[2 * r + 1], starts[2 * r + 2]
        assert torch.equal(xs[a:b], c_ref) and torch.equal(xs[b:c], u_ref)
        assert tok_row[a:b].cpu().tolist() == [2 * slots[r]] * (b - a) and tok_row[b:c].cpu().tolist() == [2 * slots[r] + 1] * (c - b)
        assert tok_idx[a:b].cpu().tolist() == list(range(b - a)) and tok_idx[b:c].cpu().tolist() == list(range(c - b))
    seq_start, seq_row = m.last_seq
    assert seq_start.cpu().tolist() == starts and seq_row.cpu().tolist() == [2 * s + h for s in slots for h in (0, 1)]


def test_get_input_embeddings_prefill_blocks_and_chunk_cases(small_model):
    """The reference's plugin protocol (t3.py:143-249 packing, :340-421 split, :542-639 the four chunk cases): whole block,
    block cut inside the text, block cut inside the conditioning prefix, and a mixed prefill + decode batch."""
    from chatterbox_vllm2_b200 import SPEECH_TOKEN_OFFSET, packing
    m, sd = small_model
    cond, ids = synthetic_request(4, SMALL, n_text=21)
    final, rows = packing.pack_prompt(ids, cond.float())
    assert final[0] == 695 and final[33] == 696 and final[-1] == 697 and final[1:33] == [ids[0]] * 32 and len(final) == 34 + 21 + 1
    assert packing.unpack_prompt(final) == ids
    c_ref, u_ref = m.prefill_embeds(cond, ids)
    want = torch.cat([c_ref, u_ref], dim=1)
    whole = m.get_input_embeddings(torch.tensor(final), [rows])
    assert whole.shape == (len(final), 2048) and torch.equal(whole, want)
    for cut in (40, 34, 20, 55):                               # inside the text / right after the prefix / inside it / BOS alone
        a = m.get_input_embeddings(torch.tensor(final[:cut]), [rows[:cut]])
        b = m.get_input_embeddings(torch.tensor(final[cut:]), [rows[cut:]])
        assert torch.equal(torch.cat([a, b], 0), want), cut
    dec = torch.tensor([10 + SPEECH_TOKEN_OFFSET, 99 + SPEECH_TOKEN_OFFSET])
    mixed = m.get_input_embeddings(torch.cat([dec, torch.tensor(final)]), [rows])
    assert torch.equal(mixed[2:], want) and torch.equal(mixed[:2], m.get_input_embeddings(dec, None))
    with pytest.raises(ValueError):
        m.get_input_embeddings(torch.tensor(final[36:50]), [rows[36:50]])          # neither start nor end (t3.py:648-652)


def test_filter_speech_tokens(small_model):
    """tts.py:492-514: subtract 2500, keep 0 <= t < 6561, order preserved."""
    m, _ = small_model
    g = torch.Generator().manual_seed(1)
    for n in (0, 1, 5, 1024, 3000):
        ids = torch.randint(0, 2500 + 8194, (n,), generator=g)
        got = m.filter_speech_tokens(ids).cpu().tolist() if n else []
        assert got == [t - 2500 for t in ids.tolist() if 0 <= t - 2500 < 6561]


# ---------------------------------------------------------------------------------------------- analyzer heuristics in the sampler
def test_sampler_analyzer_heuristics_match_the_reference_class(L):
    """Every decision the REFERENCE's AlignmentStreamAnalyzer.step takes on the recorded streams (tests/golden/analyzer_cases.json,
    minted from the reference's own class) is reproduced by the CUDA sampler: forced EOS -> the stop token is emitted whatever the
    logits say; suppressed EOS -> never emitted even when it is the argmax; untouched -> plain argmax.  Greedy and sampling."""
    gold = json.load(open(os.path.join(GOLD, "analyzer_cases.json")))
    eos, V = gold["eos"], 8194
    g = torch.Generator().manual_seed(0)
    n_checked = {analyzer_ref.FORCE: 0, analyzer_ref.SUPPRESS: 0, None: 0}
    for case in gold["cases"]:
        toks, n_text, dec = case["tokens"], case["n_text"], case["decisions"]
        assert dec == [analyzer_ref.decision(i, n_text, toks) for i in range(len(toks) + 1)]       # the CPU restatement agrees too
        B = len(toks) + 1
        st = DecodeState(B, 4, 128, 1200, DEV)
        hist = torch.zeros(B, 128, dtype=torch.int32)
        for i in range(B):
            hist[i, :i] = torch.tensor(toks[:i], dtype=torch.int32)
        st.history.copy_(hist)
        st.hist_len.copy_(torch.arange(B, dtype=torch.int32))
        st.active.fill_(1)
        st.stop_tokens.fill_(eos)
        lg = torch.randn(B, V, generator=g)
        lg[:, eos] = 9.0                                           # the EOS is the argmax unless the analyzer says otherwise
        second = lg.clone()
        second[:, eos] = -1e9
        second = second.argmax(-1).tolist()
        for temp in (0.0, 0.8):
            s = SamplingTensors(B, DEV)
            s.temperature.fill_(temp)
            s.analyzer_text_tokens.fill_(n_text)
            s.seed.copy_(torch.arange(B, dtype=torch.int64) + 5)
            out = torch.empty(B, dtype=torch.int32, device=DEV)
            lg_d = lg.to(DEV)
            _lib.check(L.t3b_sample_logits(lg_d.data_ptr(), B, V, C.byref(s.c), C.byref(st.c), out.data_ptr(), sp()))
            got = out.cpu().tolist()
            for i in range(B):
                if dec[i] == analyzer_ref.FORCE:
                    assert got[i] == eos, (n_text, i, temp)
                elif dec[i] == analyzer_ref.SUPPRESS:
                    assert got[i] != eos and (temp > 0 or got[i] == second[i]), (n_text, i, temp)
                elif temp == 0.0:
                    assert got[i] == eos, (n_text, i)
                n_checked[dec[i]] += 1
    assert all(v > 50 for v in n_checked.values()), n_checked


# ---------------------------------------------------------------------------------------------- launch-shape options
@pytest.mark.parametrize("rows", [2, 4])
def test_gemv_norm_fold_equals_separate_norm_launches(L, small_model, rows):
    """t3b_options_t: add + RMSNorm folded into the qkv / gate_up GEMVs (default at <= 2 rows) is bit-identical to the 7-launch layer."""
    m, _ = small_model
    lens = [37, 37, 100, 100][:rows]
    outs = []
    for no_norm in (0, 1):
        kv, st = random_state(SMALL, rows, lens, seed=3)
        opt = _lib.make_options(gemv_max_rows=4, gemv_no_norm=no_norm)
        ws = m.workspace(rows)
        hid = torch.empty(rows, 1024, dtype=torch.bfloat16, device=DEV)
        _lib.check(L.t3b_decode_step(C.byref(m.c), C.byref(kv.c), C.byref(st.c), None, rows, hid.data_ptr(), ws.data_ptr(), ws.numel(),
                                     C.byref(opt), sp()))
        outs.append((hid.clone(), kv.k.clone(), kv.v.clone()))
    assert all(torch.equal(a, b) for a, b in zip(outs[0], outs[1]))


@pytest.mark.parametrize("rows", [2, 4, 6, 8, 12, 16])
def test_gemv_mma_decode_chain_matches_tcgen05_chain(L, small_model, rows):
    """t3b_options_t.gemv_mma: the 5-launch GEMV chain on the tensor cores (<= 16 rows, add + RMSNorm in front of qkv / gate_up) against
    the 7-launch tcgen05 chain.  Both accumulate in fp32, in different orders: layer 0's K / V come out of one projection (all but a
    few elements identical, the rest one bf16 step apart), the hidden state after the small model's layers agrees to a few bf16
    steps; the chain is deterministic, and a request's rows do not depend on the rows beside them (rows 0..1 alone == rows 0..1 of
    the batch, bit for bit)."""
    m, _ = small_model
    assert m.enable_gemv_mma()
    g = torch.Generator().manual_seed(rows)
    lens = torch.randint(5, 300, (rows // 2,), generator=g).repeat_interleave(2).tolist()
    outs = []
    for mode in ("mma", "umma", "mma", "mma2"):
        n = 2 if mode == "mma2" else rows
        kv, st = random_state(SMALL, rows, lens, seed=5)
        opt = _lib.make_options(gemv_max_rows=-1, norm_fusion=-1, attn_team_warps=4) if mode == "umma" else _lib.make_options(gemv_max_rows=16, gemv_mma=1, attn_team_warps=4)
        ws = m.workspace(rows)
        hid = torch.zeros(rows, 1024, dtype=torch.bfloat16, device=DEV)
        _lib.check(L.t3b_decode_step(C.byref(m.c), C.byref(kv.c), C.byref(st.c), None, n, hid.data_ptr(), ws.data_ptr(), ws.numel(),
                                     C.byref(opt), sp()))
        torch.cuda.synchronize()
        outs.append((hid.clone(), kv.k.clone(), kv.v.clone()))
    mma, umma, again, alone = outs
    assert all(torch.equal(a, b) for a, b in zip(mma, again))                          # deterministic
    assert torch.equal(alone[0][:2], mma[0][:2])                                       # independent of the rows beside it
    hf, hp = mma[0].float(), umma[0].float()
    assert torch.isfinite(hf).all()
    k0m, k0u = mma[1][0].float(), umma[1][0].float()                                   # layer 0 of the pool (viewed as raw bf16)
    assert (k0m == k0u).float().mean().item() > 0.99
    same = (hf == hp).float().mean().item()
    err = (hf - hp).abs().max().item() / hp.abs().max().item()
    assert same > 0.4 and err < 2 ** -5, (same, err)


@pytest.mark.parametrize("rows", [6, 16, 18, 32])
def test_norm_folded_into_projections_matches_separate_norm_launches(L, small_model, rows):
    """t3b_options_t.norm_fusion: at up to 32 rows the decode chain runs add + RMSNorm inside the o / down epilogues and the
    qkv / gate_up operand build (5 launches per layer).  The arithmetic is the norm kernel's except for the order in which a row's
    sum of squares is added (8 tile partials instead of 128 thread partials): rstd may move by one fp32 ulp, so the comparison with
    the 7-launch chain is 'all but a few elements identical, the rest one or two bf16 steps apart'.  Layer 0's K/V (written before any
    folded norm) must be bit-identical, the folded chain must be deterministic, and a row's result must not depend on the rows
    beside it."""
    m, _ = small_model
    g = torch.Generator().manual_seed(rows)
    lens = torch.randint(5, 300, (rows // 2,), generator=g).repeat_interleave(2).tolist()
    outs = []
    for mode in (1, -1, 1):
        kv, st = random_state(SMALL, rows, lens, seed=5)
        opt = _lib.make_options(gemv_max_rows=-1, norm_fusion=mode)
        ws = m.workspace(rows)
        hid = torch.empty(rows, 1024, dtype=torch.bfloat16, device=DEV)
        _lib.check(L.t3b_decode_step(C.byref(m.c), C.byref(kv.c), C.byref(st.c), None, rows, hid.data_ptr(), ws.data_ptr(), ws.numel(),
                                     C.byref(opt), sp()))
        torch.cuda.synchronize()
        outs.append((hid.clone(), kv.k.clone(), kv.v.clone()))
    fused, plain, again = outs
    assert all(torch.equal(a, b) for a, b in zip(fused, again))                       # deterministic
    assert torch.equal(fused[1][0], plain[1][0]) and torch.equal(fused[2][0], plain[2][0])   # layer 0 K / V
    hf, hp = fused[0].float(), plain[0].float()
    assert torch.isfinite(hf).all()
    same = (hf == hp).float().mean().item()
    err = (hf - hp).abs().max().item() / hp.abs().max().item()
    assert same > 0.8 and err < 2 ** -6, (same, err)


@pytest.mark.parametrize("rows,chunk,stages", [(64, 32, 0), (64, 16, 3), (100, 48, 4), (64, 64, 2)])
def test_umma_row_chunk_and_stage_options_do_not_change_results(L, rows, chunk, stages):
    """Row chunking / ring depth only move work between thread blocks: every output element keeps its K order -> bit-identical."""
    g = torch.Generator().manual_seed(rows + chunk)
    for N, K, epi in ((1024, 4096, 0), (3072, 1024, 0), (8192, 1024, 1)):
        x = torch.randn(rows, K, generator=g).bfloat16().to(DEV)
        w = (torch.randn(N, K, generator=g) * K ** -0.5).bfloat16().to(DEV)
        lin = _lib.Linear(_lib.T3B_W_BF16, N, K, w.data_ptr(), None, None)
        ys = []
        for opt in (None, _lib.make_options(umma_stages=stages, umma_row_chunk=chunk)):
            y = torch.zeros(rows, N // (2 if epi else 1), dtype=torch.bfloat16, device=DEV)
            split = 2 if chunk % 8 else 0
            _lib.check(L.t3b_linear_bf16(x.data_ptr(), C.byref(lin), rows, y.data_ptr(), None, split, None, _lib.T3B_IMPL_UMMA, epi,
                                         C.byref(opt) if opt is not None else None, sp()))
            ys.append(y)
        assert torch.equal(ys[0], ys[1]), (N, K)



def test_row_order_hint_does_not_change_results(L, small_model):
    """t3b_state_t.row_order only changes which warp serves which (row, head) item: a decode step with the rows dealt in a random
    order (and in longest-context-first order, what the engine passes) is bit-identical to the identity order - hidden states and
    the KV pool."""
    m, _ = small_model
    rows = 96
    g = torch.Generator().manual_seed(3)
    lens = torch.randint(5, 400, (rows // 2,), generator=g).repeat_interleave(2).tolist()
    outs = []
    for mode in ("identity", "random", "longest_first"):
        kv, st = random_state(SMALL, rows, lens, seed=8)
        if mode == "random":
            st.row_order.copy_(torch.randperm(rows, generator=g).to(torch.int32))
        elif mode == "longest_first":
            st.row_order.copy_(torch.tensor(sorted(range(rows), key=lambda r: -lens[r]), dtype=torch.int32))
        view = st.view(0, ordered=(mode != "identity"))
        ws = m.workspace(rows)
        hid = torch.empty(rows, 1024, dtype=torch.bfloat16, device=DEV)
        _lib.check(L.t3b_decode_step(C.byref(m.c), C.byref(kv.c), C.byref(view), None, rows, hid.data_ptr(), ws.data_ptr(), ws.numel(), None, sp()))
        torch.cuda.synchronize()
        outs.append((hid.clone(), kv.k.clone(), kv.v.clone()))
    for o in outs[1:]:
        assert all(torch.equal(a, b) for a, b in zip(outs[0], o))


def test_prefill_hidden_states_match_oracle_for_ragged_prompts(small_model):
    """t3.py:542-561 + the multi-token pass (t3b_prefill_embed -> t3b_forward_tokens with the read-once prefill attention): the
    final-norm hidden state of EVERY prompt token of ragged cond / uncond rows against oracle.forward_tokens (bf16 mode), and the
    K/V the pass leaves in the pool against the oracle's cache (what the decode steps will attend to)."""
    from chatterbox_vllm2_b200.model import KVPool
    m, sd = small_model
    n_texts = (1, 9, 40, 64, 17)
    reqs = [synthetic_request(20 + i, SMALL, n_text=n) for i, n in enumerate(n_texts)]
    slots = [2, 0, 4, 1, 3]
    x, tok_row, tok_idx, starts = m.prefill_assemble([c for c, _ in reqs], [ids for _, ids in reqs], slots)
    seq_start, seq_row = m.last_seq
    max_blocks = 8
    kv = KVPool(SMALL, 2 * len(reqs) * max_blocks + 1, DEV)
    bt = (torch.arange(2 * len(reqs) * max_blocks, dtype=torch.int32) + 1).reshape(2 * len(reqs), max_blocks).to(DEV)
    h = m.forward_tokens(kv, x, tok_row, tok_idx, tok_idx, bt, seq_start, seq_row, max(34 + n + 1 for n in n_texts))
    torch.cuda.synchronize()
    hs = h.float().cpu()
    orc = T3Oracle(sd, SMALL, mode="bf16")
    worst, cos_min = 0.0, 1.0
    kpool = kv.as_bf16("k").float().cpu()                    # [layer, block, head, 16, 64]
    for r, (cond, ids) in enumerate(reqs):
        c_ref, u_ref = orc.prefill_embeds(cond.float(), ids)
        P = c_ref.shape[0]
        rc, ru = orc.new_rows(2)
        for half, (emb, row) in enumerate(((c_ref, rc), (u_ref, ru))):
            ref = orc.forward_tokens(emb, [row] * P, range(P))
            got = hs[starts[2 * r + half]:starts[2 * r + half + 1]]
            worst = max(worst, float((got - ref).abs().max()))
            cos_min = min(cos_min, float(torch.nn.functional.cosine_similarity(got.flatten().double(), ref.flatten().double(), dim=0)))
            # K of the last layer at every position of this row
            kv_row = 2 * slots[r] + half
            blocks = bt[kv_row].cpu().long()
            k_got = kpool[SMALL.n_layers - 1][blocks].permute(0, 2, 1, 3).reshape(-1, 16, 64)[:P]
            k_ref = orc.kc[row][SMALL.n_layers - 1][:P]
            assert (k_got - k_ref).abs().max() <= 0.0625, (r, half, float((k_got - k_ref).abs().max()))
    from helpers import record
    record("prefill_hidden_ragged_L2", {"max_abs": worst, "min_cosine": cos_min, "prompts": list(n_texts)})
    assert worst <= 0.125 and cos_min >= 0.9999, (worst, cos_min)     # measured 0.0625 = one bf16 ulp at |h| in [8, 16)


@pytest.mark.parametrize("rows", [257, 512, 1000, 2500, 6336])
def test_persistent_big_gemm_equals_tiled_kernel(L, rows):
    """csrc/linear_big.cu (persistent blocks, 4-deep ring, double-buffered TMEM accumulator; the default for > 1024 rows without
    split-K): the same tcgen05 accumulation per output element as linear_umma_kernel, so y is bit-identical - ragged last row
    chunk, more tiles than blocks (several tiles per block, both accumulators reused), SiLU epilogue, the 8194-wide head."""
    g = torch.Generator().manual_seed(rows)
    for N, K, epi in ((3072, 1024, 0), (8192, 1024, 1), (1024, 4096, 0), (8194, 1024, 0)):
        x = torch.randn(rows, K, generator=g).bfloat16().to(DEV)
        w = (torch.randn(N, K, generator=g) * K ** -0.5).bfloat16().to(DEV)
        lin = _lib.Linear(_lib.T3B_W_BF16, N, K, w.data_ptr(), None, None)
        ys = []
        for opt in (_lib.make_options(umma_big_min_rows=-1), _lib.make_options(umma_big_min_rows=1)):
            y = torch.full((rows, N // (2 if epi else 1)), float("nan"), dtype=torch.bfloat16, device=DEV)
            _lib.check(L.t3b_linear_bf16(x.data_ptr(), C.byref(lin), rows, y.data_ptr(), None, 1, None, _lib.T3B_IMPL_UMMA, epi, C.byref(opt), sp()))
            ys.append(y)
        torch.cuda.synchronize()
        assert torch.isfinite(ys[1].float()).all() and torch.equal(ys[0], ys[1]), (N, K, epi)
