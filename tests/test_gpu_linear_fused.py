"""t3b_linear_fused: F2.1 (fused_add_rms_norm, vllm:ir/ops/layernorm.py:40-59) folded into the projections around
it - the normalised activation operand built inside the GEMM from the fp32 residual stream, and the residual add
(llama.py:316-333) in the GEMM epilogue - against a torch fp32 restatement with the reference's bf16 roundings, and
the whole decode step through it against the 7-launch path (T3B_FUSE_NORM=0 semantics: t3b_add_rmsnorm + t3b_linear_bf16).
Tolerances are written beside each assert (1 bf16 ulp = 2^-8 relative)."""
import ctypes as C

import pytest
import torch

from chatterbox_vllm2_b200 import _lib
from test_gpu_kernels import DEV, L, _check_linear, bf, sp, ulp_diff  # noqa: F401  (L is a fixture)

pytestmark = pytest.mark.gpu
EPS = 1e-5


def _stream(rows, K, seed):
    """A residual stream z (fp32 sums of two bf16 values, like the kernels produce), its per-128 sums of squares."""
    g = torch.Generator().manual_seed(seed)
    z = torch.randn(rows, K, generator=g).bfloat16().float() + torch.randn(rows, K, generator=g).bfloat16().float()
    ss = (z.double() ** 2).reshape(rows, K // 128, 128).sum(-1).t().contiguous().float()      # [K/128][rows]
    return z, ss


def _norm_ref(z, nw):
    """layernorm.py:40-59: x fp32, variance over the row, x * rsqrt(var + eps) -> bf16, * weight -> bf16."""
    rstd = torch.rsqrt((z.double() ** 2).mean(-1, keepdim=True) + EPS).float()
    return ((z * rstd).bfloat16().float() * nw.float()).bfloat16()


@pytest.mark.parametrize("rows", [6, 16, 50, 64, 128, 200, 300])
@pytest.mark.parametrize("N,K,split,epi", [(3072, 1024, 4, 0), (3072, 1024, 1, 0), (8192, 1024, 2, 1), (8192, 1024, 1, 1), (1024, 4096, 8, 0)])
def test_norm_fused_into_operand_build(L, rows, N, K, split, epi):
    z, ss = _stream(rows, K, rows + N + split)
    g = torch.Generator().manual_seed(7)
    nw = (1.0 + 0.1 * torch.randn(K, generator=g)).bfloat16()
    w = (torch.randn(N, K, generator=g) * K ** -0.5).bfloat16()
    zd, ssd, nwd, wd = z.to(DEV), ss.to(DEV), bf(nw), bf(w)
    n_out = N // 2 if epi else N
    y = torch.full((rows, n_out), float("nan"), dtype=torch.bfloat16, device=DEV)
    lin = _lib.Linear(_lib.T3B_W_BF16, N, K, wd.data_ptr(), None, None)
    f = _lib.LinearFuse(zd.data_ptr(), ssd.data_ptr(), K // 128, nwd.data_ptr(), EPS, None, None, None)
    _lib.check(L.t3b_linear_fused(None, C.byref(lin), rows, y.data_ptr(), C.byref(f), split, epi, sp()), "t3b_linear_fused")
    # the same GEMM fed with the separately normalised activations must agree bit for bit (same tiles, same order)
    xn = _norm_ref(z, nw)
    y2 = torch.empty_like(y)
    xd = bf(xn)
    _lib.check(L.t3b_linear_bf16(xd.data_ptr(), C.byref(lin), rows, y2.data_ptr(), None, split, None, _lib.T3B_IMPL_UMMA, epi, sp()))
    torch.cuda.synchronize()
    assert torch.isfinite(y.float()).all()
    # rstd from fp32 partial sums vs float64: a 1-ulp flip of a normalised activation is possible but rare
    assert (y != y2).float().mean().item() < 0.02, (y != y2).float().mean().item()
    d = ulp_diff(y, y2.float())
    assert (d <= 2.0).float().mean() > 0.999
    if not epi:
        _check_linear(y, (xn.float() @ w.float().t()).bfloat16().float(), K)


@pytest.mark.parametrize("rows", [6, 16, 64, 100, 128, 256, 300])
@pytest.mark.parametrize("N,K,split", [(1024, 1024, 8), (1024, 4096, 8), (1024, 4096, 2), (1024, 1024, 1)])
def test_residual_epilogue(L, rows, N, K, split):
    g = torch.Generator().manual_seed(rows + K + split)
    x = torch.randn(rows, K, generator=g).bfloat16()
    w = (torch.randn(N, K, generator=g) * K ** -0.5).bfloat16()
    z0, _ = _stream(rows, N, rows)
    xd, wd, zd = bf(x), bf(w), z0.to(DEV)
    z_keep = zd.clone()
    ssd = torch.full((N // 128, rows), float("nan"), device=DEV)
    lin = _lib.Linear(_lib.T3B_W_BF16, N, K, wd.data_ptr(), None, None)
    f = _lib.LinearFuse(None, None, 0, None, 0.0, zd.data_ptr(), zd.data_ptr(), ssd.data_ptr())          # in place
    _lib.check(L.t3b_linear_fused(xd.data_ptr(), C.byref(lin), rows, None, C.byref(f), split, _lib.T3B_EPI_RESIDUAL, sp()), "t3b_linear_fused")
    y = torch.empty(rows, N, dtype=torch.bfloat16, device=DEV)
    _lib.check(L.t3b_linear_bf16(xd.data_ptr(), C.byref(lin), rows, y.data_ptr(), None, split, None, _lib.T3B_IMPL_UMMA, 0, sp()))
    torch.cuda.synchronize()
    # bit-exact against the same GEMM's bf16 output + bf16(residual), summed in fp32
    ref = y.float() + z_keep.bfloat16().float()
    assert torch.equal(zd, ref)
    ss_ref = (ref.double() ** 2).reshape(rows, N // 128, 128).sum(-1).t()
    assert torch.allclose(ssd.double(), ss_ref, rtol=1e-5, atol=1e-6)


def test_stream_open_close(L):
    rows, D = 10, 1024
    g = torch.Generator().manual_seed(3)
    x = torch.randn(rows, D, generator=g).bfloat16()
    nw = (1.0 + 0.1 * torch.randn(D, generator=g)).bfloat16()
    xd, nwd = bf(x), bf(nw)
    z = torch.empty(rows, D, device=DEV)
    ss = torch.empty(D // 128, rows, device=DEV)
    y = torch.empty(rows, D, dtype=torch.bfloat16, device=DEV)
    _lib.check(L.t3b_z_init(xd.data_ptr(), rows, D, z.data_ptr(), ss.data_ptr(), sp()))
    _lib.check(L.t3b_z_rmsnorm(z.data_ptr(), ss.data_ptr(), nwd.data_ptr(), rows, D, EPS, y.data_ptr(), sp()))
    torch.cuda.synchronize()
    assert torch.equal(z.cpu(), x.float())
    assert torch.allclose(ss.cpu().double(), (x.double() ** 2).reshape(rows, 8, 128).sum(-1).t(), rtol=1e-5)
    ref = _norm_ref(x.float(), nw)
    assert (ulp_diff(y, ref.float()) <= 1.0).all() and (y.cpu() != ref).float().mean() < 0.01


def test_fused_rejects_bad_arguments(L):
    w = bf(torch.zeros(1024, 1024))
    lin = _lib.Linear(_lib.T3B_W_BF16, 1024, 1024, w.data_ptr(), None, None)
    f = _lib.LinearFuse(None, None, 0, None, 0.0, None, None, None)
    assert L.t3b_linear_fused(None, C.byref(lin), 8, w.data_ptr(), C.byref(f), 0, 0, sp()) == -1                        # no input
    assert L.t3b_linear_fused(w.data_ptr(), C.byref(lin), 8, None, C.byref(f), 0, _lib.T3B_EPI_RESIDUAL, sp()) == -1    # no stream
    assert L.t3b_linear_fused(w.data_ptr(), C.byref(lin), 8, w.data_ptr(), None, 0, 7, sp()) == -1                      # bad epilogue


@pytest.mark.parametrize("rows", [6, 16, 64, 130])
def test_decode_step_fused_norm_matches_seven_launch_path(L, rows):
    """The 5-launch layer (norms inside the projections) against the 7-launch layer on the same state: same arithmetic,
    only the order of the fp32 sum of squares differs -> hidden states agree to bf16 rounding noise."""
    from chatterbox_vllm2_b200 import T3Dims
    from chatterbox_vllm2_b200.model import T3B200Model
    from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint
    from test_gpu_mega import random_state
    dims = T3Dims().small(n_layers=2, max_pos=2048)
    m = T3B200Model(dims=dims, device=DEV, cfg_scale=0.5)
    m.load_weights(synthetic_checkpoint(dims).items())
    g = torch.Generator().manual_seed(rows)
    lens = torch.randint(40, 300, (rows,), generator=g).tolist()
    kv, st = random_state(dims, rows, lens, seed=rows)
    k0, v0 = kv.k.clone(), kv.v.clone()
    outs = []
    try:
        for mode in (0, 1, 1):
            L.t3b_debug_set_fuse_norm(mode)
            kv.k.copy_(k0); kv.v.copy_(v0)
            hid = torch.zeros(rows, dims.hidden, dtype=torch.bfloat16, device=DEV)
            m.decode_step(kv, st, rows, hid)
            torch.cuda.synchronize()
            outs.append((hid, kv.k.clone()))
    finally:
        L.t3b_debug_set_fuse_norm(-1)
    a, b = outs[0][0].float(), outs[1][0].float()
    assert torch.isfinite(b).all()
    cos = torch.nn.functional.cosine_similarity(a.double().flatten(), b.double().flatten(), dim=0).item()
    assert cos >= 0.9999 and (a - b).abs().max().item() <= 0.0625, (cos, (a - b).abs().max().item())
    assert (outs[0][1].float() - outs[1][1].float()).abs().max().item() <= 0.0625
    assert torch.equal(outs[1][0].view(torch.int16), outs[2][0].view(torch.int16))          # deterministic


@pytest.mark.parametrize("rows", [6, 16, 64, 100, 128, 200])
@pytest.mark.parametrize("K", [1024, 4096])
def test_tail_norm_is_bit_identical_to_separate_add_rmsnorm(L, rows, K):
    """The add + RMSNorm that follows the o / down projection, run by the last thread block of the GEMM (tail_* fields),
    against t3b_linear_bf16 followed by t3b_add_rmsnorm: same arithmetic in the same order -> the same bits."""
    N = 1024
    g = torch.Generator().manual_seed(rows + K)
    x = torch.randn(rows, K, generator=g).bfloat16()
    w = (torch.randn(N, K, generator=g) * K ** -0.5).bfloat16()
    res = torch.randn(rows, N, generator=g).bfloat16()
    nw = (1.0 + 0.1 * torch.randn(N, generator=g)).bfloat16()
    xd, wd, nwd = bf(x), bf(w), bf(nw)
    lin = _lib.Linear(_lib.T3B_W_BF16, N, K, wd.data_ptr(), None, None)
    # reference: two launches
    y_ref = torch.empty(rows, N, dtype=torch.bfloat16, device=DEV)
    res_ref, out_ref = bf(res), torch.empty(rows, N, dtype=torch.bfloat16, device=DEV)
    _lib.check(L.t3b_linear_bf16(xd.data_ptr(), C.byref(lin), rows, y_ref.data_ptr(), None, 0, None, _lib.T3B_IMPL_UMMA, 0, sp()))
    _lib.check(L.t3b_add_rmsnorm(y_ref.data_ptr(), None, 1, res_ref.data_ptr(), nwd.data_ptr(), out_ref.data_ptr(), rows, N, EPS, 1, sp()))
    # fused tail, twice (the counters must be left re-armed)
    counters = torch.zeros(64, dtype=torch.int32, device=DEV)
    for _ in range(2):
        y = torch.empty(rows, N, dtype=torch.bfloat16, device=DEV)
        res_d, out = bf(res), torch.full((rows, N), float("nan"), dtype=torch.bfloat16, device=DEV)
        f = _lib.LinearFuse(None, None, 0, None, EPS, None, None, None, res_d.data_ptr(), nwd.data_ptr(), out.data_ptr(), counters.data_ptr())
        _lib.check(L.t3b_linear_fused(xd.data_ptr(), C.byref(lin), rows, y.data_ptr(), C.byref(f), 0, _lib.T3B_EPI_NONE, sp()), "t3b_linear_fused")
        torch.cuda.synchronize()
        assert torch.equal(y.view(torch.int16), y_ref.view(torch.int16))
        assert torch.equal(res_d.view(torch.int16), res_ref.view(torch.int16))
        assert torch.equal(out.view(torch.int16), out_ref.view(torch.int16))
        assert int(counters.abs().sum()) == 0


@pytest.mark.parametrize("rows", [6, 16, 64, 130, 256])
def test_decode_step_tail_norm_equals_seven_launch_path(L, rows):
    from chatterbox_vllm2_b200 import T3Dims
    from chatterbox_vllm2_b200.model import T3B200Model
    from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint
    from test_gpu_mega import random_state
    dims = T3Dims().small(n_layers=3, max_pos=2048)
    m = T3B200Model(dims=dims, device=DEV, cfg_scale=0.5)
    m.load_weights(synthetic_checkpoint(dims).items())
    g = torch.Generator().manual_seed(rows)
    lens = torch.randint(40, 300, (rows,), generator=g).tolist()
    kv, st = random_state(dims, rows, lens, seed=rows)
    k0, v0 = kv.k.clone(), kv.v.clone()
    outs = []
    try:
        for mode in (0, 1, 1):
            L.t3b_debug_set_tail_norm(mode)
            kv.k.copy_(k0); kv.v.copy_(v0)
            hid = torch.zeros(rows, dims.hidden, dtype=torch.bfloat16, device=DEV)
            m.decode_step(kv, st, rows, hid)
            torch.cuda.synchronize()
            outs.append((hid, kv.k.clone(), kv.v.clone()))
    finally:
        L.t3b_debug_set_tail_norm(-1)
    for a, b in zip(outs[0], outs[1]):
        assert torch.equal(a.view(torch.int16), b.view(torch.int16))
    assert torch.equal(outs[1][0].view(torch.int16), outs[2][0].view(torch.int16))


@pytest.mark.parametrize("rows", [2, 4])
def test_small_batch_norm_in_gemv_equals_seven_launch_path(L, rows):
    """rows <= 4: add + RMSNorm run in front of the qkv / gate_up GEMVs (5 launches per layer) - the same arithmetic in the same
    order as t3b_add_rmsnorm followed by the GEMV, so hidden states and the K/V written this step are bit-identical."""
    from chatterbox_vllm2_b200 import T3Dims
    from chatterbox_vllm2_b200.model import T3B200Model
    from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint
    from test_gpu_mega import random_state
    dims = T3Dims().small(n_layers=3, max_pos=2048)
    m = T3B200Model(dims=dims, device=DEV, cfg_scale=0.5)
    m.load_weights(synthetic_checkpoint(dims).items())
    lens = [37, 37, 180, 180][:rows]
    kv, st = random_state(dims, rows, lens, seed=rows)
    k0, v0 = kv.k.clone(), kv.v.clone()
    outs = []
    try:
        for mode in (0, 1, 1):
            L.t3b_debug_set_gemv_norm(mode)
            kv.k.copy_(k0); kv.v.copy_(v0)
            hid = torch.zeros(rows, dims.hidden, dtype=torch.bfloat16, device=DEV)
            m.decode_step(kv, st, rows, hid)
            torch.cuda.synchronize()
            outs.append((hid, kv.k.clone(), kv.v.clone()))
    finally:
        L.t3b_debug_set_gemv_norm(-1)
    for a, b in zip(outs[0], outs[1]):
        assert torch.equal(a.view(torch.int16), b.view(torch.int16))
    assert torch.equal(outs[1][0].view(torch.int16), outs[2][0].view(torch.int16))
