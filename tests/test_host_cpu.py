"""CPU tests of the host side: the C-ABI library loads and exports what include/t3b200.h declares (no compute
calls), the product path fails loudly without a GPU, the paged-KV allocator and request sharding behave, the
replica layout works across 2 processes (gloo), and the committed golden fixtures reproduce from the oracle."""
import ctypes as C
import os
import re
import socket
import subprocess
import sys
import textwrap

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "chatterbox-vllm2_b200", "csrc", "libt3b200.so")


@pytest.fixture(scope="module")
def built():
    if not os.path.exists(LIB):
        import __graft_entry__
        __graft_entry__.build()
    return LIB


def test_library_exports_every_header_symbol(built):
    from chatterbox_vllm2_b200 import _lib
    L = C.CDLL(built)
    names = _lib.header_symbols()
    hdr = open(os.path.join(ROOT, "include", "t3b200.h")).read()
    declared = set(re.findall(r"\b(t3b_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(names) and len(names) >= 20
    for n in names:
        assert hasattr(L, n), n
    L.t3b_version.restype = C.c_int32
    assert L.t3b_version() >= 100


def test_header_cites_the_reference_for_every_entry_point():
    hdr = open(os.path.join(ROOT, "include", "t3b200.h")).read()
    assert "t3.py:" in hdr and "llama.py:" in hdr and "sampler.py:" in hdr and "awq.py:" in hdr


def test_argument_errors_are_reported_without_a_gpu(built):
    from chatterbox_vllm2_b200 import _lib
    L = _lib.lib()
    rc = L.t3b_embed_gather(None, None, None, None, 1, 1024, 8194, 4100, 1, None, None)
    assert rc == -1 and b"null" in L.t3b_last_error().lower()
    with pytest.raises(_lib.T3BLibraryError):
        _lib.check(rc, "t3b_embed_gather")


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_product_path_has_no_cpu_fallback(built):
    from chatterbox_vllm2_b200 import T3Dims, _lib
    from chatterbox_vllm2_b200.model import T3B200Model
    from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint
    dims = T3Dims().small(n_layers=1, max_pos=64)
    m = T3B200Model(dims=dims, device="cpu")
    m.load_weights(synthetic_checkpoint(dims).items())
    with pytest.raises(_lib.T3BLibraryError):
        m.embed("speech_emb", "speech_pos_emb", torch.zeros(1, dtype=torch.int32), torch.zeros(1, dtype=torch.int32))
    assert _lib.lib().t3b_check_device() != 0


def test_kv_pool_allocator():
    from chatterbox_vllm2_b200 import T3Dims
    from chatterbox_vllm2_b200.model import KVPool
    dims = T3Dims().small(n_layers=1, max_pos=64)
    kv = KVPool(dims, 9, "cpu")
    assert KVPool.bytes_per_block(T3Dims()) == 16 * 122880
    a = kv.alloc(5)
    assert 0 not in a and len(set(a)) == 5            # block 0 is the scratch block
    with pytest.raises(MemoryError):
        kv.alloc(4)
    b = kv.alloc(3)
    assert not set(a) & set(b)
    kv.release(a)
    assert sorted(kv.alloc(5)) == sorted(a)


def test_shard_requests_partitions():
    from chatterbox_vllm2_b200.engine import shard_requests
    for n, w in ((4096, 8), (7, 2), (3, 4), (0, 2)):
        parts = [shard_requests(n, w, r) for r in range(w)]
        assert sorted(i for p in parts for i in p) == list(range(n))
        assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_replicas_world_size_2_gloo(tmp_path):
    """Two processes, gloo: each takes its shard of 7 requests, "decodes" them with the CPU oracle (standing in
    for the per-GPU engine: the multi-GPU layout has no data-path collective) and rank 0 gathers token lists
    in request order, exactly as the replica dispatcher does."""
    script = tmp_path / "rank.py"
    script.write_text(textwrap.dedent(f"""
        import os, sys
        sys.path.insert(0, {ROOT!r}); sys.path.insert(0, os.path.join({ROOT!r}, "tests"))
        import torch, torch.distributed as dist
        from chatterbox_vllm2_b200 import T3Dims
        from chatterbox_vllm2_b200.engine import shard_requests
        from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint, synthetic_request
        from helpers import oracle_generate
        from oracle.t3_ref import T3Oracle
        dist.init_process_group("gloo")
        rank, world = dist.get_rank(), dist.get_world_size()
        dims = T3Dims().small(n_layers=1, max_pos=128)
        orc = T3Oracle(synthetic_checkpoint(dims), dims, mode="fp32")
        mine = shard_requests(7, world, rank)
        out = oracle_generate(orc, [synthetic_request(i, dims, n_text=6) for i in mine], 4)["tokens"].tolist()
        gathered = [None] * world
        dist.all_gather_object(gathered, dict(zip(mine, out)))
        if rank == 0:
            merged = {{}}
            for g in gathered: merged.update(g)
            ref = oracle_generate(orc, [synthetic_request(i, dims, n_text=6) for i in range(7)], 4)["tokens"].tolist()
            assert [merged[i] for i in range(7)] == ref, (merged, ref)
            print("REPLICAS_OK")
        dist.barrier(); dist.destroy_process_group()
    """))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="2")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", str(_free_port()), str(script)], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "REPLICAS_OK" in r.stdout, r.stdout + r.stderr


def test_golden_greedy_stream_reproduces_from_oracle():
    """Re-run the oracle (2-layer model, 3 requests, 16 greedy steps) and require the committed fixture: teacher-forced with the
    fixture's tokens, its top-8 logits to 2 bf16 ulps (this host's BLAS may order fp32 sums differently from the build container's) and
    its tokens wherever its own top-2 margin exceeds the bf16 noise floor; the sampled fixture the same way."""
    import torch
    from helpers import GOLD, oracle_generate
    from chatterbox_vllm2_b200 import T3Dims
    from chatterbox_vllm2_b200.synthetic import synthetic_checkpoint, synthetic_request
    from oracle.t3_ref import T3Oracle
    g = np.load(os.path.join(GOLD, "t3_small_greedy.npz"), allow_pickle=False)
    small = T3Dims().small(n_layers=2, max_pos=512)
    sd = synthetic_checkpoint(small)
    prompts = [synthetic_request(i, small, n_text=12) for i in range(3)]
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    ref = oracle_generate(T3Oracle(sd, small, mode="bf16"), prompts, 16, rho=2.0, forced=g["tokens"])
    at = np.take_along_axis(ref["logits"], g["top_idx"].astype(np.int64), -1)
    # the oracle rounds to bf16 where the kernels round, so another host's fp32 summation order shows up as whole bf16 ulps
    # (0.03125 at |l| in [4, 8)): at most 2 ulps, and most values identical
    assert np.abs(at - g["top_val"]).max() <= 0.0625 and (at == g["top_val"]).mean() > 0.5
    bad = [(b, t) for b in range(3) for t in range(16) if ref["tokens"][b, t] != g["tokens"][b, t] and g["margins"][t, b] > 0.125]
    assert not bad, bad
    assert (ref["tokens"] == g["tokens"]).mean() > 0.9
    s = np.load(os.path.join(GOLD, "t3_small_sampled.npz"), allow_pickle=False)
    rs = oracle_generate(T3Oracle(sd, small, mode="bf16"), prompts, 16, rho=2.0, temperature=0.8, min_p=0.05, top_p=0.8,
                         seeds=[42, 43, 44], forced=s["tokens"])
    assert (rs["tokens"] == s["tokens"]).mean() > 0.9


def test_bench_reference_arm_contract():
    """`bench.py --impl reference` prints one JSON line with the keys the driver reads (tiny run)."""
    import json
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--batch", "1"],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "ms_per_step", "cpu_baseline", "e2e", "config"):
        assert k in line
    assert line["impl"] == "reference" and line["cpu_baseline"]["kind"] == "port" and line["value"] > 0


def test_dispatcher_least_outstanding_kv_tokens():
    """whole requests go to the replica with the least outstanding KV tokens; static and online forms."""
    from chatterbox_vllm2_b200.dispatch import ReplicaDispatcher, assign_requests, kv_token_cost, kv_traffic_cost
    import random
    rng = random.Random(7)
    lens = [rng.randint(100, 1000) for _ in range(4096)]            # BASELINE.json configs[4]
    costs = [kv_traffic_cost(99, n) for n in lens]
    for world in (1, 2, 4, 8):
        shares = assign_requests(costs, world)
        assert sorted(i for s in shares for i in s) == list(range(4096))             # a partition, nothing lost or duplicated
        loads = [sum(costs[i] for i in s) for s in shares]
        assert max(loads) - min(loads) <= max(costs)                                 # greedy bound: within one request of each other
        assert all(s == sorted(s) for s in shares)                                   # arrival order kept inside a replica
    assert kv_token_cost(99, 300) == 2 * 399 and kv_traffic_cost(99, 2) == 2 * (2 * 99 + 3)
    d = ReplicaDispatcher(3)
    where = [d.submit(i, 99, n) for i, n in enumerate([1000, 100, 100, 100, 100])]
    assert where[:3] == [0, 1, 2] and where[3] in (1, 2) and where[4] in (1, 2) and where[3] != where[4]
    d.complete(0)
    assert d.outstanding()[0] == 0 and d.submit(9, 99, 10) == 0                      # the drained replica takes the next request
    probed = ReplicaDispatcher(2, probe=lambda r: [500, 10][r])                      # true loads reported by the engines win
    assert probed.submit(0, 99, 100) == 1


def test_vllm_plugin_registers_architecture_and_s3gen_filter():
    """models/t3/__init__.py:5 registers "ChatterboxT3"; the shim registers the same architecture name with the installed vLLM (lazy
    form: no CUDA needed) and answers to the 0.10 and the 0.22 hook names.  tts.py:492-514: the S3Gen hand-off filter."""
    from chatterbox_vllm2_b200 import SPEECH_TOKEN_OFFSET, vllm_plugin
    from vllm import ModelRegistry
    vllm_plugin.register()
    assert "ChatterboxT3" in ModelRegistry.get_supported_archs()
    for hook in ("load_weights", "get_multimodal_embeddings", "embed_multimodal", "get_input_embeddings", "embed_input_ids", "forward",
                 "compute_logits", "get_language_model"):
        assert callable(getattr(vllm_plugin.T3B200ForVllm, hook))
    ids = [SPEECH_TOKEN_OFFSET + t for t in (0, 5, 6560, 6561, 6562, 8193)] + [17, 2499]
    assert vllm_plugin.speech_tokens_for_s3gen(ids) == [0, 5, 6560]


def test_gemv_weight_loads_stay_in_front_of_the_dependency_wait():
    """The GEMV kernels request their whole weight slice BEFORE griddepcontrol.wait (weights never depend on the previous kernel).
    ptxas is free to hoist the wait above ``ld.global.nc`` loads - it did so in one build of the tensor-core GEMV, which cost 0.9 us
    per launch (profiles/r2_notes.md r5) - and only the control flow between the loads and the wait keeps the order.  Guard it in
    the shipped binary: in every GEMV kernel the first ACQBULK (the wait) comes after all of the kernel's streaming weight loads
    (LDG.E.NA = ld.global.nc.L1::no_allocate)."""
    import re
    import shutil
    import subprocess
    from chatterbox_vllm2_b200 import _lib
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    so = _lib.lib_path() if hasattr(_lib, "lib_path") else os.path.join(ROOT, "chatterbox-vllm2_b200", "csrc", "libt3b200.so")
    sass = subprocess.run([cuobjdump, "-sass", so], capture_output=True, text=True, check=True).stdout
    fn, seen = None, {}
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1) if "linear_gemv" in m.group(1) else None
            if fn:
                seen[fn] = [0, 0, False]          # weight loads before the wait, after it, wait seen
            continue
        if fn is None:
            continue
        if "ACQBULK" in line:
            seen[fn][2] = True
        elif "LDG.E.NA" in line:
            seen[fn][1 if seen[fn][2] else 0] += 1
    assert len(seen) >= 30, len(seen)
    for fn, (pre, post, wait) in seen.items():
        assert wait and pre >= 2, (fn, pre, post)
        if "mma" in fn or fn.endswith("Lb1EEvPK13__nv_bfloat16S2_iiiPS0_8GemvNorm6PfArgs"):
            assert post == 0, (fn, pre, post)      # single-shot kernels: the whole slice is requested before the wait


def test_ctypes_mirrors_match_the_header_layout(tmp_path):
    """The Python host side mirrors the C structs of include/t3b200.h field by field (chatterbox_vllm2_b200/_lib.py).  Compile a
    C program against the header with gcc and compare sizeof / offsetof of the last field of every struct with the ctypes
    mirrors - a field added on one side only would shift every argument behind it."""
    import ctypes as C
    import shutil
    import subprocess
    from chatterbox_vllm2_b200 import _lib
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("gcc not available")
    pairs = [("t3b_linear_t", _lib.Linear, "flags"), ("t3b_layer_t", _lib.Layer, "ln2"), ("t3b_model_t", _lib.Model, None),
             ("t3b_kv_t", _lib.KV, None), ("t3b_state_t", _lib.State, "row_order"), ("t3b_sampling_t", _lib.Sampling, "combine"),
             ("t3b_options_t", _lib.Options, "gemv_mma")]
    lines = ["#include <stdio.h>", "#include <stddef.h>", '#include "t3b200.h"', "int main(void) {"]
    for cname, _, last in pairs:
        lines.append(f'  printf("{cname} %zu %zu\\n", sizeof({cname}), ' + (f"offsetof({cname}, {last})" if last else "(size_t)0") + ");")
    lines += ["  return 0;", "}"]
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run([gcc, "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()
    got = {out[i]: (int(out[i + 1]), int(out[i + 2])) for i in range(0, len(out), 3)}
    for cname, mirror, last in pairs:
        assert got[cname][0] == C.sizeof(mirror), (cname, got[cname][0], C.sizeof(mirror))
        if last:
            assert got[cname][1] == getattr(mirror, last).offset, (cname, last)
