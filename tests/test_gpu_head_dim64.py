"""GPU parity at head_dim 64: the shape of the reference's only in-tree call of the op
(examples/trianglemix_quickstart.py:149-188: B 2, S 32, 8 heads, D 64, fp32, no mask) and masked GQA cases,
against outputs of the unmodified reference (tests/golden/attn_d64.npz) and the oracle.  D = 64 is served by
the mma.sync kernel (TMX_KERNEL_AUTO picks it; the tcgen05 kernels are built for Qwen3's D = 128 and say so).
Tolerance as in test_gpu_parity.py."""
import os
import warnings

import numpy as np
import pytest
import torch

from oracle import trianglemix_oracle as orc

pytestmark = pytest.mark.gpu
ATOL, RTOL = 1.5e-2, 1e-2
DEV = "cuda:0"
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "attn_d64.npz"))


@pytest.fixture(scope="module")
def ops():
    from vllm_trianglemix_b200 import ops as o
    return o


def _bf16(a):
    return torch.from_numpy(a.copy()).view(torch.bfloat16)


def _check(out, ref, what):
    rep = orc.error_report(out, ref, ATOL, RTOL)
    assert rep["violations"] == 0, (what, rep)


def test_quickstart_demo_through_the_reference_op_signature():
    """examples/trianglemix_quickstart.py example 6 with the tensors moved to the GPU: same constructor arguments,
    fp32 inputs, no attn_mask => bidirectional attention at seq_len <= 2048 (reference :96-97)."""
    import vllm_trianglemix_b200 as tmx
    cfg = tmx.TriangleMixConfig(num_sink_tokens=4, sliding_window_size=8, num_last_tokens=16, num_triangle_layers=1)
    attn = tmx.AscendTriangleMixAttention(num_heads=8, head_dim=64, scale=1.0 / (64 ** 0.5), trianglemix_config=cfg,
                                          layer_idx=0, num_layers=4)
    q, k, v = (_bf16(GOLD["demo_" + n]).float().to(DEV) for n in "qkv")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")            # "fp32 inputs are cast to bf16"
        out = attn(q, k, v)
    assert out.shape == (2, 32, 8, 64) and out.dtype == torch.float32 and out.is_contiguous()
    _check(out.cpu(), torch.from_numpy(GOLD["demo_out"]), "demo")


@pytest.mark.parametrize("kernel", ["auto", "mma"])
@pytest.mark.parametrize("tag", ["t", "u"])
def test_attention_golden_d64(ops, tag, kernel):
    b, s, hq, hkv, d, si, sl, last = [int(x) for x in GOLD[tag + "_params"]]
    q, k, v = (_bf16(GOLD["%s_%s" % (tag, n)]) for n in "qkv")
    cu = torch.arange(0, (b + 1) * s, s, dtype=torch.int32, device=DEV)
    for name, mode in (("tri", ops.TRIANGLE), ("dense", ops.DENSE_CAUSAL)):
        out = ops.prefill_attn(q.reshape(b * s, hq, d).to(DEV), k.reshape(b * s, hkv, d).to(DEV),
                               v.reshape(b * s, hkv, d).to(DEV), cu, s, d ** -0.5, si, sl, last, mode, kernel=kernel)
        _check(out.cpu().view(b, s, hq, d), torch.from_numpy(GOLD["%s_%s_out" % (tag, name)]), (tag, name, kernel))


@pytest.mark.parametrize("lens,hq,hkv,si,sl,last", [
    ([2304], 8, 2, 4, 32, 64),                 # > 2048: split-KV dense tail of the mma.sync path + merge
    ([700, 129, 1, 64, 2500], 4, 4, 4, 32, 64),
    ([333, 4097], 6, 2, 20, 150, 300),         # GQA ratio 3, off-default triangle
    ([5000], 2, 1, 0, 0, 1),
])
@pytest.mark.parametrize("mode", [0, 1, 2])
def test_attention_varlen_d64_vs_oracle(ops, lens, hq, hkv, si, sl, last, mode):
    t = sum(lens)
    g = torch.Generator().manual_seed(t + hq + mode)
    q = torch.randn(t, hq, 64, generator=g).bfloat16()
    k = torch.randn(t, hkv, 64, generator=g).bfloat16()
    v = torch.randn(t, hkv, 64, generator=g).bfloat16()
    cu_list = [0] + np.cumsum(lens).tolist()
    cu = torch.tensor(cu_list, dtype=torch.int32, device=DEV)
    out, lse = ops.prefill_attn(q.to(DEV), k.to(DEV), v.to(DEV), cu, max(lens), 0.125, si, sl, last, mode,
                                return_lse=True)
    if mode == 2:      # TMX_MODE_NONE: bidirectional, every key of the own sequence
        ref = torch.empty(t, hq, 64)
        for b in range(len(lens)):
            s0, s1 = cu_list[b], cu_list[b + 1]
            ref[s0:s1] = orc.attention_fp32(q[None, s0:s1], k[None, s0:s1], v[None, s0:s1], None, 0.125)[0]
    else:
        ref = orc.attention_varlen_fp32(q, k, v, cu_list, 0.125, si, sl, last, mode)
    _check(out.cpu(), ref, (lens, hq, hkv, mode))
    assert torch.isfinite(lse).all()


def test_d64_bitmask_and_errors(ops):
    from vllm_trianglemix_b200._lib import TmxError
    n, hq, hkv = 777, 4, 2
    g = torch.Generator().manual_seed(5)
    q = torch.randn(n, hq, 64, generator=g).bfloat16().to(DEV)
    k = torch.randn(n, hkv, 64, generator=g).bfloat16().to(DEV)
    v = torch.randn(n, hkv, 64, generator=g).bfloat16().to(DEV)
    cu = torch.tensor([0, n], dtype=torch.int32, device=DEV)
    a = ops.prefill_attn(q, k, v, cu, n, 0.125, 4, 32, 64, ops.TRIANGLE)
    bits = ops.mask_bits(n, 4, 32, 64, "triangle", DEV)
    b = ops.prefill_attn(q, k, v, cu, n, 0.125, 4, 32, 64, ops.TRIANGLE, bitmask=bits)
    assert torch.equal(a, b)                       # predicate path == bitmask path
    # the tcgen05 kernels are built for head_dim 128: an explicit request is refused, never mis-served
    for kernel in ("tcgen05", "pair"):
        with pytest.raises(TmxError):
            ops.prefill_attn(q, k, v, cu, n, 0.125, 4, 32, 64, ops.TRIANGLE, kernel=kernel)
    with pytest.raises(TmxError):                  # any other head_dim stays a shape error
        ops.prefill_attn(q[..., :32].contiguous(), k[..., :32].contiguous(), v[..., :32].contiguous(), cu, n,
                         0.125, 4, 32, 64, ops.TRIANGLE)


@pytest.mark.parametrize("mode", [0, 1])
def test_context_rows_d64_contiguous_and_paged(ops, mode):
    """Chunked prefill / decode rows at D = 64: keys from a contiguous buffer and from a paged cache give the same
    bits, and match the oracle's rows of the full sequence (as tests/test_gpu_chunked.py does at D = 128)."""
    seq_lens, q_lens, hq, hkv, si, sl, last, page = [1000, 300, 2500], [200, 300, 1], 8, 2, 4, 32, 64, 16
    g = torch.Generator().manual_seed(64 + mode)
    seqs = [(torch.randn(n, hq, 64, generator=g).bfloat16(), torch.randn(n, hkv, 64, generator=g).bfloat16(),
             torch.randn(n, hkv, 64, generator=g).bfloat16()) for n in seq_lens]
    cu = lambda lens: [0] + np.cumsum(lens).tolist()
    q = torch.cat([s[0][-ql:] for s, ql in zip(seqs, q_lens)]).to(DEV)
    cu_q = torch.tensor(cu(q_lens), dtype=torch.int32, device=DEV)
    lens = torch.tensor(seq_lens, dtype=torch.int32, device=DEV)
    k = torch.cat([s[1] for s in seqs]).to(DEV)
    v = torch.cat([s[2] for s in seqs]).to(DEV)
    out = ops.chunked_prefill_attn(q, cu_q, lens, max(q_lens), 0.125, si, sl, last, mode, k=k, v=v,
                                   cu_seqlens_k=torch.tensor(cu(seq_lens), dtype=torch.int32, device=DEV))
    # paged copy of the same keys behind a shuffled block table (page 0 unused)
    need = [(n + page - 1) // page for n in seq_lens]
    nb = sum(need) + 4
    perm = (torch.randperm(nb - 1, generator=g) + 1).tolist()
    cache = torch.zeros(2, nb, page, hkv, 64, dtype=torch.bfloat16)
    bt = torch.zeros(len(seqs), max(need) + 2, dtype=torch.int32)
    for b, (_, kk, vv) in enumerate(seqs):
        for j in range(need[b]):
            pg = perm.pop()
            bt[b, j] = pg
            rows = slice(j * page, min(seq_lens[b], (j + 1) * page))
            cache[0, pg, :rows.stop - rows.start] = kk[rows]
            cache[1, pg, :rows.stop - rows.start] = vv[rows]
    cache, bt = cache.to(DEV), bt.to(DEV)
    paged = ops.chunked_prefill_attn(q, cu_q, lens, max(q_lens), 0.125, si, sl, last, mode, key_cache=cache[0],
                                     value_cache=cache[1], block_table=bt)
    assert torch.equal(out, paged)
    cq = cu(q_lens)
    for b, ((qq, kk, vv), ql) in enumerate(zip(seqs, q_lens)):
        n = seq_lens[b]
        ref = orc.attention_rows_fp32(qq, kk, vv, 0.125, n, si, sl, last, mode, n - ql, n)
        _check(out[cq[b]:cq[b + 1]].cpu(), ref, ("context", b, mode))
