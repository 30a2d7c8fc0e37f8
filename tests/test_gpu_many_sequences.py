"""Batches of MANY short sequences (hundreds of requests per step, as a serving scheduler produces them):
whole-sequence prefill and a decode step (one query row per sequence, keys from the paged cache) --
tests/test_gpu_parity.py's fuzz stops at 5 sequences per batch."""
import numpy as np
import pytest
import torch

from oracle import trianglemix_oracle as orc

pytestmark = pytest.mark.gpu
ATOL, RTOL = 1.5e-2, 1e-2
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ops():
    from vllm_trianglemix_b200 import ops as o
    return o


@pytest.mark.parametrize("kernel", ["auto", "mma", "tcgen05"])
@pytest.mark.parametrize("mode", [0, 1])
def test_prefill_600_sequences(ops, mode, kernel):
    rng = np.random.RandomState(600 + mode)
    lens = rng.randint(1, 200, size=600).tolist()
    lens[17], lens[300] = 2304, 513                  # two long ones in between (split-KV tail, streaming path)
    hq, hkv, si, sl, last = 8, 2, 4, 32, 64
    t = sum(lens)
    g = torch.Generator().manual_seed(t)
    q = torch.randn(t, hq, 128, generator=g).bfloat16()
    k = torch.randn(t, hkv, 128, generator=g).bfloat16()
    v = torch.randn(t, hkv, 128, generator=g).bfloat16()
    cu_list = [0] + np.cumsum(lens).tolist()
    cu = torch.tensor(cu_list, dtype=torch.int32, device=DEV)
    out = torch.full((t, hq, 128), float("nan"), dtype=torch.bfloat16, device=DEV)
    ops.prefill_attn(q.to(DEV), k.to(DEV), v.to(DEV), cu, max(lens), 0.088, si, sl, last, mode, out=out, kernel=kernel)
    ref = orc.attention_varlen_fp32(q, k, v, cu_list, 0.088, si, sl, last, mode)
    rep = orc.error_report(out.cpu(), ref, ATOL, RTOL)
    assert rep["violations"] == 0, (mode, kernel, rep)


@pytest.mark.parametrize("kernel", ["auto", "mma", "tcgen05"])
@pytest.mark.parametrize("mode", [0, 1])
def test_decode_step_500_sequences(ops, mode, kernel):
    """One new token per sequence, contexts of 1..3000 tokens behind a shuffled block table (page 16)."""
    rng = np.random.RandomState(500 + mode)
    seq_lens = rng.randint(1, 3000, size=500).tolist()
    hq, hkv, si, sl, last, page = 8, 2, 4, 32, 64, 16
    g = torch.Generator().manual_seed(sum(seq_lens))
    need = [(n + page - 1) // page for n in seq_lens]
    nb = sum(need) + 2
    perm = (torch.randperm(nb - 1, generator=g) + 1).tolist()
    cache = torch.randn(2, nb, page, hkv, 128, generator=g).bfloat16()       # unused slots hold garbage on purpose
    bt = torch.zeros(len(seq_lens), max(need) + 1, dtype=torch.int32)
    for b in range(len(seq_lens)):
        for j in range(need[b]):
            bt[b, j] = perm.pop()
    q = torch.randn(len(seq_lens), hq, 128, generator=g).bfloat16()
    cu_q = torch.arange(len(seq_lens) + 1, dtype=torch.int32, device=DEV)
    out = ops.chunked_prefill_attn(q.to(DEV), cu_q, torch.tensor(seq_lens, dtype=torch.int32, device=DEV), 1, 0.088,
                                   si, sl, last, mode, key_cache=cache[0].to(DEV), value_cache=cache[1].to(DEV),
                                   block_table=bt.to(DEV), kernel=kernel)
    out = out.cpu()
    for b in range(0, len(seq_lens), 7):                                     # every 7th sequence against the oracle
        n = seq_lens[b]
        pages = bt[b, :need[b]].long()
        kk = cache[0][pages].reshape(-1, hkv, 128)[:n]
        vv = cache[1][pages].reshape(-1, hkv, 128)[:n]
        qq = torch.zeros(n, hq, 128, dtype=torch.bfloat16)
        qq[n - 1] = q[b]
        ref = orc.attention_rows_fp32(qq, kk, vv, 0.088, n, si, sl, last, mode, n - 1, n)
        rep = orc.error_report(out[b:b + 1], ref, ATOL, RTOL)
        assert rep["violations"] == 0, (b, n, mode, kernel, rep)
