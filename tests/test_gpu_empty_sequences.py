"""Ragged batches with EMPTY sequences (cu_seqlens with repeated offsets) and degenerate calls.

The reference op takes (B, S, H, D) and has no varlen form; vLLM's varlen layout (query_start_loc,
vllm/v1/attention/backends/flash_attn.py:110-140) allows a request to contribute no rows.  An empty sequence
must neither disturb its neighbours (its slot of the block index stays a padding slot) nor touch memory."""
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


@pytest.mark.parametrize("kernel", ["auto", "mma", "tcgen05", "pair"])
@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("lens", [[0, 700], [700, 0], [0, 0, 2304, 0, 129, 0], [1, 0, 1], [300, 0, 0, 0, 2500]])
def test_empty_sequences_in_a_batch(ops, lens, mode, kernel):
    hq, hkv, si, sl, last = 8, 2, 4, 32, 64
    t = sum(lens)
    g = torch.Generator().manual_seed(t + mode)
    q = torch.randn(t, hq, 128, generator=g).bfloat16()
    k = torch.randn(t, hkv, 128, generator=g).bfloat16()
    v = torch.randn(t, hkv, 128, generator=g).bfloat16()
    cu_list = [0]
    for n in lens:
        cu_list.append(cu_list[-1] + n)
    cu = torch.tensor(cu_list, dtype=torch.int32, device=DEV)
    out = torch.full((t, hq, 128), float("nan"), dtype=torch.bfloat16, device=DEV)
    ops.prefill_attn(q.to(DEV), k.to(DEV), v.to(DEV), cu, max(lens), 0.088, si, sl, last, mode, out=out, kernel=kernel)
    if mode == 2:
        ref = torch.empty(t, hq, 128)
        for b, n in enumerate(lens):
            if n:
                s0 = cu_list[b]
                ref[s0:s0 + n] = orc.attention_fp32(q[None, s0:s0 + n], k[None, s0:s0 + n], v[None, s0:s0 + n], None, 0.088)[0]
    else:
        ref = orc.attention_varlen_fp32(q, k, v, cu_list, 0.088, si, sl, last, mode)
    rep = orc.error_report(out.cpu(), ref, ATOL, RTOL)
    assert rep["violations"] == 0, (lens, mode, kernel, rep)      # (NaN left in `out` = a row nobody wrote)
    assert not torch.isnan(out.float()).any()


def test_all_sequences_empty_and_no_sequences(ops):
    q = torch.empty(0, 8, 128, dtype=torch.bfloat16, device=DEV)
    k = torch.empty(0, 2, 128, dtype=torch.bfloat16, device=DEV)
    for cu in ([0, 0, 0], [0]):
        out = ops.prefill_attn(q, k, k, torch.tensor(cu, dtype=torch.int32, device=DEV), 0, 0.088, 4, 32, 64, 1)
        assert out.shape == (0, 8, 128)
    torch.cuda.synchronize()
