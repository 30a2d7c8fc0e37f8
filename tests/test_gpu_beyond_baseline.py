"""GPU parity BEYOND BASELINE.json's largest size (128K tokens): maximum-size cases whose element
offsets cross 2^31 (1M tokens x 32 heads x 128 = 2^32 elements of q / out; a 256K bitmask is
2^31 words), checked against the oracle's row form on sampled row ranges and against the
closed-form pair counts.  The reference itself cannot run any of these (its N x N fp32 mask
would be 4 TiB at 1M tokens, trianglemix_attention.py:53-91); the oracle evaluates the same
arithmetic for chosen rows against all keys.  Tolerance as in test_gpu_parity.py."""
import pytest
import torch

from oracle import trianglemix_oracle as orc

pytestmark = pytest.mark.gpu
ATOL, RTOL = 1.5e-2, 1e-2
DEV = "cuda:0"
SI, SL, LAST = 4, 32, 64
SCALE = 128 ** -0.5


@pytest.fixture(scope="module")
def ops():
    from vllm_trianglemix_b200 import ops as o
    return o


def _rand_dev(t, hq, hkv, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    q = torch.randn(t, hq, 128, generator=g, device=DEV, dtype=torch.float32).bfloat16()
    k = torch.randn(t, hkv, 128, generator=g, device=DEV, dtype=torch.float32).bfloat16()
    v = torch.randn(t, hkv, 128, generator=g, device=DEV, dtype=torch.float32).bfloat16()
    return q, k, v


def _check_ranges(out, q, k, v, n, mode, kvh, ranges, what):
    """Row ranges of the one sequence [0, n) for KV head kvh (and its query heads) against the oracle."""
    g = q.shape[1] // k.shape[1]
    qs = q[:, kvh * g:(kvh + 1) * g].cpu()
    ks = k[:, kvh:kvh + 1].cpu()
    vs = v[:, kvh:kvh + 1].cpu()
    for r0, r1 in ranges:
        ref = orc.attention_rows_fp32(qs, ks, vs, SCALE, n, SI, SL, LAST, mode, r0, r1)
        got = out[r0:r1, kvh * g:(kvh + 1) * g].cpu()
        rep = orc.error_report(got, ref, ATOL, RTOL)
        assert rep["violations"] == 0, (what, "kv head", kvh, "rows", (r0, r1), rep)


def test_triangle_1m_tokens(ops):
    """One 1,048,576-token sequence, Qwen3-8B heads, triangle layer: q and out hold 2^32 elements each."""
    n, hq, hkv = 1 << 20, 32, 8
    q, k, v = _rand_dev(n, hq, hkv, seed=101)
    cu = torch.tensor([0, n], dtype=torch.int32, device=DEV)
    out = ops.prefill_attn(q, k, v, cu, n, SCALE, SI, SL, LAST, 1)
    torch.cuda.synchronize()
    lrow = n - LAST + 1
    # first rows (sink / window boundaries, first tiles), the 2^31-element boundary of q (row n/2), the rows
    # around the first dense row, and the last rows of the dense tail (which attend all 1M keys)
    ranges = [(0, 40), (126, 131), (n // 2 - 2, n // 2 + 2), (lrow - 2, lrow + 2), (n - 3, n)]
    _check_ranges(out, q, k, v, n, 1, hkv - 1, ranges, "triangle 1M")
    _check_ranges(out, q, k, v, n, 1, 0, [(n // 2 - 1, n // 2 + 1), (n - 1, n)], "triangle 1M")
    # rows i <= sl of a triangle layer attend exactly the causal prefix: identical to a short dense call
    short = ops.prefill_attn(q[:128], k[:128], v[:128], torch.tensor([0, 128], dtype=torch.int32, device=DEV),
                             128, SCALE, SI, SL, LAST, 0)
    rep = orc.error_report(out[:SL + 1].cpu(), short[:SL + 1].cpu().float(), ATOL, RTOL)
    assert rep["violations"] == 0, rep


def test_dense_512k_tokens_one_kv_head(ops):
    """One 524,288-token sequence with the per-rank heads of Qwen3-8B at P = 8 (4 q + 1 kv): 4097 KV tiles
    per query tile at the end of the sequence, 0.28 PFLOP."""
    n, hq, hkv = 1 << 19, 4, 1
    q, k, v = _rand_dev(n, hq, hkv, seed=102)
    cu = torch.tensor([0, n], dtype=torch.int32, device=DEV)
    out = ops.prefill_attn(q, k, v, cu, n, SCALE, SI, SL, LAST, 0)
    torch.cuda.synchronize()
    ranges = [(0, 3), (127, 130), (n // 2 - 1, n // 2 + 1), (n - 130, n - 127), (n - 2, n)]
    _check_ranges(out, q, k, v, n, 0, 0, ranges, "dense 512K")
    # constant V: every output row is that constant (softmax weights sum to 1), over all 512K rows
    outc = ops.prefill_attn(q, k, torch.full_like(v, 0.5), cu, n, SCALE, SI, SL, LAST, 0)
    assert float((outc.float() - 0.5).abs().max()) <= 2 ** -8


def test_mask_bits_256k(ops):
    """256K bitmask = 8 GiB = 2^31 words: population counts equal the closed-form pair counts."""
    n = 1 << 18
    lut = torch.tensor([bin(i).count("1") for i in range(256)], dtype=torch.int32, device=DEV)
    for piece, want in (("triangle", orc.attended_pairs(n, SI, SL, LAST, orc.TRIANGLE)),
                        ("dense", n * (n + 1) // 2)):
        bits = ops.mask_bits(n, SI, SL, LAST, piece, DEV)
        b8 = bits.view(torch.uint8)
        total = 0
        for r0 in range(0, n, 4096):
            total += int(lut[b8[r0:r0 + 4096].long()].sum(dtype=torch.int64))
        assert total == want, (piece, total, want)
        # the last row (index 2^18 - 1, word offset beyond 2^31 - 2^13) attends every column in both pieces
        assert int(lut[b8[n - 1].long()].sum()) == n
        del bits, b8
