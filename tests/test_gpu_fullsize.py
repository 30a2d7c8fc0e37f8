"""GPU parity at BASELINE.json's FULL sizes (configs 2-5), where the oracle cannot
materialise S x S scores: (i) sampled query rows against the oracle's row form
(attention_rows_fp32: the reference arithmetic for chosen rows against ALL keys),
(ii) size-independent properties that need no oracle at all -- V-linearity with a power
of two (bit-exact), rows of a constant V, head-permutation equivariance, the two
independent kernels (tcgen05 / mma.sync) agreeing, and dense == triangle on the rows where
the masks coincide.  Tolerance as in test_gpu_parity.py: atol 1.5e-2 / rtol 1e-2."""
import numpy as np
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


def _sample_rows(n, last):
    lrow = max(0, n - last + 1)
    rows = {0, 1, SI, SI + 1, SL, SL + 1, 63, 64, 127, 128, 129, n // 2 - 1, n // 2, n // 3,
            lrow - 129, lrow - 1, lrow, lrow + 1, n - 2, n - 1}
    return sorted(r for r in rows if 0 <= r < n)


def _check_rows(out_dev, q, k, v, n, s0, mode, kv_heads_checked, what):
    """out rows of ONE sequence [s0, s0+n) against the oracle, for a subset of KV heads."""
    g = q.shape[1] // k.shape[1]
    rows = _sample_rows(n, LAST)
    worst = 0.0
    for h in kv_heads_checked:
        qs = q[s0:s0 + n, h * g:(h + 1) * g].cpu()
        ks = k[s0:s0 + n, h:h + 1].cpu()
        vs = v[s0:s0 + n, h:h + 1].cpu()
        for r in rows:
            ref = orc.attention_rows_fp32(qs, ks, vs, SCALE, n, SI, SL, LAST, mode, r, r + 1)
            got = out_dev[s0 + r:s0 + r + 1, h * g:(h + 1) * g].cpu()
            rep = orc.error_report(got, ref, ATOL, RTOL)
            assert rep["violations"] == 0, (what, "kv head", h, "row", r, rep)
            worst = max(worst, rep["max_abs"])
    return worst


@pytest.mark.parametrize("n,hq,hkv", [(32768, 32, 8),      # BASELINE configs[1]
                                      (131072, 32, 8),     # configs[2]
                                      (131072, 64, 8)])    # configs[3] (Qwen3-32B heads, unsharded)
@pytest.mark.parametrize("mode", [1, 0])
def test_fullsize_sampled_rows_vs_oracle(ops, n, hq, hkv, mode):
    if mode == 0 and hq == 64:
        pytest.skip("dense 128K x 64 heads: 0.25 s of GPU per call, covered by the 32-head case")
    q, k, v = _rand_dev(n, hq, hkv, seed=n + hq)
    cu = torch.tensor([0, n], dtype=torch.int32, device=DEV)
    out = ops.prefill_attn(q, k, v, cu, n, SCALE, SI, SL, LAST, mode)
    torch.cuda.synchronize()
    _check_rows(out, q, k, v, n, 0, mode, (0, hkv - 1), ("full", n, hq, mode))


def test_fullsize_varlen_config5(ops):
    """BASELINE configs[4]: 8 sequences of 8K-64K tokens, Qwen3-8B heads (SURVEY 8(d) length list)."""
    lens = [8192, 12288, 16384, 24576, 32768, 40960, 49152, 65536]
    t = sum(lens)
    q, k, v = _rand_dev(t, 32, 8, seed=5)
    cu_list = np.concatenate([[0], np.cumsum(lens)]).tolist()
    cu = torch.tensor(cu_list, dtype=torch.int32, device=DEV)
    for mode in (1, 0):
        out = ops.prefill_attn(q, k, v, cu, max(lens), SCALE, SI, SL, LAST, mode)
        torch.cuda.synchronize()
        for b in (0, 3, 7):
            _check_rows(out, q, k, v, lens[b], cu_list[b], mode, (b % 8,), ("varlen", b, mode))
        if mode == 1:
            # a sequence's result does not depend on its neighbours in the batch
            b = 2
            alone = ops.prefill_attn(q[cu_list[b]:cu_list[b + 1]], k[cu_list[b]:cu_list[b + 1]],
                                     v[cu_list[b]:cu_list[b + 1]],
                                     torch.tensor([0, lens[b]], dtype=torch.int32, device=DEV), lens[b],
                                     SCALE, SI, SL, LAST, mode)
            both = out[cu_list[b]:cu_list[b + 1]]
            tail0 = (lens[b] - LAST + 1) // 128 * 128      # tail tiles: KV split count depends on the batch
            assert torch.equal(alone[:tail0], both[:tail0])
            rep = orc.error_report(alone[tail0:].cpu(), both[tail0:].cpu().float(), ATOL, RTOL)
            assert rep["violations"] == 0, rep


@pytest.mark.parametrize("mode", [1, 0])
def test_fullsize_properties_32k(ops, mode):
    n, hq, hkv = 32768, 32, 8
    q, k, v = _rand_dev(n, hq, hkv, seed=17)
    cu = torch.tensor([0, n], dtype=torch.int32, device=DEV)
    out = ops.prefill_attn(q, k, v, cu, n, SCALE, SI, SL, LAST, mode)
    # (1) linear in V; scaling by a power of two is exact in bf16 and fp32 => bit-exact
    out4 = ops.prefill_attn(q, k, v * 4, cu, n, SCALE, SI, SL, LAST, mode)
    # (outputs that are bf16 subnormals, ~1e-40, may flush differently: allow 1e-30)
    assert float((out4.float() - out.float() * 4).abs().max()) < 1e-30
    # (2) constant V: every output row is that constant (softmax weights sum to 1)
    ones = torch.full_like(v, 0.5)
    outc = ops.prefill_attn(q, k, ones, cu, n, SCALE, SI, SL, LAST, mode)
    assert float((outc.float() - 0.5).abs().max()) <= 2 ** -8
    # (3) permuting KV-head groups permutes the output (heads never interact)
    perm = torch.randperm(hkv, device=DEV)
    g = hq // hkv
    qperm = (perm[:, None] * g + torch.arange(g, device=DEV)[None, :]).reshape(-1)
    outp = ops.prefill_attn(q[:, qperm].contiguous(), k[:, perm].contiguous(), v[:, perm].contiguous(), cu, n,
                            SCALE, SI, SL, LAST, mode)
    assert torch.equal(outp, out[:, qperm])
    # (4) the mma.sync kernel (independent code path) agrees within the oracle tolerance
    if mode == 1:
        outm = ops.prefill_attn(q, k, v, cu, n, SCALE, SI, SL, LAST, mode, kernel="mma")
        rep = orc.error_report(outm.cpu(), out.cpu().float(), ATOL, RTOL)
        assert rep["violations"] == 0, rep
    # (5) masks coincide on rows i <= sl and on the dense tail rows => same attended sets there
    other = ops.prefill_attn(q, k, v, cu, n, SCALE, SI, SL, LAST, 1 - mode)
    lrow = n - LAST + 1
    for rows in (slice(0, SL + 1), slice(lrow, n)):
        rep = orc.error_report(other[rows].cpu(), out[rows].cpu().float(), ATOL, RTOL)
        assert rep["violations"] == 0, (rows, rep)


def test_fullsize_mask_bits_properties(ops):
    """128K bitmask: population count per piece equals the closed-form pair counts
    (13,233,222 attended pairs in the triangle mask), and the pieces partition
    the causal triangle: streaming | last | middle == dense, pairwise disjoint."""
    n = 131072
    counts = {}
    acc = None
    for piece in ("streaming", "last", "middle", "triangle", "dense"):
        bits = ops.mask_bits(n, SI, SL, LAST, piece, DEV)
        b8 = bits.view(torch.uint8)
        lut = torch.tensor([bin(i).count("1") for i in range(256)], dtype=torch.int64, device=DEV)
        total = 0
        for r0 in range(0, n, 8192):
            total += int(lut[b8[r0:r0 + 8192].long()].sum())
        counts[piece] = total
        if piece in ("streaming", "last", "middle"):
            if acc is None:
                acc = bits.clone()
            else:
                assert int((acc & bits).ne(0).sum()) == 0          # disjoint
                acc |= bits
        elif piece == "dense":
            assert torch.equal(acc, bits)                          # partition of the causal triangle
        del bits, b8
    assert counts["triangle"] == 13233222 == orc.attended_pairs(n, SI, SL, LAST, orc.TRIANGLE)
    assert counts["dense"] == n * (n + 1) // 2
    assert counts["streaming"] + counts["last"] == counts["triangle"]
    assert counts["streaming"] + counts["last"] + counts["middle"] == counts["dense"]
