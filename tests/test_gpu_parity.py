"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle and the
golden vectors produced by the unmodified reference.  Masks: bit-exact.  Attention:
bf16 kernel vs the fp32 oracle on the same bf16-rounded inputs, atol 1.5e-2 / rtol 1e-2
(the reference repo's own bar for bf16 attention kernels, tests/kernels/attention/
test_flash_attn.py:179,192)."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch

from oracle import trianglemix_oracle as orc

pytestmark = pytest.mark.gpu
ATOL, RTOL = 1.5e-2, 1e-2
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
with open(os.path.join(GOLDEN, "masks.json")) as f:
    G = json.load(f)
ATTN = np.load(os.path.join(GOLDEN, "attn_small.npz"))
KERNELS = ["mma", "tcgen05", "pair", "auto"]


def _skip_unpaired(kernel, hq, hkv):
    # the tcgen05 kernel pairs two query heads of one KV head per CTA
    if kernel == "tcgen05" and (hq // hkv) % 2:
        pytest.skip("tcgen05 kernel needs an even number of query heads per KV head")
    if kernel == "pair" and (hq // hkv) % 4:
        pytest.skip("the CTA-pair kernel takes four query heads of a KV head per cluster")


@pytest.fixture(scope="module")
def tmx():
    import vllm_trianglemix_b200 as m
    return m


@pytest.fixture(scope="module")
def ops():
    from vllm_trianglemix_b200 import ops as o
    return o


DEV = "cuda:0"


def _u32(t):
    return t.cpu().numpy().view(np.uint32)


# ------------------------------------------------------------------ masks
@pytest.mark.parametrize("case", G["cases"], ids=lambda c: "N%d_%d_%d_%d" % (c["N"], c["si"], c["sl"], c["num_last"]))
def test_mask_bits_golden(ops, case):
    n, si, sl, last = case["N"], case["si"], case["sl"], case["num_last"]
    words = (n + 31) // 32
    for piece in orc.PIECES:
        bits = _u32(ops.mask_bits(n, si, sl, last, piece, DEV))
        assert not bits[:, words:].any()
        got = np.ascontiguousarray(bits[:, :words])
        assert hashlib.sha256(got.tobytes()).hexdigest() == case[piece + "_sha"], piece
        assert int(np.unpackbits(got.view(np.uint8)).sum()) == case[piece]


@pytest.mark.parametrize("case", G["big"], ids=lambda c: "N%d" % c["N"])
def test_mask_bits_golden_big(ops, case):
    n, si, sl, last = case["N"], case["si"], case["sl"], case["num_last"]
    bits = _u32(ops.mask_bits(n, si, sl, last, "triangle", DEV))[:, :(n + 31) // 32]
    assert hashlib.sha256(np.ascontiguousarray(bits).tobytes()).hexdigest() == case["triangle_sha"]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("n,si,sl,last", [(300, 4, 32, 64), (97, 4, 32, 64), (130, 0, 0, 1), (1, 4, 32, 64),
                                          (2, 0, 0, 0), (64, 7, 3, 200), (64, 100, 3, 5), (1000, 4, 32, 64)])
def test_generator_api_matches_oracle(tmx, dtype, n, si, sl, last):
    gen = tmx.TriangleMixMaskGenerator(tmx.TriangleMixConfig(si, sl, last))
    dev = torch.device(DEV)
    for piece in orc.PIECES:
        m = getattr(gen, "_create_%s_mask" % piece)(n, dev, dtype)
        assert m.dtype == dtype and tuple(m.shape) == (n, n)
        assert np.array_equal(m.float().cpu().numpy(), orc.piece_closed(piece, n, si, sl, last).astype(np.float32))
    for tri, mode in ((True, orc.TRIANGLE), (False, orc.DENSE_CAUSAL)):
        a = gen.get_attention_mask(n, 3, dev, dtype, use_triangle=tri)
        assert a.dtype == dtype
        assert np.array_equal(a.float().cpu().numpy(), orc.additive(orc.mask_closed(n, si, sl, last, mode)))
        assert gen.get_attention_mask(n, 3, dev, dtype, use_triangle=tri) is a        # cached object
        assert (n, 3, str(dev), str(dtype), tri) in gen._mask_cache
    gen.clear_cache()
    assert not gen._mask_cache


@pytest.mark.parametrize("n", [32768, 131072])
def test_mask_bits_long_by_row_chunks(ops, n):
    si, sl, last = 4, 32, 64
    bits = ops.mask_bits(n, si, sl, last, "triangle", DEV)
    stride = bits.shape[1]
    total = int(torch.ops.aten.sum(_popcount(bits)).item())
    assert total == orc.attended_pairs(n, si, sl, last, orc.TRIANGLE)
    for r0 in (0, 64, n // 2 - 7, n - last - 70, n - 256):
        r1 = min(n, r0 + 256)
        want = orc.pack_bits(orc.mask_closed(n, si, sl, last, orc.TRIANGLE, r0, r1), stride)
        assert np.array_equal(_u32(bits[r0:r1]), want)


def _popcount(x):
    x = x.to(torch.int64) & 0xFFFFFFFF
    x = x - ((x >> 1) & 0x55555555)
    x = (x & 0x33333333) + ((x >> 2) & 0x33333333)
    x = (x + (x >> 4)) & 0x0F0F0F0F
    return ((x * 0x01010101) & 0xFFFFFFFF) >> 24


def test_bits_from_additive_and_rewrite(tmx, ops):
    n, si, sl, last = 777, 3, 17, 40
    dev = torch.device(DEV)
    for dtype in (torch.float32, torch.bfloat16, torch.float16):
        add = ops.mask_additive(n, si, sl, last, ops.TRIANGLE, dev, dtype)
        bits, bad = ops.bits_from_additive(add)
        assert int(bad.item()) == 0
        assert torch.equal(bits, ops.mask_bits(n, si, sl, last, "triangle", dev))
        rew = tmx.AscendNPUTriangleMixOptimizer.optimize_mask_for_npu(add)
        if dtype != torch.float16:      # -1e6 overflows to -inf in fp16, in the reference too
            assert not torch.isinf(rew).any()
        want = orc.rewrite_neg_inf(orc.additive(orc.mask_closed(n, si, sl, last, orc.TRIANGLE)))
        assert np.array_equal(rew.float().cpu().numpy(), want.astype(np.float32) if dtype == torch.float32
                              else torch.from_numpy(want).to(dtype).float().numpy())
    add = ops.mask_additive(64, si, sl, last, ops.TRIANGLE, dev, torch.float32)
    add[5, 3] = -2.5
    _, bad = ops.bits_from_additive(add)
    assert int(bad.item()) == 1


@pytest.mark.parametrize("bm,bn", [(64, 64), (128, 64), (128, 128), (32, 32)])
def test_block_index_vs_bruteforce(ops, bm, bn):
    lens = [1, 63, 64, 65, 300, 1000, 2049]
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    total = int(cu[-1])
    for (si, sl, last) in ((4, 32, 64), (0, 0, 1), (7, 3, 200), (100, 3, 5), (1, 70, 130)):
        for mode in (orc.TRIANGLE, orc.DENSE_CAUSAL):
            ranges, counts, full = ops.block_index(torch.from_numpy(cu).to(DEV), total, bm, bn, si, sl, last, mode)
            ranges, counts, full = ranges.cpu().numpy(), counts.cpu().numpy(), full.cpu().numpy()
            used = set()
            for b, n in enumerate(lens):
                m = orc.mask_closed(n, si, sl, last, mode)
                for q in range((n + bm - 1) // bm):
                    slot = cu[b] // bm + b + q
                    used.add(slot)
                    rows = m[q * bm:(q + 1) * bm]
                    touched = [j for j in range((n + bn - 1) // bn) if rows[:, j * bn:(j + 1) * bn].any()]
                    lo0, hi0, lo1, hi1 = ranges[slot]
                    listed = list(range(lo0, hi0)) + list(range(lo1, hi1))
                    assert set(touched) <= set(listed)
                    if not (lo1 == hi1 and mode == orc.TRIANGLE):
                        assert listed == touched            # exact when two ranges / dense
                    assert counts[slot] == len(listed)
                    nfull = sum(1 for j in listed if rows[:, j * bn:(j + 1) * bn].all()
                                and (j + 1) * bn <= n and rows.shape[0] > 0)
                    assert full[slot].sum() <= nfull       # FULL is sufficient, never over-claimed
            for slot in range(total // bm + len(lens)):
                if slot not in used:
                    assert counts[slot] == 0


# ------------------------------------------------------------------ attention
def _bf16(a):
    return torch.from_numpy(a.copy()).view(torch.bfloat16)


def _check(out, ref, what):
    rep = orc.error_report(out, ref, ATOL, RTOL)
    assert rep["violations"] == 0, (what, rep)
    return rep


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_attention_golden(ops, tag, kernel):
    s, hq, hkv, d, si, sl, last = [int(x) for x in ATTN[tag + "_params"]]
    _skip_unpaired(kernel, hq, hkv)
    q, k, v = _bf16(ATTN[tag + "_q"])[0], _bf16(ATTN[tag + "_k"])[0], _bf16(ATTN[tag + "_v"])[0]
    cu = torch.tensor([0, s], dtype=torch.int32, device=DEV)
    for name, mode in (("tri", ops.TRIANGLE), ("dense", ops.DENSE_CAUSAL)):
        out = ops.prefill_attn(q.to(DEV), k.to(DEV), v.to(DEV), cu, s, d ** -0.5, si, sl, last, mode, kernel=kernel)
        _check(out.cpu(), torch.from_numpy(ATTN["%s_%s_out" % (tag, name)])[0], (tag, name, kernel))
    if tag == "a":
        out = ops.prefill_attn(q.to(DEV), k.to(DEV), v.to(DEV), cu, s, d ** -0.5, 0, 0, 0, ops.NONE, kernel=kernel)
        _check(out.cpu(), torch.from_numpy(ATTN["a_nomask_out"])[0], (tag, "nomask", kernel))


def _rand(t, hq, hkv, d=128, seed=0, qscale=1.0):
    g = torch.Generator().manual_seed(seed)
    q = (torch.randn(t, hq, d, generator=g) * qscale).bfloat16()
    k = torch.randn(t, hkv, d, generator=g).bfloat16()
    v = torch.randn(t, hkv, d, generator=g).bfloat16()
    return q, k, v


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("n,hq,hkv,si,sl,last", [
    (2048, 16, 8, 4, 32, 64),          # BASELINE config 1
    (1, 4, 4, 4, 32, 64), (2, 2, 1, 0, 0, 0), (63, 2, 2, 4, 32, 64), (65, 4, 1, 4, 32, 64),
    (129, 8, 2, 0, 0, 1), (1000, 4, 2, 7, 3, 200), (513, 2, 1, 100, 3, 5), (3000, 4, 2, 1, 70, 130),
    (4500, 2, 1, 4, 32, 64),           # long enough for split-KV of the dense tail rows
    # streaming kernel (kernel="auto", triangle): GQA ratios 2 / 4 / 8, widest window / sink it takes,
    # and the settings just outside it (si = 16, window too wide, > 512 dense tail rows) that fall back
    (3000, 8, 2, 15, 80, 64), (3000, 8, 1, 3, 5, 33), (5000, 4, 2, 4, 32, 64), (4200, 16, 2, 0, 0, 2),
    (700, 8, 2, 16, 32, 64), (1300, 8, 2, 4, 81, 64), (9000, 8, 2, 4, 32, 600), (37, 8, 2, 4, 32, 64),
])
def test_attention_vs_oracle(ops, n, hq, hkv, si, sl, last, mode, kernel):
    _skip_unpaired(kernel, hq, hkv)
    q, k, v = _rand(n, hq, hkv, seed=n)
    cu = torch.tensor([0, n], dtype=torch.int32, device=DEV)
    scale = 128 ** -0.5
    out, lse = ops.prefill_attn(q.to(DEV), k.to(DEV), v.to(DEV), cu, n, scale, si, sl, last, mode,
                                return_lse=True, kernel=kernel)
    ref = orc.attention_varlen_fp32(q, k, v, [0, n], scale, si, sl, last, mode)
    _check(out.cpu(), ref, (n, mode, kernel))
    # lse against the oracle's scores on a few rows
    rows = sorted(set([0, n // 2, n - 1]))
    m = torch.from_numpy(orc.additive(orc.mask_closed(n, si, sl, last, mode)))[rows]
    kk = k.float().repeat_interleave(hq // hkv, dim=1)
    sc = torch.einsum("rhd,jhd->hrj", q[rows].float(), kk) * scale + m
    assert torch.allclose(lse.cpu()[:, rows], torch.logsumexp(sc, dim=-1), atol=2e-2, rtol=1e-3)


@pytest.mark.parametrize("kernel", KERNELS)
def test_attention_bitmask_source_identical(ops, kernel):
    n, hq, hkv, si, sl, last = 1500, 4, (1 if kernel == "pair" else 2), 4, 32, 64
    q, k, v = _rand(n, hq, hkv, seed=3)
    cu = torch.tensor([0, n], dtype=torch.int32, device=DEV)
    for mode, piece in ((ops.TRIANGLE, "triangle"), (ops.DENSE_CAUSAL, "dense")):
        a = ops.prefill_attn(q.to(DEV), k.to(DEV), v.to(DEV), cu, n, 0.088, si, sl, last, mode, kernel=kernel)
        bits = ops.mask_bits(n, si, sl, last, piece, DEV)
        b = ops.prefill_attn(q.to(DEV), k.to(DEV), v.to(DEV), cu, n, 0.088, si, sl, last, mode, bitmask=bits,
                             kernel=kernel)
        if kernel == "auto" and mode == ops.TRIANGLE:
            # without a bitmask the streaming kernel serves the layer, with one the mma kernel does
            _check(b.cpu().float(), a.cpu().float(), "bitmask vs streaming kernel")
        else:
            assert torch.equal(a, b)
        c = ops.prefill_attn(q.to(DEV), k.to(DEV), v.to(DEV), cu, n, 0.088, 0, 0, 0, ops.NONE, bitmask=bits,
                             kernel=kernel)
        _check(c.cpu().float(), a.cpu().float(), "all-tiles bitmask")


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("mode", [0, 1])
def test_attention_varlen(ops, mode, kernel):
    lens = [700, 1, 64, 4500, 129, 2048]
    t = sum(lens)
    hq, hkv, si, sl, last = 8, 2, 4, 32, 64
    _skip_unpaired(kernel, hq, hkv)
    q, k, v = _rand(t, hq, hkv, seed=11)
    cu_list = [0]
    for n in lens:
        cu_list.append(cu_list[-1] + n)
    cu = torch.tensor(cu_list, dtype=torch.int32, device=DEV)
    # strided inputs: q/k/v are slices of a fused qkv buffer, as vLLM's qkv_proj output is
    fused = torch.cat([q.reshape(t, -1), k.reshape(t, -1), v.reshape(t, -1)], dim=1).to(DEV)
    qd = fused[:, :hq * 128].view(t, hq, 128)
    kd = fused[:, hq * 128:(hq + hkv) * 128].view(t, hkv, 128)
    vd = fused[:, (hq + hkv) * 128:].view(t, hkv, 128)
    out = ops.prefill_attn(qd, kd, vd, cu, max(lens), 128 ** -0.5, si, sl, last, mode, kernel=kernel)
    ref = orc.attention_varlen_fp32(q, k, v, cu_list, 128 ** -0.5, si, sl, last, mode)
    _check(out.cpu(), ref, ("varlen", mode, kernel))


@pytest.mark.parametrize("kernel", KERNELS)
def test_attention_peaky_softmax(ops, kernel):
    n, hq, hkv = 1024, 4, 1
    _skip_unpaired(kernel, hq, hkv)
    q, k, v = _rand(n, hq, hkv, seed=5, qscale=4.0)
    cu = torch.tensor([0, n], dtype=torch.int32, device=DEV)
    for mode in (0, 1):
        out = ops.prefill_attn(q.to(DEV), k.to(DEV), v.to(DEV), cu, n, 128 ** -0.5, 4, 32, 64, mode, kernel=kernel)
        ref = orc.attention_varlen_fp32(q, k, v, [0, n], 128 ** -0.5, 4, 32, 64, mode)
        _check(out.cpu(), ref, ("peaky", mode, kernel))


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("step", [3.0, 12.0, -6.0])
def test_attention_drifting_scores(ops, kernel, step):
    """Scores that drift by `step` (log2 units) per 128-key tile along the key axis: with a rising drift every tile
    makes the running reference of the max-free softmax pass stale (the late-refresh path that re-bases the scores
    it holds in registers); a falling drift leaves the first tile's maximum standing for the whole row."""
    n, hq, hkv = 2304, 4, 1
    _skip_unpaired(kernel, hq, hkv)
    g = torch.Generator().manual_seed(11)
    u = torch.randn(128, generator=g)
    u = u / u.norm()
    ramp = torch.arange(n, dtype=torch.float32) / 128.0 * step / 1.4427          # natural-log units per key
    k = (torch.randn(n, hkv, 128, generator=g) * 0.5 + (ramp[:, None, None] * u) * (128 ** 0.5) / 4.0).bfloat16()
    q = (torch.randn(n, hq, 128, generator=g) * 0.5 + 4.0 * u).bfloat16()
    v = torch.randn(n, hkv, 128, generator=g).bfloat16()
    cu = torch.tensor([0, n], dtype=torch.int32, device=DEV)
    for mode, prm in ((0, (0, 0, 0)), (1, (20, 300, 40)), (2, (0, 0, 0))):
        out = ops.prefill_attn(q.to(DEV), k.to(DEV), v.to(DEV), cu, n, 128 ** -0.5, *prm, mode, kernel=kernel)
        if mode == 2:       # no mask: plain fp32 softmax(Q K^T * scale) V (ascend_trianglemix.py:86-106 without a mask)
            kf = k.float().repeat_interleave(hq // hkv, dim=1).transpose(0, 1)
            vf = v.float().repeat_interleave(hq // hkv, dim=1).transpose(0, 1)
            w = torch.softmax(torch.matmul(q.float().transpose(0, 1), kf.transpose(-2, -1)) * 128 ** -0.5, dim=-1)
            ref = torch.matmul(w, vf).transpose(0, 1).contiguous()
        else:
            ref = orc.attention_varlen_fp32(q, k, v, [0, n], 128 ** -0.5, *prm, mode)
        _check(out.cpu(), ref, ("drift", step, mode, kernel))


# ------------------------------------------------------------------ seeded fuzz over the whole parameter space
def _fuzz_case(seed):
    import random
    r = random.Random(seed)
    b = r.choice([1, 1, 2, 3, 5])
    lens = []
    for _ in range(b):
        kind = r.random()
        if kind < 0.15:
            lens.append(r.choice([1, 2, 31, 32, 33, 127, 128, 129, 255, 256, 257, 512]))
        elif kind < 0.3:
            lens.append(128 * r.randint(1, 10))
        else:
            lens.append(r.randint(1, 1400))
    hq, hkv = r.choice([(2, 1), (2, 2), (4, 2), (4, 1), (8, 2), (8, 1), (16, 8), (32, 8), (6, 2), (3, 1)])
    nmax = max(lens)
    si = r.choice([0, 1, 4, 4, 4, 15, 16, 20, nmax + 5])
    sl = r.choice([0, 1, 32, 32, 32, 63, 64, 80, 81, 200])
    last = r.choice([0, 1, 2, 64, 64, 64, 65, 129, 300, 513, nmax + 3])
    mode = r.choice([0, 1, 1])
    kernel = r.choice(["auto", "auto", "mma"])
    if (hq // hkv) % 4 == 0 and r.random() < 0.5:
        kernel = "pair"
    scale = r.choice([128 ** -0.5, 0.05, 0.2])
    return lens, hq, hkv, si, sl, last, mode, kernel, scale


@pytest.mark.parametrize("seed", range(int(os.environ.get("TMX_FUZZ_SEEDS", "60"))))
def test_attention_fuzz_vs_oracle(ops, seed):
    lens, hq, hkv, si, sl, last, mode, kernel, scale = _fuzz_case(seed)
    t = sum(lens)
    q, k, v = _rand(t, hq, hkv, seed=1000 + seed)
    cu_list = [0]
    for n in lens:
        cu_list.append(cu_list[-1] + n)
    cu = torch.tensor(cu_list, dtype=torch.int32, device=DEV)
    out = ops.prefill_attn(q.to(DEV), k.to(DEV), v.to(DEV), cu, max(lens), scale, si, sl, last, mode, kernel=kernel)
    ref = orc.attention_varlen_fp32(q, k, v, cu_list, scale, si, sl, last, mode)
    _check(out.cpu(), ref, (seed, lens, hq, hkv, si, sl, last, mode, kernel))


# ------------------------------------------------------------------ torch.library boundary
def test_registered_torch_op_matches_ctypes_path(ops):
    """torch.ops.trianglemix.prefill_attn (TORCH_LIBRARY TU) and the ctypes binding call the same entry point."""
    from vllm_trianglemix_b200 import torch_ops
    n, hq, hkv = 3000, 8, 2
    q, k, v = _rand(n, hq, hkv, seed=21)
    qd, kd, vd = q.to(DEV), k.to(DEV), v.to(DEV)
    cu = torch.tensor([0, n], dtype=torch.int32, device=DEV)
    for mode in (ops.TRIANGLE, ops.DENSE_CAUSAL):
        a = ops.prefill_attn(qd, kd, vd, cu, n, 0.088, 4, 32, 64, mode)
        b = torch_ops.prefill_attn(qd, kd, vd, cu, n, 0.088, 4, 32, 64, mode)
        assert torch.equal(a, b)
    t = torch_ops.load()
    bits = torch.empty((n, ops.bitmask_stride_words(n)), dtype=torch.int32, device=DEV)
    t.mask_bits(bits, n, 4, 32, 64, ops.PIECES["triangle"])
    assert torch.equal(bits, ops.mask_bits(n, 4, 32, 64, "triangle", DEV))
    # the op is traceable: torch.compile sees a registered, mutating op instead of an opaque ctypes call
    ws = torch.empty(t.attn_workspace_bytes(1, n, n, hq, hkv, 128, 4, 32, 64, 1, 0), dtype=torch.uint8, device=DEV)

    def f(qq, kk, vv):
        o = torch.empty_like(qq)
        torch.ops.trianglemix.prefill_attn(o, qq, kk, vv, cu, n, 0.088, 4, 32, 64, 1, None, None, ws, 0)
        return o * 1.0

    c = torch.compile(f, fullgraph=True, backend="eager")(qd, kd, vd)
    assert torch.equal(c, ops.prefill_attn(qd, kd, vd, cu, n, 0.088, 4, 32, 64, ops.TRIANGLE))


# ------------------------------------------------------------------ o_proj GEMM (the step after attention)
@pytest.mark.parametrize("m,n,k", [(128, 256, 64), (1000, 512, 256), (4096, 4096, 512), (333, 1024, 4096), (1, 256, 128),
                                   (8192, 5120, 1024)])
def test_linear_vs_fp32_matmul(ops, m, n, k):
    """C = A W^T on tcgen05 against the fp32 matmul of the same bf16 inputs (what RowParallelLinear computes,
    linear.py:1197-1359).  bf16 output of an fp32 accumulation: |err| <= 2^-8 |ref| + accumulation noise."""
    g = torch.Generator().manual_seed(m + n + k)
    a = torch.randn(m, k, generator=g).bfloat16()
    w = (torch.randn(n, k, generator=g) * k ** -0.5).bfloat16()
    out = ops.linear(a.to(DEV), w.to(DEV)).cpu().float()
    ref = a.float() @ w.float().t()
    assert out.shape == (m, n)
    assert torch.allclose(out, ref, atol=1e-2, rtol=1e-2), float((out - ref).abs().max())
    assert float((out - ref).norm() / ref.norm()) < 4e-3
    # strided input: a slice of a wider buffer, as attn_output is a view of the attention layer's output
    wide = torch.randn(m, k + 64, generator=g).bfloat16().to(DEV)
    out2 = ops.linear(wide[:, :k], w.to(DEV)).cpu().float()
    assert torch.allclose(out2, wide[:, :k].cpu().float() @ w.float().t(), atol=1e-2, rtol=1e-2)


def test_linear_scatter_and_sum_single_process(ops):
    """The fused reduce-scatter epilogue with local buffers standing in for the peers: 4 'ranks' x 256 rows."""
    world, rows, n, k = 4, 256, 512, 128
    g = torch.Generator().manual_seed(5)
    ws = [(torch.randn(n, k, generator=g) * k ** -0.5).bfloat16().to(DEV) for _ in range(world)]
    xs = [torch.randn(world * rows, k, generator=g).bfloat16().to(DEV) for _ in range(world)]
    staging = [torch.zeros(world, rows, n, dtype=torch.bfloat16, device=DEV) for _ in range(world)]
    for r in range(world):          # "rank" r pushes its partial of every row block into the owner's slot r
        ops.linear_scatter(xs[r], ws[r], [s.data_ptr() for s in staging], r, world, rows)
    total = sum(x.float() @ w.float().t() for x, w in zip(xs, ws))
    for o in range(world):
        got = ops.sum_partials(staging[o]).cpu().float()
        ref = total[o * rows:(o + 1) * rows].cpu()
        assert torch.allclose(got, ref, atol=3e-2, rtol=2e-2), float((got - ref).abs().max())
