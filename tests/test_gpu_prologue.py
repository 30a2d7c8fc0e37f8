"""GPU parity of the step in front of attention: fused q/k RMSNorm +
NeoX RoPE (+ KV-cache write) through the C ABI against the oracle and the reference-produced
golden vectors.

Tolerance (stated, floating point): every operation of the reference's forward_native is
reproduced with its bf16 rounding, so the only freedom is the fp32 summation order of the 128
squares in the variance.  That can move a head row's rsqrt by 1 fp32 ulp, which flips the bf16
rounding of x * rsqrt only for elements sitting within 1 fp32 ulp of a rounding midpoint (about
1 in 2^15; equal |x| values flip together) -- measured: a flipped row is seen in ~0.1 % of head
rows, e.g. fuzz seed 1385, reproduced on the CPU by adding 1 ulp to the sum of squares.  A flip is
1 bf16 ulp of a rotation operand, which x1*cos - x2*sin carries into an output that may be much
smaller than the operand (fuzz seed 1858: |diff| 2^-5 next to operands of 4..8).  Bar: every
element within atol 1.5e-2 + rtol 1e-2 x |rotary pair (e, e+64)| of the oracle, and at most
max(2, 1 %) of the head rows (128 elements) contain any element that is not bit-identical.
Cache rows and untouched slots: bit-exact."""
import os

import numpy as np
import pytest
import torch

from oracle import qknorm_rope_oracle as orc
from oracle import trianglemix_oracle as aorc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "qknorm_rope.npz"))


@pytest.fixture(scope="module")
def ops():
    from vllm_trianglemix_b200 import ops as o
    return o


@pytest.fixture(scope="module")
def tmx():
    import vllm_trianglemix_b200 as m
    return m


def bf(a):
    return torch.from_numpy(a.copy()).view(torch.bfloat16)


def close_enough(got, want, what):
    got, want = got.cpu(), want.cpu()
    ne = (got.reshape(-1, 128).view(torch.int16) != want.reshape(-1, 128).view(torch.int16))
    rows, rows_bad = ne.shape[0], int(ne.any(dim=1).sum())
    assert rows_bad <= max(2, rows // 100), "%s: %d of %d head rows differ from the oracle" % (what, rows_bad, rows)
    # a flip is 1 bf16 ulp of a ROTATION OPERAND, so the relative part is taken against the norm of the
    # rotary pair (e, e + 64) the element belongs to, not against the element (x1*cos - x2*sin may cancel)
    g, w = got.float().reshape(-1, 2, 64), want.float().reshape(-1, 2, 64)
    pair = (w * w).sum(dim=1, keepdim=True).sqrt()
    bad = ((g - w).abs() > 1.5e-2 + 1e-2 * pair).sum().item()
    assert bad == 0, "%s: %d elements outside tolerance (max abs %.3e)" % (what, bad, (g - w).abs().max().item())
    return 1.0 - float(ne.float().mean())


@pytest.mark.parametrize("tag", ["a", "b"])
def test_golden_from_reference(ops, tag):
    t, hq, hkv, d, max_pos = (int(x) for x in G[tag + "_params"])
    cache = orc.cos_sin_cache(d, max_pos, 1000000.0)
    pos = torch.from_numpy(G[tag + "_pos"])
    q2, k2 = ops.qk_norm_rope(pos.to(DEV), bf(G[tag + "_q"]).to(DEV), bf(G[tag + "_k"]).to(DEV),
                              bf(G[tag + "_qw"]).to(DEV), bf(G[tag + "_kw"]).to(DEV), cache.to(DEV), 1e-6, d)
    close_enough(q2, bf(G[tag + "_qo"]), "q")
    close_enough(k2, bf(G[tag + "_ko"]), "k")
    assert ops.last_launch_count() == 1


def _inputs(t, hq, hkv, max_pos, seed, qscale=3.0):
    g = torch.Generator().manual_seed(seed)
    d = 128
    qkv = torch.randn(t, (hq + 2 * hkv) * d, generator=g)
    qkv[:, :hq * d] *= qscale
    qkv = qkv.bfloat16()
    qw = (1.0 + 0.1 * torch.randn(d, generator=g)).bfloat16()
    kw = (1.0 + 0.1 * torch.randn(d, generator=g)).bfloat16()
    pos = torch.randint(0, max_pos, (t,), generator=g, dtype=torch.int64)
    return qkv, qw, kw, pos


@pytest.mark.parametrize("t,hq,hkv", [(1, 2, 1), (37, 4, 2), (1000, 16, 8), (4099, 32, 8), (513, 8, 1)])
@pytest.mark.parametrize("inplace", [False, True])
def test_vs_oracle_packed_qkv_views(ops, t, hq, hkv, inplace):
    """q/k are strided views of the packed qkv buffer, as qkv.split gives them (qwen3.py:161)."""
    d, max_pos = 128, 8192
    qkv, qw, kw, pos = _inputs(t, hq, hkv, max_pos, seed=t)
    cache = orc.cos_sin_cache(d, max_pos, 10000.0)
    q, k, v = qkv.split([hq * d, hkv * d, hkv * d], dim=-1)
    wq, wk = orc.qk_norm_rope(q, k, pos, qw, kw, cache, 1e-6, d)
    qkv_d = qkv.to(DEV)
    qd, kd, vd = qkv_d.split([hq * d, hkv * d, hkv * d], dim=-1)
    gq, gk = ops.qk_norm_rope(pos.to(DEV), qd, kd, qw.to(DEV), kw.to(DEV), cache.to(DEV), 1e-6, d, inplace=inplace)
    torch.cuda.synchronize()
    close_enough(gq, wq, "q")
    close_enough(gk, wk, "k")
    assert torch.equal(vd.cpu(), v)                                  # v untouched
    if inplace:
        assert gq.data_ptr() == qd.data_ptr() and gk.data_ptr() == kd.data_ptr()
    else:
        assert torch.equal(qd.cpu(), q) and torch.equal(kd.cpu(), k)  # inputs untouched


def test_empty_and_bad_arguments(ops):
    from vllm_trianglemix_b200 import TmxError
    d = 128
    cache = orc.cos_sin_cache(d, 16, 10000.0).to(DEV)
    w = torch.ones(d, dtype=torch.bfloat16, device=DEV)
    q0 = torch.empty(0, 4 * d, dtype=torch.bfloat16, device=DEV)
    k0 = torch.empty(0, 2 * d, dtype=torch.bfloat16, device=DEV)
    gq, gk = ops.qk_norm_rope(torch.empty(0, dtype=torch.int64, device=DEV), q0, k0, w, w, cache, 1e-6)
    assert gq.shape == q0.shape and gk.shape == k0.shape
    q = torch.zeros(3, 4 * d, dtype=torch.bfloat16, device=DEV)
    k = torch.zeros(3, 2 * d, dtype=torch.bfloat16, device=DEV)
    pos = torch.zeros(3, dtype=torch.int64, device=DEV)
    with pytest.raises(TmxError):
        ops.qk_norm_rope(pos, q.float(), k, w, w, cache, 1e-6)           # dtype
    with pytest.raises(TmxError):
        ops.qk_norm_rope(pos[:2], q, k, w, w, cache, 1e-6)               # positions length
    with pytest.raises(TmxError):
        ops.qk_norm_rope(pos, q[:, :64 * 7], k, w, w, cache, 1e-6)       # not a multiple of head_dim
    with pytest.raises(TmxError):
        ops.qk_norm_rope(pos, q.cpu(), k.cpu(), w.cpu(), w.cpu(), cache.cpu(), 1e-6)   # no CPU path
    # all-zero rows: rsqrt(eps) * 0 = 0, no NaN
    gq, gk = ops.qk_norm_rope(pos, q, k, w, w, cache, 1e-6)
    assert not gq.float().abs().sum().item() and not gk.float().abs().sum().item()


@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("t,hkv,block_size,num_blocks", [(42, 8, 16, 8), (1, 1, 16, 1), (777, 2, 32, 40)])
def test_reshape_and_cache_flash_vs_reference_loop(ops, layout, t, hkv, block_size, num_blocks):
    """Mirrors the reference's own test (tests/kernels/attention/test_cache.py:230-370):
    random unique slots, key/value as views of a packed qkv tensor, both cache layouts."""
    g = torch.Generator().manual_seed(t)
    d = 128
    qkv = torch.randn(t, 3, hkv, d, generator=g).bfloat16()
    _, key, value = qkv.unbind(dim=1)
    slots = torch.randperm(block_size * num_blocks, generator=g)[:t].to(torch.int64)
    if t > 4:
        slots[3] = -1                                                    # a padding token
    shape = (2, num_blocks, block_size, hkv, d) if layout == "NHD" else (2, num_blocks, hkv, block_size, d)
    kv = torch.randn(shape, generator=g).bfloat16()
    want = kv.clone()
    wk, wv = want.unbind(0)
    if layout == "HND":
        wk, wv = wk.permute(0, 2, 1, 3), wv.permute(0, 2, 1, 3)
    orc.reshape_and_cache_flash(key, value, wk, wv, slots)
    kv_d = kv.to(DEV)
    kc, vc = kv_d.unbind(0)
    if layout == "HND":
        kc, vc = kc.permute(0, 2, 1, 3), vc.permute(0, 2, 1, 3)
    qkv_d = qkv.to(DEV)
    _, key_d, value_d = qkv_d.unbind(dim=1)
    ops.reshape_and_cache_flash(key_d, value_d, kc, vc, slots.to(DEV), "auto", None, None)
    torch.cuda.synchronize()
    assert torch.equal(kv_d.cpu(), want)
    from vllm_trianglemix_b200 import TmxError
    with pytest.raises(TmxError):
        ops.reshape_and_cache_flash(key_d, value_d, kc, vc, slots.to(DEV), "fp8", None, None)


def test_fused_cache_write_matches_separate_steps(ops):
    """norm+rope+cache in one launch == norm+rope, then the cache write of its k and the raw v."""
    t, hq, hkv, d, max_pos, bs, nb = 2500, 32, 8, 128, 40960, 16, 200
    qkv, qw, kw, pos = _inputs(t, hq, hkv, max_pos, seed=7)
    cache = orc.cos_sin_cache(d, max_pos, 1000000.0).to(DEV)
    g = torch.Generator().manual_seed(8)
    slots = torch.randperm(bs * nb, generator=g)[:t].to(torch.int64)
    slots[5] = -1
    a = qkv.to(DEV)
    qa, ka, va = a.split([hq * d, hkv * d, hkv * d], dim=-1)
    kv1 = torch.zeros(2, nb, bs, hkv, d, dtype=torch.bfloat16, device=DEV)
    kv2 = torch.zeros_like(kv1)
    args = (pos.to(DEV), qa, ka, qw.to(DEV), kw.to(DEV), cache, 1e-6, d)
    q1, k1 = ops.qk_norm_rope(*args, v=va, key_cache=kv1[0], value_cache=kv1[1], slot_mapping=slots.to(DEV))
    assert ops.last_launch_count() == 1
    q2, k2 = ops.qk_norm_rope(*args)
    ops.reshape_and_cache_flash(k2.view(t, hkv, d), va.view(t, hkv, d), kv2[0], kv2[1], slots.to(DEV))
    torch.cuda.synchronize()
    assert torch.equal(q1, q2) and torch.equal(k1, k2) and torch.equal(kv1, kv2)
    # and against the oracle: cache rows hold exactly the k the kernel returned and the raw v
    wkc = torch.zeros(nb, bs, hkv, d, dtype=torch.bfloat16)
    wvc = torch.zeros_like(wkc)
    orc.reshape_and_cache_flash(k1.cpu().view(t, hkv, d), qkv[:, (hq + hkv) * d:].reshape(t, hkv, d), wkc, wvc, slots)
    assert torch.equal(kv1[0].cpu(), wkc) and torch.equal(kv1[1].cpu(), wvc)


@pytest.mark.parametrize("layer_idx", [0, 1])
def test_qwen3_block_end_to_end(tmx, layer_idx):
    """Qwen3Attention.forward between the projections (qwen3.py:160-190), hook wired: packed qkv in,
    attn_output out, against oracle prologue -> oracle masked attention; KV cache rows written."""
    t, hq, hkv, d, max_pos = 3000, 8, 2, 128, 4096
    si, sl, last = 4, 32, 64
    qkv, qw, kw, _ = _inputs(t, hq, hkv, max_pos, seed=11, qscale=1.0)
    pos = torch.arange(t, dtype=torch.int64)
    cfg = tmx.TriangleMixConfig(si, sl, last, triangle_layer_indices=[1])
    blk = tmx.Qwen3TriangleMixAttention(hq, hkv, d, max_position=max_pos, rms_norm_eps=1e-6, rope_theta=1000000.0,
                                        trianglemix_config=cfg, layer_idx=layer_idx, num_layers=2, device=DEV)
    blk.q_norm_weight.data.copy_(qw)
    blk.k_norm_weight.data.copy_(kw)
    assert torch.equal(blk.cos_sin_cache.cpu(), orc.cos_sin_cache(d, max_pos, 1000000.0))
    meta = tmx.TriangleMixMetadataBuilder.from_seq_lens([t], DEV)
    bs, nb = 16, (t + 15) // 16
    meta.slot_mapping = torch.arange(t, dtype=torch.int64, device=DEV)
    kv_cache = torch.zeros(tmx.TriangleMixBackend.get_kv_cache_shape(nb, bs, hkv, d), dtype=torch.bfloat16, device=DEV)
    out = blk(pos.to(DEV), qkv.to(DEV), kv_cache, meta)
    torch.cuda.synchronize()
    q, k, v = qkv.split([hq * d, hkv * d, hkv * d], dim=-1)
    wq, wk = orc.qk_norm_rope(q, k, pos, qw, kw, orc.cos_sin_cache(d, max_pos, 1000000.0), 1e-6, d)
    mode = aorc.TRIANGLE if layer_idx == 1 else aorc.DENSE_CAUSAL
    ref = aorc.attention_varlen_fp32(wq.view(t, hq, d), wk.view(t, hkv, d), v.reshape(t, hkv, d), [0, t], d ** -0.5,
                                     si, sl, last, mode)
    rep = aorc.error_report(out.cpu().view(t, hq, d), ref)
    assert rep["violations"] == 0, rep
    close_enough(kv_cache[0].view(-1, hkv * d)[:t], wk, "k cache")
    assert torch.equal(kv_cache[1].view(-1, hkv * d)[:t].cpu(), v)


@pytest.mark.parametrize("seed", range(int(os.environ.get("TMX_FUZZ_SEEDS", "30"))))
def test_prologue_fuzz_vs_oracle(ops, seed):
    """Random token counts, head counts, cache geometry, slot maps with padding, in place or not."""
    import random
    r = random.Random(seed)
    d = 128
    t = r.choice([1, 2, 7, 33, 64, 100, 257, 1000, r.randint(1, 3000)])
    hq, hkv = r.choice([(1, 1), (2, 1), (4, 1), (4, 2), (8, 1), (8, 2), (16, 8), (32, 8), (5, 5), (7, 1), (64, 8)])
    max_pos = r.choice([16, 4096, 40960])
    bs = r.choice([16, 32, 48, 64])
    nb = (t + bs - 1) // bs + r.randint(0, 3)
    inplace, with_cache, hnd = r.random() < 0.5, r.random() < 0.7, r.random() < 0.3
    qkv, qw, kw, pos = _inputs(t, hq, hkv, max_pos, seed=500 + seed)
    cache = orc.cos_sin_cache(d, max_pos, r.choice([10000.0, 1000000.0]))
    eps = r.choice([1e-6, 1e-5])
    g = torch.Generator().manual_seed(seed)
    slots = torch.randperm(bs * nb, generator=g)[:t].to(torch.int64)
    slots[torch.rand(t, generator=g) < 0.1] = -1
    q, k, v = qkv.split([hq * d, hkv * d, hkv * d], dim=-1)
    wq, wk = orc.qk_norm_rope(q, k, pos, qw, kw, cache, eps, d)
    shape = (2, nb, hkv, bs, d) if hnd else (2, nb, bs, hkv, d)
    kv = torch.randn(shape, generator=g).bfloat16()
    kv_d = kv.to(DEV)
    kc, vc = kv_d.unbind(0)
    if hnd:
        kc, vc = kc.permute(0, 2, 1, 3), vc.permute(0, 2, 1, 3)
    qkv_d = qkv.to(DEV)
    qd, kd, vd = qkv_d.split([hq * d, hkv * d, hkv * d], dim=-1)
    extra = dict(v=vd, key_cache=kc, value_cache=vc, slot_mapping=slots.to(DEV)) if with_cache else {}
    gq, gk = ops.qk_norm_rope(pos.to(DEV), qd, kd, qw.to(DEV), kw.to(DEV), cache.to(DEV), eps, d, inplace=inplace,
                              **extra)
    torch.cuda.synchronize()
    close_enough(gq, wq, "q")
    close_enough(gk, wk, "k")
    assert torch.equal(vd.cpu(), v)
    want = kv.clone()
    if with_cache:
        wkc, wvc = want.unbind(0)
        if hnd:
            wkc, wvc = wkc.permute(0, 2, 1, 3), wvc.permute(0, 2, 1, 3)
        orc.reshape_and_cache_flash(gk.cpu().view(t, hkv, d), v.reshape(t, hkv, d), wkc, wvc, slots)
    assert torch.equal(kv_d.cpu(), want)          # written rows exact, every other slot untouched
