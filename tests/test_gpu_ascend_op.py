"""GPU parity of the reference's own op signature: AscendTriangleMixAttention.forward
(vllm/attention/ascend_trianglemix.py:52-108) called through the drop-in class, against the
reference op's arithmetic in fp32 (oracle.attention_fp32 = :86-106) with the mask the reference
helper would hand out (oracle.helper_mask = trianglemix_attention.py:309-340).

Covers what the class adds on top of the kernel: (B, S, H, D) batches, the `None` mask at
seq_len <= 2048 (=> BIDIRECTIONAL attention, reference :96-97), a caller attn_mask alone and
added to the layer's mask (reference :76-84), masks tagged by the generator, the -1e6 masks of
optimize_mask_for_npu, fp32 inputs, and the error behaviour."""
import warnings

import numpy as np
import pytest
import torch

from oracle import trianglemix_oracle as orc

pytestmark = pytest.mark.gpu
ATOL, RTOL = 1.5e-2, 1e-2
DEV = "cuda:0"


@pytest.fixture(scope="module")
def tmx():
    import vllm_trianglemix_b200 as m
    return m


def _rand(b, s, hq, hkv, d=128, seed=0, dtype=torch.bfloat16):
    g = torch.Generator().manual_seed(seed)
    q = torch.randn(b, s, hq, d, generator=g).to(dtype)
    k = torch.randn(b, s, hkv, d, generator=g).to(dtype)
    v = torch.randn(b, s, hkv, d, generator=g).to(dtype)
    return q, k, v


def _check(out, ref, what):
    rep = orc.error_report(out, ref, ATOL, RTOL)
    assert rep["violations"] == 0, (what, rep)


def _ref(q, k, v, add_mask, scale):
    """The reference op in fp32 on the bf16-rounded inputs, one batch row at a time (memory)."""
    m = None if add_mask is None else torch.from_numpy(np.ascontiguousarray(add_mask))
    outs = [orc.attention_fp32(q[b:b + 1].bfloat16(), k[b:b + 1].bfloat16(), v[b:b + 1].bfloat16(), m, scale)
            for b in range(q.shape[0])]
    return torch.cat(outs, dim=0)


# (i) + (ii): batches, GQA, both sides of the 2048 threshold, triangle and dense layers
@pytest.mark.parametrize("kernel", ["auto", "mma", "tcgen05"])
@pytest.mark.parametrize("b,s,hq,hkv", [(2, 96, 8, 2), (3, 96, 4, 4), (2, 2048, 4, 2), (3, 2049, 4, 2),
                                        (2, 5000, 4, 2), (3, 33, 2, 1), (2, 300, 8, 1)])
@pytest.mark.parametrize("layer_idx", [0, 1])
def test_forward_batches_and_threshold(tmx, b, s, hq, hkv, layer_idx, kernel):
    if kernel == "tcgen05" and (hq // hkv) % 2:
        pytest.skip("tcgen05 kernel pairs two query heads of one KV head")
    si, sl, last = 4, 32, 64
    cfg = tmx.TriangleMixConfig(si, sl, last, num_triangle_layers=1)      # layer 0 triangle, layer 1 dense
    scale = 128 ** -0.5
    op = tmx.AscendTriangleMixAttention(hq, 128, scale, trianglemix_config=cfg, layer_idx=layer_idx, num_layers=2,
                                        num_kv_heads=hkv, kernel=kernel)
    q, k, v = _rand(b, s, hq, hkv, seed=s + b)
    out = op(q.to(DEV), k.to(DEV), v.to(DEV))
    assert out.shape == (b, s, hq, 128) and out.dtype == torch.bfloat16 and out.is_contiguous()
    # what the reference helper returns: None at s <= 2048 (=> no mask at all, bidirectional)
    add = orc.helper_mask(s, layer_idx, si, sl, last, 1, None)
    assert (add is None) == (s <= 2048)
    _check(out.cpu(), _ref(q, k, v, add, scale), (b, s, hq, hkv, layer_idx, kernel))


def test_forward_without_helper_is_bidirectional(tmx):
    # no trianglemix_config / num_layers => no helper => plain unmasked attention (reference :45-50, :96-97)
    b, s, h = 2, 700, 4
    op = tmx.AscendTriangleMixAttention(h, 128, 0.09)
    assert op.trianglemix_attn is None and op.layer_idx == 0
    q, k, v = _rand(b, s, h, h, seed=5)
    out = op(q.to(DEV), k.to(DEV), v.to(DEV))
    _check(out.cpu(), _ref(q, k, v, None, 0.09), "no helper")
    causal = _ref(q, k, v, orc.additive(orc.mask_closed(s, 0, 0, 0, orc.DENSE_CAUSAL)), 0.09)
    assert orc.error_report(out.cpu(), causal)["violations"] > 0          # it really is NOT causal


def _random_additive(s, seed, keep=0.3):
    g = np.random.default_rng(seed)
    m = g.random((s, s)) < keep
    m[np.arange(s), np.arange(s)] = True          # no fully masked row (the reference would return NaN)
    return np.where(m, np.float32(0), np.float32(-np.inf))


# (iii) caller attn_mask: alone, and added to the layer's mask
@pytest.mark.parametrize("mask_dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("kernel", ["auto", "mma"])
def test_forward_user_mask_alone(tmx, kernel, mask_dtype):
    b, s, hq, hkv = 2, 333, 4, 2
    add = _random_additive(s, 1)
    op = tmx.AscendTriangleMixAttention(hq, 128, 0.1, num_kv_heads=hkv, kernel=kernel)
    q, k, v = _rand(b, s, hq, hkv, seed=7)
    out = op(q.to(DEV), k.to(DEV), v.to(DEV), attn_mask=torch.from_numpy(add).to(mask_dtype).to(DEV))
    _check(out.cpu(), _ref(q, k, v, add, 0.1), ("user mask", kernel))


@pytest.mark.parametrize("layer_idx", [0, 1])
def test_forward_user_mask_plus_layer_mask(tmx, layer_idx):
    b, s, hq, hkv = 2, 2100, 4, 2
    si, sl, last = 4, 32, 64
    cfg = tmx.TriangleMixConfig(si, sl, last, num_triangle_layers=1)
    op = tmx.AscendTriangleMixAttention(hq, 128, 0.1, trianglemix_config=cfg, layer_idx=layer_idx, num_layers=2,
                                        num_kv_heads=hkv)
    user = _random_additive(s, 2, keep=0.6)
    tri = orc.helper_mask(s, layer_idx, si, sl, last, 1, None)
    q, k, v = _rand(b, s, hq, hkv, seed=9)
    out = op(q.to(DEV), k.to(DEV), v.to(DEV), attn_mask=torch.from_numpy(user).to(DEV))
    _check(out.cpu(), _ref(q, k, v, user + tri, 0.1), ("user + layer mask", layer_idx))   # reference :82-84


def test_forward_generator_mask_and_inplace_edit(tmx):
    # BASELINE config 1's route: the caller fetches the mask from the generator (no threshold there)
    # and passes it as attn_mask.  The tagged tensor takes the analytic path; an in-place edit voids the tag.
    s, hq, hkv = 1024, 4, 2
    si, sl, last = 4, 32, 64
    cfg = tmx.TriangleMixConfig(si, sl, last)
    gen = tmx.TriangleMixMaskGenerator(cfg)
    op = tmx.AscendTriangleMixAttention(hq, 128, 0.1, num_kv_heads=hkv)
    q, k, v = _rand(1, s, hq, hkv, seed=3)
    mask = gen.get_attention_mask(s, 0, torch.device(DEV), torch.float32, use_triangle=True)
    out = op(q.to(DEV), k.to(DEV), v.to(DEV), attn_mask=mask)
    add = orc.additive(orc.mask_closed(s, si, sl, last, orc.TRIANGLE))
    _check(out.cpu(), _ref(q, k, v, add, 0.1), "generator mask")
    mask[:, 2] = float("-inf")                       # caller edits the cached tensor in place
    mask[2, 2] = 0.0
    add[:, 2] = -np.inf
    add[2, 2] = 0.0
    out2 = op(q.to(DEV), k.to(DEV), v.to(DEV), attn_mask=mask)
    _check(out2.cpu(), _ref(q, k, v, add, 0.1), "edited generator mask")
    assert not torch.equal(out, out2)


def test_forward_accepts_optimized_mask(tmx):
    # optimize_mask_for_npu rewrites -inf to -1e6 (reference :135-139); the reference op accepts the result
    s, h = 400, 2
    si, sl, last = 4, 32, 64
    gen = tmx.TriangleMixMaskGenerator(tmx.TriangleMixConfig(si, sl, last))
    mask = gen.get_attention_mask(s, 0, torch.device(DEV), torch.float32, use_triangle=True)
    opt = tmx.AscendNPUTriangleMixOptimizer.optimize_mask_for_npu(mask)
    assert not torch.isinf(opt).any()
    op = tmx.AscendTriangleMixAttention(h, 128, 0.1)
    q, k, v = _rand(2, s, h, h, seed=4)
    out = op(q.to(DEV), k.to(DEV), v.to(DEV), attn_mask=opt)
    _check(out.cpu(), _ref(q, k, v, opt.cpu().numpy(), 0.1), "-1e6 mask")


# (iv) fp32 inputs: computed in bf16, returned in the input dtype
def test_forward_fp32_inputs(tmx):
    b, s, h = 2, 500, 4
    op = tmx.AscendTriangleMixAttention(h, 128, 0.1)
    q, k, v = _rand(b, s, h, h, seed=6, dtype=torch.float32)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        out = op(q.to(DEV), k.to(DEV), v.to(DEV))
    assert out.dtype == torch.float32 and any("bfloat16" in str(x.message) for x in w)
    _check(out.cpu(), _ref(q, k, v, None, 0.1), "fp32 inputs")


# (v) error behaviour: no silent fallback
def test_forward_errors(tmx):
    s, h = 64, 4
    op = tmx.AscendTriangleMixAttention(h, 128, 0.1)
    q, k, v = _rand(1, s, h, h, seed=8)
    bad = torch.zeros(s, s, device=DEV)
    bad[3, 1] = -2.5
    with pytest.raises(tmx.TmxError, match="other than 0"):
        op(q.to(DEV), k.to(DEV), v.to(DEV), attn_mask=bad)
    with pytest.raises(tmx.TmxError, match="seq_len, seq_len"):
        op(q.to(DEV), k.to(DEV), v.to(DEV), attn_mask=torch.zeros(s, s + 1, device=DEV))
    with pytest.raises(tmx.TmxError, match="num_kv_heads"):
        op(q.to(DEV), k[:, :, :2].to(DEV), v[:, :, :2].to(DEV))            # GQA without num_kv_heads=
    with pytest.raises(tmx.TmxError, match="CUDA"):
        op(q, k, v)                                                         # CPU tensors: no CPU path
    q32 = torch.randn(1, s, h, 32).bfloat16().to(DEV)
    with pytest.raises(tmx.TmxError, match="head_dim|shape"):
        op(q32, q32, q32)                                                   # head_dim must be 128 or 64
