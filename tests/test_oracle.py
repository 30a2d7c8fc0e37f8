"""Pins oracle/ against the golden vectors produced by the unmodified reference
(tests/golden/make_golden.py) and the values its own test script prints."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch

from oracle import trianglemix_oracle as orc

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
with open(os.path.join(GOLDEN, "masks.json")) as f:
    G = json.load(f)
SMALL = np.load(os.path.join(GOLDEN, "masks_small.npz"))
ATTN = np.load(os.path.join(GOLDEN, "attn_small.npz"))


def _sha(bits):
    return hashlib.sha256(np.ascontiguousarray(bits).tobytes()).hexdigest()


@pytest.mark.parametrize("case", G["cases"], ids=lambda c: "N%d_%d_%d_%d" % (c["N"], c["si"], c["sl"], c["num_last"]))
def test_closed_form_matches_reference(case):
    n, si, sl, last = case["N"], case["si"], case["sl"], case["num_last"]
    for piece in orc.PIECES:
        m = orc.piece_closed(piece, n, si, sl, last)
        assert int(m.sum()) == case[piece], piece
        assert _sha(orc.pack_bits(m)) == case[piece + "_sha"], piece
    add = orc.additive(orc.mask_closed(n, si, sl, last, orc.TRIANGLE))
    assert int(np.isinf(add).sum()) == case["additive_neg_inf"]
    assert int((add == 0).sum()) == case["additive_zero"]
    assert orc.attended_pairs(n, si, sl, last, orc.TRIANGLE) == case["triangle"]
    assert orc.attended_pairs(n, si, sl, last, orc.DENSE_CAUSAL) == case["dense"]


@pytest.mark.parametrize("case", [c for c in G["cases"] if c["N"] <= 130],
                         ids=lambda c: "N%d_%d_%d_%d" % (c["N"], c["si"], c["sl"], c["num_last"]))
def test_loops_match_reference_and_closed_form(case):
    n, si, sl, last = case["N"], case["si"], case["sl"], case["num_last"]
    for piece in ("streaming", "last", "middle"):
        loops = orc.piece_loops(piece, n, si, sl, last)
        key = "%s_%d_%d_%d_%d" % (piece, n, si, sl, last)
        assert np.array_equal(orc.pack_bits(loops > 0), SMALL[key])
        assert np.array_equal(loops > 0, orc.piece_closed(piece, n, si, sl, last))
    tri = orc.triangle01_loops(n, si, sl, last)
    assert np.array_equal(orc.pack_bits(tri > 0), SMALL["triangle_%d_%d_%d_%d" % (n, si, sl, last)])
    assert np.array_equal(orc.dense01(n) > 0, orc.piece_closed("dense", n, si, sl, last))
    # partition identity
    parts = sum(orc.piece_closed(p, n, si, sl, last).astype(np.int32) for p in ("streaming", "last", "middle"))
    assert np.array_equal(parts, orc.piece_closed("dense", n, si, sl, last).astype(np.int32))


@pytest.mark.parametrize("case", G["big"], ids=lambda c: "N%d" % c["N"])
def test_closed_form_big(case):
    n, si, sl, last = case["N"], case["si"], case["sl"], case["num_last"]
    m = orc.mask_closed(n, si, sl, last, orc.TRIANGLE)
    assert int(m.sum()) == case["triangle"]
    assert _sha(orc.pack_bits(m)) == case["triangle_sha"]


def test_reference_test_script_known_answers():
    # values the reference's tests/trianglemix_attention_test.py prints
    assert orc.attended_pairs(12, 2, 4, 6, orc.TRIANGLE) == 78
    assert orc.attended_pairs(16, 3, 6, 8, orc.TRIANGLE) == 136
    assert orc.attended_pairs(24, 4, 8, 16, orc.TRIANGLE) == 300
    assert orc.attended_pairs(4096, 4, 8, 16, orc.TRIANGLE) == 118378
    assert orc.attended_pairs(32768, 4, 32, 64, orc.TRIANGLE) == 3304518
    assert orc.attended_pairs(131072, 4, 32, 64, orc.TRIANGLE) == 13233222
    assert orc.piece_closed("streaming", 24, 4, 8, 16)[15, 1]     # the script's comment is wrong


def test_rows_subset_of_loops():
    full = orc.triangle01_loops(97, 4, 32, 64)
    rows = [0, 5, 33, 34, 96]
    assert np.array_equal(orc.triangle01_loops(97, 4, 32, 64, rows), full[rows])


def test_switch_and_threshold():
    assert [orc.should_use_triangle(i, 4, list(range(4))) for i in range(8)] == G["switch_first4_of_8"]
    assert [orc.should_use_triangle(i, 2, [5, 7]) for i in range(8)] == G["switch_explicit_5_7"]
    assert [orc.should_use_triangle(i, None, []) for i in range(4)] == G["switch_unset"]
    for n, is_none in G["threshold_none_at"].items():
        assert (orc.helper_mask(int(n), 0, 4, 32, 64, 4, None) is None) == is_none
    m = orc.helper_mask(2049, 5, 4, 32, 64, 4, None)
    assert m.shape == (2049, 2049) and np.array_equal(m == 0, orc.piece_closed("dense", 2049, 4, 32, 64))
    m = orc.helper_mask(2049, 3, 4, 32, 64, 4, None)
    assert np.array_equal(m == 0, orc.piece_closed("triangle", 2049, 4, 32, 64))


def test_rewrite_neg_inf():
    add = orc.additive(orc.mask_closed(16, 3, 6, 8, orc.TRIANGLE))
    r = orc.rewrite_neg_inf(add)
    assert bool(np.isinf(r).any()) == G["rewrite16"]["any_inf"]
    assert int((r == np.float32(-1e6)).sum()) == G["rewrite16"]["count_-1e6"]
    assert int((r == 0).sum()) == G["rewrite16"]["count_0"]
    assert orc.rewrite_neg_inf(np.zeros((0, 0), np.float32)).size == 0


def _bf16(a):
    return torch.from_numpy(a.copy()).view(torch.bfloat16)


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_attention_matches_reference(tag):
    s, hq, hkv, d, si, sl, last = [int(x) for x in ATTN[tag + "_params"]]
    q, k, v = _bf16(ATTN[tag + "_q"]), _bf16(ATTN[tag + "_k"]), _bf16(ATTN[tag + "_v"])
    scale = d ** -0.5
    for name, mode in (("tri", orc.TRIANGLE), ("dense", orc.DENSE_CAUSAL)):
        add = torch.from_numpy(orc.additive(orc.mask_closed(s, si, sl, last, mode)))
        out = orc.attention_fp32(q, k, v, add, scale)
        ref = torch.from_numpy(ATTN["%s_%s_out" % (tag, name)])
        assert torch.allclose(out, ref, atol=2e-6, rtol=1e-5)
        rows = orc.attention_rows_fp32(q[0], k[0], v[0], scale, s, si, sl, last, mode, 7, min(s, 71))
        assert torch.allclose(rows, ref[0, 7:min(s, 71)], atol=2e-6, rtol=1e-5)
        cu = [0, s]
        var = orc.attention_varlen_fp32(q[0], k[0], v[0], cu, scale, si, sl, last, mode, max_full=64)
        assert torch.allclose(var, ref[0], atol=2e-6, rtol=1e-5)
    if tag == "a":
        out = orc.attention_fp32(q, k, v, None, scale)
        assert torch.allclose(out, torch.from_numpy(ATTN["a_nomask_out"]), atol=2e-6, rtol=1e-5)


def test_error_report():
    a = torch.zeros(4, 4)
    b = torch.zeros(4, 4)
    b[0, 0] = 1.0
    rep = orc.error_report(a, b)
    assert rep["violations"] == 1 and rep["max_abs"] == 1.0 and abs(rep["fro_rel"] - 1.0) < 1e-6


def test_middle_gate_gradient_closed_form_equals_autograd():
    """The identity the GPU producer uses (vllm_trianglemix_b200/layer_selection.py): dL/dg = dO . m (1 - m)
    (O_mid - O_tri) with m = sigmoid(lse_mid - lse_tri), against autograd over the reference op's arithmetic."""
    import torch
    torch.manual_seed(0)
    n, hq, hkv, si, sl, last = 200, 4, 2, 3, 9, 17
    q, k, v, do = (torch.randn(n, h, 128).bfloat16() for h in (hq, hkv, hkv, hq))
    scale = 128 ** -0.5
    g = orc.middle_gate_gradient_autograd(q, k, v, do, scale, si, sl, last)
    tri = torch.from_numpy(orc.additive(orc.piece_closed("triangle", n, si, sl, last).astype(np.float32)))
    mid = torch.from_numpy(orc.additive(orc.piece_closed("middle", n, si, sl, last).astype(np.float32)))
    qf = q.float().transpose(0, 1)
    kf = k.float().repeat_interleave(hq // hkv, 1).transpose(0, 1)
    vf = v.float().repeat_interleave(hq // hkv, 1).transpose(0, 1)
    s = torch.matmul(qf, kf.transpose(-2, -1)) * scale
    m = torch.nan_to_num(torch.sigmoid(torch.logsumexp(s + mid, -1) - torch.logsumexp(s + tri, -1)))
    o_t = torch.matmul(torch.softmax(s + tri, -1), vf)
    o_m = torch.nan_to_num(torch.matmul(torch.softmax(s + mid, -1), vf))
    g2 = (m * (1 - m) * ((o_m - o_t) * do.float().transpose(0, 1)).sum(-1)).transpose(0, 1)
    assert float((g - g2).abs().max()) < 1e-4
    assert float(g.abs().max()) > 1e-2


# ---- head_dim 64 (the reference's own demo shape, tests/golden/make_golden_d64.py) ------------------------
ATTN64 = np.load(os.path.join(GOLDEN, "attn_d64.npz"))


def test_attention_d64_matches_reference():
    """The oracle is head_dim-agnostic: the quickstart demo (no mask => bidirectional, B = 2) and two masked GQA
    cases at D = 64 against outputs of the unmodified reference op."""
    q, k, v = (_bf16(ATTN64["demo_" + n]) for n in "qkv")
    out = orc.attention_fp32(q, k, v, None, 64 ** -0.5)
    assert torch.allclose(out, torch.from_numpy(ATTN64["demo_out"]), atol=2e-6, rtol=1e-5)
    for tag in ("t", "u"):
        b, s, hq, hkv, d, si, sl, last = [int(x) for x in ATTN64[tag + "_params"]]
        q, k, v = (_bf16(ATTN64["%s_%s" % (tag, n)]) for n in "qkv")
        for name, mode in (("tri", orc.TRIANGLE), ("dense", orc.DENSE_CAUSAL)):
            add = torch.from_numpy(orc.additive(orc.mask_closed(s, si, sl, last, mode)))
            out = orc.attention_fp32(q, k, v, add, d ** -0.5)
            assert torch.allclose(out, torch.from_numpy(ATTN64["%s_%s_out" % (tag, name)]), atol=2e-6, rtol=1e-5)
