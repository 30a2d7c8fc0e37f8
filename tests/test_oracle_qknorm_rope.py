"""CPU: the q/k RMSNorm + RoPE oracle is pinned bit for bit to golden vectors that the
reference's own forward_native methods produced (tests/golden/make_golden_qknorm_rope.py),
and the product's host-side cos/sin table equals the reference's."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import qknorm_rope_oracle as orc            # noqa: E402
import vllm_trianglemix_b200 as tmx                     # noqa: E402

G = np.load(os.path.join(ROOT, "tests", "golden", "qknorm_rope.npz"))


def bf(a):
    return torch.from_numpy(a.copy()).view(torch.bfloat16)


@pytest.mark.parametrize("tag", ["a", "b"])
def test_oracle_matches_reference_golden(tag):
    t, hq, hkv, d, max_pos = (int(x) for x in G[tag + "_params"])
    cache = orc.cos_sin_cache(d, max_pos, 1000000.0)
    pos = torch.from_numpy(G[tag + "_pos"])
    assert torch.equal(cache[pos].view(torch.int16), torch.from_numpy(G[tag + "_cache"]))
    q2, k2 = orc.qk_norm_rope(bf(G[tag + "_q"]), bf(G[tag + "_k"]), pos, bf(G[tag + "_qw"]), bf(G[tag + "_kw"]),
                              cache, 1e-6, d)
    assert q2.shape == (t, hq * d) and k2.shape == (t, hkv * d)
    assert torch.equal(q2.view(torch.int16), torch.from_numpy(G[tag + "_qo"]))
    assert torch.equal(k2.view(torch.int16), torch.from_numpy(G[tag + "_ko"]))


def test_cos_sin_cache_matches_reference():
    for fn in (orc.cos_sin_cache, tmx.compute_cos_sin_cache):
        cache = fn(128, 40960, 1000000.0)
        assert cache.dtype == torch.bfloat16 and cache.shape == (40960, 128)
        assert torch.equal(cache[[0, 1, 2, 1000, 40959]].view(torch.int16), torch.from_numpy(G["cache_rows"]))
        assert float(cache.float().sum()) == float(G["cache_sum"][0])


def test_rope_preserves_pair_norm_and_identity_at_position_zero():
    torch.manual_seed(1)
    cache = orc.cos_sin_cache(128, 64, 10000.0)
    x = torch.randn(5, 3, 128).bfloat16()
    y = orc.rope_neox(x, *cache[torch.zeros(5, dtype=torch.int64)].chunk(2, dim=-1))
    assert torch.equal(x, y)                      # cos 1, sin 0


def test_reshape_and_cache_oracle_roundtrip():
    torch.manual_seed(2)
    k = torch.randn(7, 2, 128).bfloat16()
    v = torch.randn(7, 2, 128).bfloat16()
    kc = torch.zeros(4, 16, 2, 128, dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    slots = torch.tensor([3, 17, -1, 63, 0, 32, 31])
    orc.reshape_and_cache_flash(k, v, kc, vc, slots)
    flat_k, flat_v = kc.view(64, 2, 128), vc.view(64, 2, 128)
    for i, s in enumerate(slots.tolist()):
        if s >= 0:
            assert torch.equal(flat_k[s], k[i]) and torch.equal(flat_v[s], v[i])
    assert int((flat_k.abs().sum(dim=(1, 2)) > 0).sum()) == 6
