"""The analyzer's gradient producer: GPU middle_qk_gradient against autograd over the
reference op's arithmetic with a gate on the Middle Q-K block (oracle.middle_gate_gradient_autograd), and the
selector on top of the unchanged TriangleMixAnalyzer."""
import pytest
import torch

from oracle import trianglemix_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def tmx():
    import vllm_trianglemix_b200 as m
    return m


def _inputs(n, hq, hkv, seed, qscale=1.0):
    g = torch.Generator().manual_seed(seed)
    return ((torch.randn(n, hq, 128, generator=g) * qscale).bfloat16(), torch.randn(n, hkv, 128, generator=g).bfloat16(),
            torch.randn(n, hkv, 128, generator=g).bfloat16(), torch.randn(n, hq, 128, generator=g).bfloat16())


@pytest.mark.parametrize("kernel", ["auto", "mma"])
@pytest.mark.parametrize("n,hq,hkv,si,sl,last", [(700, 4, 2, 4, 32, 64), (1500, 8, 2, 4, 32, 64), (333, 2, 2, 1, 7, 20),
                                                 (64, 4, 1, 4, 32, 64)])
def test_middle_qk_gradient_vs_autograd(tmx, n, hq, hkv, si, sl, last, kernel):
    q, k, v, do = _inputs(n, hq, hkv, seed=n)
    scale = 128 ** -0.5
    got = tmx.middle_qk_gradient(q.to(DEV), k.to(DEV), v.to(DEV), do.to(DEV), [n], scale, si, sl, last, kernel=kernel).cpu()
    ref = orc.middle_gate_gradient_autograd(q, k, v, do, scale, si, sl, last)
    assert got.shape == ref.shape == (n, hq)
    # bf16 outputs enter a 128-term dot product with dO ~ N(0, 1): tolerance 5e-2 absolute + 5 % relative,
    # and 2 % in the Frobenius norm
    assert bool(((got - ref).abs() <= 5e-2 + 5e-2 * ref.abs()).all()), float((got - ref).abs().max())
    if float(ref.norm()) > 0:
        assert float((got - ref).norm() / ref.norm()) < 2e-2
    # rows that have no middle keys (the first sl + si + 2 rows and the last - 1 dense rows) give exactly 0
    mid = orc.piece_closed("middle", n, si, sl, last).any(axis=1)
    assert float(got[torch.from_numpy(~mid)].abs().max()) == 0.0


def test_selector_picks_the_layers_the_oracle_picks(tmx):
    n, hq, hkv, layers = 900, 4, 2, 5
    cfg = tmx.TriangleMixConfig(4, 32, 64)
    sel = tmx.TriangleLayerSelector(layers, cfg)
    ana = tmx.TriangleMixAnalyzer(layers)
    for layer in range(layers):
        # different score scales per layer: sharper attention concentrates on few keys
        q, k, v, do = _inputs(n, hq, hkv, seed=10 + layer, qscale=[0.2, 3.0, 1.0, 6.0, 0.5][layer])
        for cu in ([n], [400, 500]):                      # one sequence, then two
            g = sel.observe(layer, q.to(DEV), k.to(DEV), v.to(DEV), do.to(DEV), cu, 128 ** -0.5)
            s0, refs = 0, []
            for m in cu:
                refs.append(orc.middle_gate_gradient_autograd(q[s0:s0 + m], k[s0:s0 + m], v[s0:s0 + m], do[s0:s0 + m],
                                                              128 ** -0.5, 4, 32, 64))
                s0 += m
            ana.record_middle_qk_gradient(layer, torch.cat(refs).abs())
            assert g.shape == (n, hq)
    assert sel.select(2) == ana.get_triangle_layers(2)
    assert sel.to_config(3).triangle_layer_indices == ana.get_triangle_layers(3)
