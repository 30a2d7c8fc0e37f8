"""Golden vectors at head_dim 64 from the UNMODIFIED reference (build container only).

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_d64.py  ->  tests/golden/attn_d64.npz

"demo" is the reference's only in-tree call of the op, examples/trianglemix_quickstart.py:149-188:
AscendTriangleMixAttention(num_heads=8, head_dim=64, scale=64**-0.5, TriangleMixConfig(4, 8, 16,
num_triangle_layers=1), layer_idx=0, num_layers=4) on (2, 32, 8, 64) tensors with no attn_mask -- at
seq_len <= 2048 the helper hands out no mask, so the op attends bidirectionally
(ascend_trianglemix.py:96-97).  "t" / "u" are masked GQA cases like attn_small.npz.  Inputs are bf16-rounded
and the reference runs in fp32."""
import os

import numpy as np
import torch

from make_golden import load_reference

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    mods = load_reference()
    tm, asc = mods["trianglemix_attention"], mods["ascend_trianglemix"]
    dev = torch.device("cpu")
    torch.manual_seed(64)
    out = {}
    # the quickstart demo, verbatim constructor arguments
    cfg = tm.TriangleMixConfig(num_sink_tokens=4, sliding_window_size=8, num_last_tokens=16, num_triangle_layers=1)
    op = asc.AscendTriangleMixAttention(num_heads=8, head_dim=64, scale=1.0 / (64 ** 0.5), trianglemix_config=cfg,
                                        layer_idx=0, num_layers=4)
    q = torch.randn(2, 32, 8, 64).bfloat16()
    k = torch.randn(2, 32, 8, 64).bfloat16()
    v = torch.randn(2, 32, 8, 64).bfloat16()
    out["demo_out"] = op(q.float(), k.float(), v.float()).numpy()
    for n, t in (("q", q), ("k", k), ("v", v)):
        out["demo_" + n] = t.view(torch.int16).numpy()
    # masked GQA cases (mask passed explicitly, K/V heads repeated: the reference op is MHA-only)
    for tag, (b, s, hq, hkv, si, sl, last) in {"t": (1, 300, 4, 2, 4, 32, 64), "u": (2, 130, 4, 1, 2, 8, 12)}.items():
        d = 64
        q = torch.randn(b, s, hq, d).bfloat16()
        k = torch.randn(b, s, hkv, d).bfloat16()
        v = torch.randn(b, s, hkv, d).bfloat16()
        op = asc.AscendTriangleMixAttention(hq, d, d ** -0.5)
        gen = tm.TriangleMixMaskGenerator(tm.TriangleMixConfig(si, sl, last))
        kk = k.float().repeat_interleave(hq // hkv, dim=2)
        vv = v.float().repeat_interleave(hq // hkv, dim=2)
        for mode, tri in (("tri", True), ("dense", False)):
            mask = gen.get_attention_mask(s, 0, dev, torch.float32, use_triangle=tri)
            out["%s_%s_out" % (tag, mode)] = op.forward(q.float(), kk, vv, attn_mask=mask).numpy()
        for n, t in (("q", q), ("k", k), ("v", v)):
            out["%s_%s" % (tag, n)] = t.view(torch.int16).numpy()
        out[tag + "_params"] = np.array([b, s, hq, hkv, d, si, sl, last])
    np.savez_compressed(os.path.join(HERE, "attn_d64.npz"), **out)
    print({k_: v_.shape for k_, v_ in out.items()})


if __name__ == "__main__":
    main()
