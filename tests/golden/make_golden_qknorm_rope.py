"""Golden vectors for the fused q/k RMSNorm + RoPE prologue, produced by the
reference's OWN code in the build container:
  RMSNorm.forward_native          vllm/model_executor/layers/layernorm.py:170-206
  RotaryEmbedding.forward_native  vllm/model_executor/layers/rotary_embedding/base.py:66-95
  (+ _compute_cos_sin_cache :47-58, apply_rotary_emb_torch common.py:29-47)
called exactly as Qwen3Attention.forward does (vllm/model_executor/models/qwen3.py:166-174).
The classes are CustomOps whose constructors want a vLLM platform config (no device here), so
the unbound reference methods are run on a plain namespace carrying the attributes they read.

    cd /tmp && PYTHONDONTWRITEBYTECODE=1 PYTHONPATH=/root/reference python /root/repo/tests/golden/make_golden_qknorm_rope.py
"""
import os
import types

import numpy as np
import torch
from vllm.model_executor.layers.layernorm import RMSNorm
from vllm.model_executor.layers.rotary_embedding.base import RotaryEmbedding

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "qknorm_rope.npz")


def ref_norm(weight, eps):
    ns = types.SimpleNamespace(hidden_size=weight.shape[0], variance_size_override=None,
                               variance_epsilon=eps, has_weight=True, weight=weight)
    return lambda x: RMSNorm.forward_native(ns, x)


def ref_rope(head, max_pos, base, dtype):
    ns = types.SimpleNamespace(head_size=head, rotary_dim=head, max_position_embeddings=max_pos, base=base,
                               is_neox_style=True, dtype=dtype)
    ns._compute_inv_freq = lambda b: RotaryEmbedding._compute_inv_freq(ns, b)
    ns.cos_sin_cache = RotaryEmbedding._compute_cos_sin_cache(ns).to(dtype)
    return ns, (lambda pos, q, k: RotaryEmbedding.forward_native(ns, pos, q, k))


def u16(t):
    return t.contiguous().view(torch.int16).numpy().copy()


def main():
    torch.manual_seed(0)
    out = {}
    for tag, (t, hq, hkv, max_pos) in {"a": (37, 4, 2, 4096), "b": (64, 8, 2, 40960)}.items():
        d = 128
        qw = (1.0 + 0.1 * torch.randn(d)).bfloat16()
        kw = (1.0 + 0.1 * torch.randn(d)).bfloat16()
        q = (torch.randn(t, hq * d) * 3).bfloat16()
        k = (torch.randn(t, hkv * d) * 0.5).bfloat16()
        pos = torch.randint(0, max_pos, (t,), dtype=torch.int64)
        pos[0], pos[-1] = 0, max_pos - 1
        qn, kn = ref_norm(qw, 1e-6), ref_norm(kw, 1e-6)
        ns, rope = ref_rope(d, max_pos, 1000000.0, torch.bfloat16)
        # qwen3.py:166-174
        q1 = qn(q.view(t, hq, d)).view(q.shape)
        k1 = kn(k.view(t, hkv, d)).view(k.shape)
        q2, k2 = rope(pos, q1, k1)
        out.update({tag + "_q": u16(q), tag + "_k": u16(k), tag + "_qw": u16(qw), tag + "_kw": u16(kw),
                    tag + "_pos": pos.numpy(), tag + "_cache": u16(ns.cos_sin_cache[pos]),
                    tag + "_qo": u16(q2), tag + "_ko": u16(k2),
                    tag + "_params": np.array([t, hq, hkv, d, max_pos])})
    # the cos/sin cache itself (first and last rows + checksum) to pin the cache construction
    ns, _ = ref_rope(128, 40960, 1000000.0, torch.bfloat16)
    out["cache_rows"] = u16(ns.cos_sin_cache[[0, 1, 2, 1000, 40959]])
    out["cache_sum"] = np.array([float(ns.cos_sin_cache.float().sum())])
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
