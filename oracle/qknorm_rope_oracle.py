"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the step in front of attention in the
reference's Qwen3Attention.forward: per-head q/k RMSNorm followed by
NeoX rotary embedding, in the reference's own bf16 rounding order.

Follows (all under /root/reference):
  vllm/model_executor/models/qwen3.py:166-174                         (call order)
  vllm/model_executor/layers/layernorm.py:170-206                     (RMSNorm.forward_native)
  vllm/model_executor/layers/rotary_embedding/base.py:42-58,66-95     (cache + forward_native)
  vllm/model_executor/layers/rotary_embedding/common.py:29-47         (apply_rotary_emb_torch)

Parity status: PINNED -- tests/test_oracle_qknorm_rope.py checks this file bit for bit
against tests/golden/qknorm_rope.npz, which tests/golden/make_golden_qknorm_rope.py
produced by running the reference's unmodified methods in the build container.
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
import torch


def cos_sin_cache(head_size: int, max_pos: int, base: float, dtype=torch.bfloat16):
    """[max_pos, head_size]: cos | sin halves, computed in fp32 then cast (base.py:42-58, :36-37)."""
    inv_freq = 1.0 / (base ** (torch.arange(0, head_size, 2, dtype=torch.float) / head_size))
    t = torch.arange(max_pos, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(dtype)


def rms_norm(x, weight, eps: float):
    """layernorm.py:176-204: fp32 variance and scaling, cast to the input dtype, THEN times weight."""
    orig = x.dtype
    xf = x.to(torch.float32)
    var = xf.pow(2).mean(dim=-1, keepdim=True)
    xf = xf * torch.rsqrt(var + eps)
    return xf.to(orig) * weight


def rope_neox(x, cos, sin):
    """common.py:35-45 with is_neox_style=True: every product and the sum round to x.dtype."""
    cos = cos.unsqueeze(-2).to(x.dtype)
    sin = sin.unsqueeze(-2).to(x.dtype)
    x1, x2 = torch.chunk(x, 2, dim=-1)
    return torch.cat((x1 * cos - x2 * sin, x2 * cos + x1 * sin), dim=-1)


def qk_norm_rope(q, k, positions, q_weight, k_weight, cache, eps: float, head_size: int = 128):
    """q: [T, Hq*D], k: [T, Hkv*D] (bf16); positions int64 [T]; cache [max_pos, D].
    Returns new (q, k) of the same shapes -- qwen3.py:166-174."""
    t = q.shape[0]
    qn = rms_norm(q.view(t, -1, head_size), q_weight, eps)
    kn = rms_norm(k.view(t, -1, head_size), k_weight, eps)
    cs = cache.index_select(0, positions.flatten())
    cos, sin = cs.chunk(2, dim=-1)
    return rope_neox(qn, cos, sin).reshape(q.shape), rope_neox(kn, cos, sin).reshape(k.shape)


def reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping):
    """The reference's own restatement of its cache-write kernel, as its test spells it
    (tests/kernels/attention/test_cache.py:339-349): block = slot // block_size, offset =
    slot % block_size, cache[block, offset] = row; negative slots are padding and skipped
    (csrc/cache_kernels.cu:292-297).  key/value [T, Hkv, D];
    caches [num_blocks, block_size, Hkv, D], modified in place."""
    block_size = key_cache.shape[1]
    for i, slot in enumerate(slot_mapping.tolist()):
        if slot < 0:
            continue
        key_cache[slot // block_size, slot % block_size] = key[i]
        value_cache[slot // block_size, slot % block_size] = value[i]
