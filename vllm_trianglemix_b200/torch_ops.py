"""torch.library view of the C ABI: `torch.ops.trianglemix.*` (csrc/tmx_torch_ops.cpp, a TORCH_LIBRARY TU in the
style of the reference's csrc/torch_bindings.cpp:19,92-101).  The reference reaches attention backends through a
registered op so that piecewise CUDA graphs and torch.compile can see them (vllm/attention/layer.py:598-645);
these ops give the same visibility to callers that bypass vLLM's own `unified_attention_with_output`.
The ctypes path (ops.py) stays the default host binding; both call the same extern "C" entry points.
No fallback: a missing libtmx_torch.so raises."""
from __future__ import annotations

import os

import torch

from ._lib import TmxError

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtmx_torch.so")
_loaded = False


def load():
    """Register torch.ops.trianglemix.* (idempotent)."""
    global _loaded
    if _loaded:
        return torch.ops.trianglemix
    if not os.path.exists(LIB_PATH):
        raise TmxError("libtmx_torch.so not found at %s -- build it with `make -C vllm_trianglemix_b200/csrc torch` "
                       "(or __graft_entry__.build())" % LIB_PATH)
    torch.ops.load_library(LIB_PATH)
    # fake (meta) implementations: every op writes into caller-provided tensors and returns nothing, which is all
    # torch.compile / FakeTensor tracing needs to know
    for name in ("prefill_attn", "chunked_prefill_attn", "qk_norm_rope", "reshape_and_cache_flash", "mask_bits",
                 "mask_additive"):
        torch.library.register_fake("trianglemix::" + name)(lambda *a, **k: None)
    _loaded = True
    return torch.ops.trianglemix


def prefill_attn(q, k, v, cu_seqlens, max_seqlen, scale, si, sl, last, mode, out=None, bitmask=None, lse=None,
                 kernel: int = 0):
    """ops.prefill_attn through the registered op (same arguments, `kernel` as the TMX_KERNEL_* integer)."""
    t = load()
    if out is None:
        out = torch.empty_like(q)
    need = t.attn_workspace_bytes(cu_seqlens.numel() - 1, q.shape[0], max_seqlen, q.shape[1], k.shape[1], q.shape[2],
                                  si, sl, last, mode, kernel)
    ws = torch.empty(max(need, 256), dtype=torch.uint8, device=q.device)
    t.prefill_attn(out, q, k, v, cu_seqlens, max_seqlen, float(scale), si, sl, last, mode, bitmask, lse, ws, kernel)
    return out
