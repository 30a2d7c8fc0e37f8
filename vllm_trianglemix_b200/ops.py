"""Thin torch wrappers over the C ABI (include/tmx.h): allocate outputs, pass raw
device pointers and the current CUDA stream, turn status codes into exceptions.

Style reference: vllm/_custom_ops.py:153-200 of the reference ("Python wrapper
allocates outputs, calls the native op").  No eager/CPU fallback exists.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib

DENSE_CAUSAL, TRIANGLE, NONE = 0, 1, 2
PIECES = {"streaming": 0, "last": 1, "middle": 2, "triangle": 3, "dense": 4}
KERNELS = {"auto": 0, "mma": 1, "tcgen05": 2, "pair": 3}
_DTYPES = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}


def _stream(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _require_cuda(device) -> torch.device:
    device = torch.device(device)
    if device.type != "cuda":
        raise _lib.TmxError(
            "vllm_trianglemix_b200 runs on CUDA (sm_100a) only; got device %r. There is no CPU path." % (device,))
    return device


def _dtype_code(dtype) -> int:
    if dtype not in _DTYPES:
        raise _lib.TmxError("mask dtype must be float32, bfloat16 or float16, got %r" % (dtype,))
    return _DTYPES[dtype]


def bitmask_stride_words(n: int) -> int:
    """Words per bitmask row: ceil(n/32) rounded up to a multiple of 4 (16-byte rows)."""
    return max(4, ((n + 31) // 32 + 3) // 4 * 4)


def mask_bits(n: int, si: int, sl: int, last: int, piece="triangle", device="cuda",
              out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Packed bitmask [n, stride] int32 (bit b of word w of row i = piece[i, 32w+b])."""
    device = _require_cuda(device)
    lib = _lib.load()
    stride = bitmask_stride_words(n)
    if out is None:
        out = torch.empty((n, stride), dtype=torch.int32, device=device)
    with torch.cuda.device(device):
        st = lib.tmx_mask_bits(out.data_ptr(), stride, n, si, sl, last,
                               PIECES[piece] if isinstance(piece, str) else int(piece), _stream(device))
    _lib.check(st, "tmx_mask_bits")
    return out


def mask_dense01(n: int, si: int, sl: int, last: int, piece: str, device, dtype) -> torch.Tensor:
    device = _require_cuda(device)
    lib = _lib.load()
    out = torch.empty((n, n), dtype=dtype, device=device)
    with torch.cuda.device(device):
        st = lib.tmx_mask_dense01(out.data_ptr(), _dtype_code(dtype), n, n, si, sl, last,
                                  PIECES[piece], _stream(device))
    _lib.check(st, "tmx_mask_dense01")
    return out


def mask_additive(n: int, si: int, sl: int, last: int, mode: int, device, dtype) -> torch.Tensor:
    device = _require_cuda(device)
    lib = _lib.load()
    out = torch.empty((n, n), dtype=dtype, device=device)
    with torch.cuda.device(device):
        st = lib.tmx_mask_additive(out.data_ptr(), _dtype_code(dtype), n, n, si, sl, last, int(mode),
                                   _stream(device))
    _lib.check(st, "tmx_mask_additive")
    return out


def rewrite_neg_inf(mask: torch.Tensor, fill: float = -1e6) -> torch.Tensor:
    device = _require_cuda(mask.device)
    lib = _lib.load()
    src = mask.contiguous()
    out = torch.empty_like(src)
    with torch.cuda.device(device):
        st = lib.tmx_mask_rewrite_neg_inf(out.data_ptr(), src.data_ptr(), _dtype_code(src.dtype),
                                          src.numel(), float(fill), _stream(device))
    _lib.check(st, "tmx_mask_rewrite_neg_inf")
    return out


def bits_from_additive(mask: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """[N, N] additive mask (0 / -inf) -> (packed bitmask, device int32 count of other values)."""
    device = _require_cuda(mask.device)
    lib = _lib.load()
    if mask.dim() != 2 or mask.shape[0] != mask.shape[1]:
        raise _lib.TmxError("additive mask must be square [N, N], got %r" % (tuple(mask.shape),))
    if mask.stride(1) != 1:
        mask = mask.contiguous()
    n = mask.shape[0]
    stride = bitmask_stride_words(n)
    out = torch.empty((n, stride), dtype=torch.int32, device=device)
    bad = torch.zeros(1, dtype=torch.int32, device=device)
    with torch.cuda.device(device):
        st = lib.tmx_mask_bits_from_additive(out.data_ptr(), stride, mask.data_ptr(),
                                             _dtype_code(mask.dtype), mask.stride(0), n,
                                             bad.data_ptr(), _stream(device))
    _lib.check(st, "tmx_mask_bits_from_additive")
    return out, bad


def block_index(cu_seqlens: torch.Tensor, total_t: int, block_m: int, block_n: int, si: int, sl: int,
                last: int, mode: int):
    """(ranges [slots,4], counts [slots], full [slots,2]) int32 on the device of cu_seqlens."""
    device = _require_cuda(cu_seqlens.device)
    lib = _lib.load()
    b = cu_seqlens.numel() - 1
    slots = total_t // block_m + b
    ranges = torch.empty((slots, 4), dtype=torch.int32, device=device)
    counts = torch.empty((slots,), dtype=torch.int32, device=device)
    full = torch.empty((slots, 2), dtype=torch.int32, device=device)
    cu = cu_seqlens.to(torch.int32).contiguous()
    with torch.cuda.device(device):
        st = lib.tmx_block_index(ranges.data_ptr(), counts.data_ptr(), full.data_ptr(), cu.data_ptr(),
                                 b, total_t, block_m, block_n, si, sl, last, int(mode), _stream(device))
    _lib.check(st, "tmx_block_index")
    return ranges, counts, full


_workspaces = {}


def _workspace(device, nbytes: int) -> torch.Tensor:
    key = (device.index if device.index is not None else torch.cuda.current_device(),
           torch.cuda.current_stream(device).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def prefill_attn(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cu_seqlens: torch.Tensor,
                 max_seqlen: int, scale: float, si: int, sl: int, last: int, mode: int,
                 out: Optional[torch.Tensor] = None, bitmask: Optional[torch.Tensor] = None,
                 return_lse: bool = False, kernel: str = "auto"):
    """Masked GQA prefill attention in vLLM's varlen layout.

    q [T, Hq, D], k/v [T, Hkv, D] bf16 (last two dims contiguous), cu_seqlens int32 [B+1] on
    the same device.  Returns out [T, Hq, D] bf16 (and lse [Hq, T] fp32 if asked).
    """
    device = _require_cuda(q.device)
    lib = _lib.load()
    if q.dtype != torch.bfloat16 or k.dtype != torch.bfloat16 or v.dtype != torch.bfloat16:
        raise _lib.TmxError("prefill_attn needs bfloat16 q/k/v, got %s/%s/%s" % (q.dtype, k.dtype, v.dtype))
    if q.dim() != 3 or k.dim() != 3 or v.shape != k.shape or k.shape[0] != q.shape[0]:
        raise _lib.TmxError("expected q [T,Hq,D], k/v [T,Hkv,D]; got %r %r %r" %
                            (tuple(q.shape), tuple(k.shape), tuple(v.shape)))
    t, hq, d = q.shape
    hkv = k.shape[1]

    def tok_major(x):
        return x if (x.stride(2) == 1 and x.stride(1) == x.shape[2]) else x.contiguous()

    q, k, v = tok_major(q), tok_major(k), tok_major(v)
    if k.stride(0) != v.stride(0):
        k, v = k.contiguous(), v.contiguous()
    if out is None:
        out = torch.empty((t, hq, d), dtype=torch.bfloat16, device=device)
    elif out.dtype != torch.bfloat16 or out.stride(2) != 1 or out.stride(1) != d:
        raise _lib.TmxError("out must be bf16 [T,Hq,D] with contiguous heads")
    cu = cu_seqlens
    if cu.dtype != torch.int32 or not cu.is_contiguous():
        cu = cu.to(torch.int32).contiguous()
    b = cu.numel() - 1
    kcode = KERNELS[kernel]
    lse = torch.empty((hq, t), dtype=torch.float32, device=device) if return_lse else None
    with torch.cuda.device(device):
        need = lib.tmx_attn_workspace_bytes(b, t, max_seqlen, hq, hkv, d, si, sl, last, int(mode), kcode)
        ws = _workspace(device, need)
        st = lib.tmx_prefill_attn(
            out.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), cu.data_ptr(), b, t, max_seqlen,
            hq, hkv, d, q.stride(0), k.stride(0), out.stride(0), float(scale), si, sl, last, int(mode),
            None if bitmask is None else bitmask.data_ptr(),
            0 if bitmask is None else bitmask.stride(0),
            None if lse is None else lse.data_ptr(), ws.data_ptr(), ws.numel(), kcode, _stream(device))
    _lib.check(st, "tmx_prefill_attn")
    return (out, lse) if return_lse else out


def chunked_prefill_attn(q: torch.Tensor, cu_seqlens_q: torch.Tensor, seq_lens: torch.Tensor, max_query_len: int,
                         scale: float, si: int, sl: int, last: int, mode: int, *,
                         key_cache: Optional[torch.Tensor] = None, value_cache: Optional[torch.Tensor] = None,
                         block_table: Optional[torch.Tensor] = None,
                         k: Optional[torch.Tensor] = None, v: Optional[torch.Tensor] = None,
                         cu_seqlens_k: Optional[torch.Tensor] = None,
                         out: Optional[torch.Tensor] = None, return_lse: bool = False, kernel: str = "auto"):
    """Attention of the LAST q_len rows of each sequence over all seq_lens[b] keys (chunked prefill,
    prefix-cache hits, decode): the context path of the reference backend,
    flash_attn_varlen_func(q, key_cache, value_cache, cu_seqlens_q, seqused_k, block_table)
    (vllm/v1/attention/backends/flash_attn.py:518-548).

    q [Tq, Hq, D] bf16; seq_lens int32 [B] (context + new tokens).  K/V either from the paged cache
    (key_cache / value_cache [num_blocks, block_size, Hkv, D] NHD views + block_table int32
    [B, max_blocks]) or contiguous (k / v [sum seq_lens, Hkv, D] + cu_seqlens_k)."""
    device = _require_cuda(q.device)
    lib = _lib.load()
    if q.dtype != torch.bfloat16 or q.dim() != 3:
        raise _lib.TmxError("chunked_prefill_attn needs bf16 q [Tq, Hq, D], got %s %r" % (q.dtype, tuple(q.shape)))
    t, hq, d = q.shape
    if not (q.stride(2) == 1 and q.stride(1) == d):
        q = q.contiguous()
    if out is None:
        out = torch.empty((t, hq, d), dtype=torch.bfloat16, device=device)
    elif out.dtype != torch.bfloat16 or out.stride(2) != 1 or out.stride(1) != d:
        raise _lib.TmxError("out must be bf16 [Tq,Hq,D] with contiguous heads")

    def i32(x):
        return x if (x.dtype == torch.int32 and x.is_contiguous()) else x.to(torch.int32).contiguous()

    cu_q, lens = i32(cu_seqlens_q), i32(seq_lens)
    b = lens.numel()
    if cu_q.numel() != b + 1:
        raise _lib.TmxError("cu_seqlens_q has %d entries for %d sequences" % (cu_q.numel(), b))
    if block_table is not None:
        if key_cache is None or value_cache is None:
            raise _lib.TmxError("block_table needs key_cache and value_cache")
        blk, tok, head, page = _cache_strides(key_cache, value_cache)
        hkv = key_cache.shape[2]
        if head != d or blk != page * tok:
            raise _lib.TmxError("the paged read path needs the NHD cache layout [num_blocks, block_size, Hkv, D] "
                                "(FlashAttention's, flash_attn.py:78-87); got strides %r" % (key_cache.stride(),))
        bt = block_table if (block_table.dtype == torch.int32 and block_table.stride(-1) == 1) \
            else block_table.to(torch.int32).contiguous()
        if bt.dim() != 2 or bt.shape[0] < b:
            raise _lib.TmxError("block_table must be int32 [B, max_blocks]")
        kptr, vptr, kv_stride, kv_rows = key_cache.data_ptr(), value_cache.data_ptr(), tok, key_cache.shape[0] * page
        cuk_ptr, bt_ptr, bt_stride = None, bt.data_ptr(), bt.stride(0)
    else:
        if k is None or v is None or cu_seqlens_k is None:
            raise _lib.TmxError("pass either (key_cache, value_cache, block_table) or (k, v, cu_seqlens_k)")
        if k.dtype != torch.bfloat16 or v.dtype != torch.bfloat16 or k.dim() != 3 or v.shape != k.shape:
            raise _lib.TmxError("k / v must be bf16 [sum seq_lens, Hkv, D]")

        def tok_major(x):
            return x if (x.stride(2) == 1 and x.stride(1) == x.shape[2]) else x.contiguous()
        k, v = tok_major(k), tok_major(v)
        if k.stride(0) != v.stride(0):
            k, v = k.contiguous(), v.contiguous()
        hkv, page = k.shape[1], 0
        cu_k = i32(cu_seqlens_k)
        kptr, vptr, kv_stride, kv_rows = k.data_ptr(), v.data_ptr(), k.stride(0), k.shape[0]
        cuk_ptr, bt_ptr, bt_stride = cu_k.data_ptr(), None, 0
    kcode = KERNELS[kernel]
    lse = torch.empty((hq, t), dtype=torch.float32, device=device) if return_lse else None
    with torch.cuda.device(device):
        need = lib.tmx_attn_workspace_bytes(b, t, max_query_len, hq, hkv, d, si, sl, last, int(mode), kcode)
        ws = _workspace(device, need)
        st = lib.tmx_chunked_prefill_attn(
            out.data_ptr(), q.data_ptr(), cu_q.data_ptr(), lens.data_ptr(), b, t, max_query_len, hq, hkv, d,
            q.stride(0), out.stride(0), kptr, vptr, kv_stride, kv_rows, cuk_ptr, bt_ptr, bt_stride, page,
            float(scale), si, sl, last, int(mode), None if lse is None else lse.data_ptr(),
            ws.data_ptr(), ws.numel(), kcode, _stream(device))
    _lib.check(st, "tmx_chunked_prefill_attn")
    return (out, lse) if return_lse else out


def _cache_strides(key_cache: torch.Tensor, value_cache: torch.Tensor):
    """(block, token, head) element strides of a [num_blocks, block_size, Hkv, D] cache view
    (NHD as FlashAttention allocates it, flash_attn.py:78-87, or an HND-permuted view)."""
    if key_cache.dim() != 4 or key_cache.shape != value_cache.shape or key_cache.stride() != value_cache.stride():
        raise _lib.TmxError("key_cache / value_cache must be [num_blocks, block_size, Hkv, D] views with equal "
                            "strides; got %r %r" % (tuple(key_cache.shape), tuple(value_cache.shape)))
    if key_cache.dtype != torch.bfloat16 or value_cache.dtype != torch.bfloat16:
        raise _lib.TmxError("only a bfloat16 KV cache (kv_cache_dtype='auto') is supported, got %s" % key_cache.dtype)
    if key_cache.stride(3) != 1:
        raise _lib.TmxError("KV cache head_dim must be contiguous")
    return key_cache.stride(0), key_cache.stride(1), key_cache.stride(2), key_cache.shape[1]


def _slots(slot_mapping: torch.Tensor) -> torch.Tensor:
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous():
        slot_mapping = slot_mapping.to(torch.int64).contiguous()
    return slot_mapping


def reshape_and_cache_flash(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                            value_cache: torch.Tensor, slot_mapping: torch.Tensor, kv_cache_dtype: str = "auto",
                            k_scale: Optional[torch.Tensor] = None, v_scale: Optional[torch.Tensor] = None) -> None:
    """Same signature as the reference's ops.reshape_and_cache_flash (vllm/_custom_ops.py:1605-1618):
    key/value [T, Hkv, D] -> cache rows at slot_mapping (len(slot_mapping) tokens, slot < 0 skipped).
    bf16 cache only; the scales are accepted and unused, as in the reference's 'auto' path."""
    device = _require_cuda(key.device)
    lib = _lib.load()
    if kv_cache_dtype not in ("auto", "bfloat16"):
        raise _lib.TmxError("kv_cache_dtype %r is not supported (bf16 cache only)" % (kv_cache_dtype,))
    if key.dtype != torch.bfloat16 or value.dtype != torch.bfloat16:
        raise _lib.TmxError("reshape_and_cache_flash needs bfloat16 key/value")
    if key.dim() != 3 or value.shape != key.shape or key.stride(2) != 1 or key.stride(1) != key.shape[2] \
            or value.stride(2) != 1 or value.stride(1) != value.shape[2]:
        raise _lib.TmxError("key/value must be [T, Hkv, D] with contiguous heads")
    blk, tok, head, block_size = _cache_strides(key_cache, value_cache)
    slots = _slots(slot_mapping)
    t = min(slots.numel(), key.shape[0])           # slot_mapping's length = actual tokens (flash_attn.py:493-497)
    with torch.cuda.device(device):
        st = lib.tmx_reshape_and_cache_flash(key.data_ptr(), value.data_ptr(), key_cache.data_ptr(),
                                             value_cache.data_ptr(), slots.data_ptr(), t, key.shape[1],
                                             key.shape[2], key.stride(0), value.stride(0), blk, tok, head,
                                             block_size, _stream(device))
    _lib.check(st, "tmx_reshape_and_cache_flash")


def qk_norm_rope(positions: torch.Tensor, q: torch.Tensor, k: torch.Tensor, q_weight: torch.Tensor,
                 k_weight: torch.Tensor, cos_sin_cache: torch.Tensor, eps: float, head_dim: int = 128,
                 inplace: bool = False, v: Optional[torch.Tensor] = None,
                 key_cache: Optional[torch.Tensor] = None, value_cache: Optional[torch.Tensor] = None,
                 slot_mapping: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Fused per-head q/k RMSNorm + NeoX RoPE (+ optional KV-cache write) -- one kernel for
    qwen3.py:166-174 (+ flash_attn.py:498-507).  q [T, Hq*D], k [T, Hkv*D] bf16 (row-contiguous);
    positions int64 [T]; cos_sin_cache [max_pos, D] bf16.  Returns (q, k) like rotary_emb does."""
    device = _require_cuda(q.device)
    lib = _lib.load()
    for name, x in (("q", q), ("k", k), ("q_weight", q_weight), ("k_weight", k_weight), ("cos_sin_cache", cos_sin_cache)):
        if x.dtype != torch.bfloat16:
            raise _lib.TmxError("qk_norm_rope needs bfloat16 %s, got %s" % (name, x.dtype))
    if q.dim() != 2 or k.dim() != 2 or q.shape[0] != k.shape[0] or q.shape[1] % head_dim or k.shape[1] % head_dim:
        raise _lib.TmxError("expected q [T, Hq*D], k [T, Hkv*D]; got %r %r" % (tuple(q.shape), tuple(k.shape)))
    if q.stride(1) != 1:
        q = q.contiguous()
    if k.stride(1) != 1:
        k = k.contiguous()
    t = q.shape[0]
    hq, hkv = q.shape[1] // head_dim, k.shape[1] // head_dim
    pos = positions.flatten()
    if pos.dtype != torch.int64 or not pos.is_contiguous():
        pos = pos.to(torch.int64).contiguous()
    if pos.numel() != t:
        raise _lib.TmxError("positions has %d entries for %d tokens" % (pos.numel(), t))
    cache = cos_sin_cache if cos_sin_cache.is_contiguous() else cos_sin_cache.contiguous()
    if cache.dim() != 2 or cache.shape[1] != head_dim:
        raise _lib.TmxError("cos_sin_cache must be [max_pos, head_dim] (rotary_dim == head_dim)")
    q_out = q if inplace else torch.empty((t, hq * head_dim), dtype=torch.bfloat16, device=device)
    k_out = k if inplace else torch.empty((t, hkv * head_dim), dtype=torch.bfloat16, device=device)
    cargs = (None, 0, None, None, None, 0, 0, 0, 0)
    if key_cache is not None:
        if v is None or value_cache is None or slot_mapping is None:
            raise _lib.TmxError("the fused cache write needs v, key_cache, value_cache and slot_mapping")
        if v.dtype != torch.bfloat16 or v.shape != k.shape:
            raise _lib.TmxError("v must be bf16 with k's shape")
        if v.stride(1) != 1:
            v = v.contiguous()
        blk, tok, head, block_size = _cache_strides(key_cache, value_cache)
        slots = _slots(slot_mapping)
        if slots.numel() < t:                     # padded q/k beyond the actual tokens: no slot -> skip
            slots = torch.cat([slots, slots.new_full((t - slots.numel(),), -1)])
        cargs = (v.data_ptr(), v.stride(0), key_cache.data_ptr(), value_cache.data_ptr(), slots.data_ptr(),
                 blk, tok, head, block_size)
    with torch.cuda.device(device):
        st = lib.tmx_qk_norm_rope(q_out.data_ptr(), k_out.data_ptr(), q.data_ptr(), k.data_ptr(),
                                  q_weight.data_ptr(), k_weight.data_ptr(), pos.data_ptr(), cache.data_ptr(),
                                  cache.shape[0], t, hq, hkv, head_dim, q.stride(0), k.stride(0),
                                  q_out.stride(0), k_out.stride(0), float(eps), *cargs, _stream(device))
    _lib.check(st, "tmx_qk_norm_rope")
    return q_out, k_out


def linear(x: torch.Tensor, weight: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y = x @ weight.T in bf16 (fp32 accumulation on tcgen05): the matmul of the reference's o_proj,
    RowParallelLinear.forward (vllm/model_executor/layers/linear.py:1197-1359), for this rank's shard.
    x [M, K], weight [N, K] (torch.nn.Linear layout), both bf16 with contiguous rows; N % 256 == 0, K % 64 == 0."""
    device = _require_cuda(x.device)
    lib = _lib.load()
    if x.dtype != torch.bfloat16 or weight.dtype != torch.bfloat16 or x.dim() != 2 or weight.dim() != 2 \
            or x.shape[1] != weight.shape[1]:
        raise _lib.TmxError("linear needs bf16 x [M, K] and weight [N, K]; got %r %r" % (tuple(x.shape), tuple(weight.shape)))
    if x.stride(1) != 1:
        x = x.contiguous()
    if weight.stride(1) != 1:
        weight = weight.contiguous()
    m, k = x.shape
    n = weight.shape[0]
    if out is None:
        out = torch.empty((m, n), dtype=torch.bfloat16, device=device)
    import ctypes
    ptrs = (ctypes.c_void_p * 1)(out.data_ptr())
    with torch.cuda.device(device):
        st = lib.tmx_linear_bf16(ptrs, 1, 0, 1, m, x.data_ptr(), weight.data_ptr(), m, n, k, x.stride(0),
                                 weight.stride(0), out.stride(0), _stream(device))
    _lib.check(st, "tmx_linear_bf16")
    return out


def linear_scatter(x: torch.Tensor, weight: torch.Tensor, dest_ptrs, slot: int, n_slots: int, rows_per_dest: int) -> None:
    """The same GEMM with the reduce-scatter fused into its epilogue: rows [d * rows_per_dest, ...) of the
    result are TMA-stored into dest_ptrs[d] (device pointers, possibly peer memory), a [n_slots, rows_per_dest, N]
    bf16 staging buffer, at slot `slot`.  See head_shard.RowParallelOProj."""
    device = _require_cuda(x.device)
    lib = _lib.load()
    import ctypes
    m, k = x.shape
    n = weight.shape[0]
    ptrs = (ctypes.c_void_p * len(dest_ptrs))(*[int(p) for p in dest_ptrs])
    with torch.cuda.device(device):
        st = lib.tmx_linear_bf16(ptrs, len(dest_ptrs), slot, n_slots, rows_per_dest, x.data_ptr(), weight.data_ptr(),
                                 m, n, k, x.stride(0), weight.stride(0), n, _stream(device))
    _lib.check(st, "tmx_linear_bf16")


def sum_partials(partials: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[S, R, N] bf16 -> [R, N] bf16, fp32 accumulation (the reduction half of the fused reduce-scatter)."""
    device = _require_cuda(partials.device)
    lib = _lib.load()
    s, r, n = partials.shape
    if out is None:
        out = torch.empty((r, n), dtype=torch.bfloat16, device=device)
    with torch.cuda.device(device):
        st = lib.tmx_sum_partials(out.data_ptr(), partials.data_ptr(), s, r, n, out.stride(0), _stream(device))
    _lib.check(st, "tmx_sum_partials")
    return out


def last_launch_count() -> int:
    return _lib.load().tmx_last_launch_count()
