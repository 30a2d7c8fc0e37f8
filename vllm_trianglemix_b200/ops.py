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
KERNELS = {"auto": 0, "mma": 1, "tcgen05": 2}
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


def last_launch_count() -> int:
    return _lib.load().tmx_last_launch_count()
