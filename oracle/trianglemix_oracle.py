"""CPU oracle for the TriangleMix prefill hot path.  TEST INFRASTRUCTURE ONLY.

This file restates, on the CPU, the algorithm of the reference fork
(ackerm0n/vllm-trianglemix).  It is the checker the CUDA path is compared with.
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it.  Nothing under ``vllm_trianglemix_b200/``
imports it, and the product path raises when the CUDA library is missing.

Parity status: PINNED.  ``tests/golden/`` holds known answers produced by
importing the reference's own modules in the build container
(``tests/golden/make_golden.py``), and ``tests/test_oracle.py`` checks every
function below against them, plus the values the reference's own test script
prints (``tests/trianglemix_attention_test.py:18-220`` in the reference).

Reference lines each function follows (paths relative to the reference root):

* ``piece_loops``          vllm/attention/trianglemix_attention.py:82-89 (streaming),
                           :126-129 (last), :166-169 (middle)
* ``triangle01_loops``     :192-196 (streaming + last, clamp)
* ``dense01``              :218-219 (tril of ones)
* ``additive``             :254-258 (0 where attended, -inf elsewhere)
* ``should_use_triangle``  :263-278 and the first-K rewrite at :305-307
* ``helper_mask``          :328-340 (``None`` at seq_len <= 2048)
* ``attention_fp32``       vllm/attention/ascend_trianglemix.py:86-106
* ``rewrite_neg_inf``      vllm/attention/ascend_trianglemix.py:135-139

The ``*_closed`` functions are the vectorised closed form of the same
predicates; ``tests/test_oracle.py`` proves them equal to the loops.
"""
from __future__ import annotations

import math
from typing import Iterable, Optional, Sequence

import numpy as np

DENSE_CAUSAL = 0
TRIANGLE = 1

PIECES = ("streaming", "last", "middle", "triangle", "dense")


# --------------------------------------------------------------------------
# element predicates -- one per piece, written exactly as the reference's
# ``if`` conditions read (0-indexed query row i, key column j).
# --------------------------------------------------------------------------
def _in_streaming(i: int, j: int, n: int, si: int, sl: int, last: int) -> bool:
    if i >= j and j <= si:
        return True
    return i >= j and (i - j) <= sl


def _in_last(i: int, j: int, n: int, si: int, sl: int, last: int) -> bool:
    return i >= j and (n - i) < last and j > si and (i - j) > sl


def _in_middle(i: int, j: int, n: int, si: int, sl: int, last: int) -> bool:
    return i >= j and (n - i) >= last and j > si and (i - j) > sl


_ELEMENT = {"streaming": _in_streaming, "last": _in_last, "middle": _in_middle}


def piece_loops(piece: str, n: int, si: int, sl: int, last: int,
                rows: Optional[Iterable[int]] = None) -> np.ndarray:
    """The reference's O(N^2) scalar double loop for one piece (float32 0/1).

    ``rows`` restricts the outer loop to some query rows (used by the bounded
    CPU-baseline sample at 128K); the returned array then has len(rows) rows.
    """
    pred = _ELEMENT[piece]
    row_ids = range(n) if rows is None else list(rows)
    out = np.zeros((len(row_ids), n), dtype=np.float32)
    for r, i in enumerate(row_ids):
        for j in range(n):
            if pred(i, j, n, si, sl, last):
                out[r, j] = 1
    return out


def triangle01_loops(n: int, si: int, sl: int, last: int,
                     rows: Optional[Iterable[int]] = None) -> np.ndarray:
    rows = None if rows is None else list(rows)
    s = piece_loops("streaming", n, si, sl, last, rows)
    l = piece_loops("last", n, si, sl, last, rows)
    return np.clip(s + l, 0, 1)


def dense01(n: int, dtype=np.float32) -> np.ndarray:
    return np.tril(np.ones((n, n), dtype=dtype))


# --------------------------------------------------------------------------
# closed forms (vectorised; any N, evaluated for a row range)
# --------------------------------------------------------------------------
def piece_closed(piece: str, n: int, si: int, sl: int, last: int,
                 row0: int = 0, row1: Optional[int] = None) -> np.ndarray:
    """Boolean [row1-row0, n] array of one piece, rows row0..row1-1."""
    row1 = n if row1 is None else row1
    i = np.arange(row0, row1, dtype=np.int64)[:, None]
    j = np.arange(n, dtype=np.int64)[None, :]
    causal = j <= i
    stream = causal & ((j <= si) | (i - j <= sl))
    rest = causal & (j > si) & (i - j > sl)
    tail = (n - i) < last
    if piece == "streaming":
        return stream
    if piece == "last":
        return rest & tail
    if piece == "middle":
        return rest & ~tail
    if piece == "triangle":
        return causal & ((j <= si) | (i - j <= sl) | tail)
    if piece == "dense":
        return causal
    raise ValueError(piece)


def mask_closed(n: int, si: int, sl: int, last: int, mode: int,
                row0: int = 0, row1: Optional[int] = None) -> np.ndarray:
    return piece_closed("triangle" if mode == TRIANGLE else "dense",
                        n, si, sl, last, row0, row1)


def attended_pairs(n: int, si: int, sl: int, last: int, mode: int) -> int:
    """Number of attended (i, j) pairs, O(N) integer arithmetic."""
    if mode == DENSE_CAUSAL:
        return n * (n + 1) // 2
    total = 0
    for i in range(n):
        if (n - i) < last:
            total += i + 1
        else:
            lo = max(0, i - sl)
            sink_hi = min(i, si)
            total += (i - lo + 1) + max(0, min(sink_hi, lo - 1) + 1)
    return total


def pack_bits(mask: np.ndarray, row_stride_words: Optional[int] = None) -> np.ndarray:
    """Pack a boolean [R, N] mask LSB-first into uint32 words (bit b of word w
    of a row = mask[row, 32*w + b]); rows padded with zero words."""
    r, n = mask.shape
    words = (n + 31) // 32
    stride = words if row_stride_words is None else row_stride_words
    padded = np.zeros((r, stride * 32), dtype=np.uint8)
    padded[:, :n] = mask
    by = np.packbits(padded, axis=1, bitorder="little")
    return by.view("<u4").reshape(r, stride)


def additive(mask01: np.ndarray, dtype=np.float32) -> np.ndarray:
    return np.where(mask01 > 0, dtype(0.0), dtype(-np.inf)).astype(dtype)


def rewrite_neg_inf(mask: np.ndarray, value: float = -1e6) -> np.ndarray:
    if mask.size == 0:
        return mask
    return np.where(np.isinf(mask) & (mask < 0), np.asarray(value, dtype=mask.dtype), mask)


# --------------------------------------------------------------------------
# layer switch
# --------------------------------------------------------------------------
def should_use_triangle(layer_idx: int, num_triangle_layers: Optional[int],
                        triangle_layer_indices: Optional[Sequence[int]]) -> bool:
    if triangle_layer_indices:
        return layer_idx in triangle_layer_indices
    if num_triangle_layers is not None:
        return layer_idx < num_triangle_layers
    return False


def helper_mask(n: int, layer_idx: int, si: int, sl: int, last: int,
                num_triangle_layers: Optional[int],
                triangle_layer_indices: Optional[Sequence[int]],
                dtype=np.float32) -> Optional[np.ndarray]:
    """What TriangleMixAttention.get_attn_mask returns (None at n <= 2048)."""
    idx = list(triangle_layer_indices or [])
    if num_triangle_layers is not None and not idx:
        idx = list(range(num_triangle_layers))
    tri = should_use_triangle(layer_idx, num_triangle_layers, idx)
    if n <= 2048:
        return None
    m = mask_closed(n, si, sl, last, TRIANGLE if tri else DENSE_CAUSAL)
    return additive(m, dtype)


# --------------------------------------------------------------------------
# masked attention (the reference op, fp32, materialised scores)
# --------------------------------------------------------------------------
def attention_fp32(q, k, v, add_mask, scale: float):
    """softmax(q k^T * scale + add_mask) v, exactly the reference's sequence of
    torch ops, in float32.

    q: [B, S, Hq, D]; k, v: [B, S, Hkv, D] (torch CPU tensors, any float dtype,
    upcast to fp32).  GQA adapter: the reference op is MHA-only, so K/V heads are
    repeat_interleave'd to Hq first.  add_mask: None or [S, S] additive (0/-inf).
    Returns [B, S, Hq, D] float32, contiguous.
    """
    import torch
    q = q.float()
    k = k.float()
    v = v.float()
    hq, hkv = q.shape[2], k.shape[2]
    if hkv != hq:
        k = k.repeat_interleave(hq // hkv, dim=2)
        v = v.repeat_interleave(hq // hkv, dim=2)
    qt, kt, vt = q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)
    scores = torch.matmul(qt, kt.transpose(-2, -1)) * scale
    if add_mask is not None:
        scores = scores + add_mask
    w = torch.softmax(scores, dim=-1)
    out = torch.matmul(w, vt)
    return out.transpose(1, 2).contiguous()


def attention_rows_fp32(q, k, v, scale: float, n: int, si: int, sl: int, last: int,
                        mode: int, row0: int, row1: int):
    """Same arithmetic for query rows [row0, row1) of ONE sequence of length n.

    q: [S, Hq, D]; k, v: [S, Hkv, D].  Used at lengths where the full (H,S,S)
    score tensor does not fit (>= 16K): the mask rows come from the closed form,
    the scores of the chosen rows against ALL n keys are materialised, as the
    reference would.  Returns [row1-row0, Hq, D] float32.
    """
    import torch
    m = torch.from_numpy(additive(mask_closed(n, si, sl, last, mode, row0, row1)))
    qf = q[row0:row1].float().transpose(0, 1)            # [Hq, R, D]
    hq, hkv = q.shape[1], k.shape[1]
    kf = k[:n].float().repeat_interleave(hq // hkv, dim=1).transpose(0, 1)
    vf = v[:n].float().repeat_interleave(hq // hkv, dim=1).transpose(0, 1)
    scores = torch.matmul(qf, kf.transpose(-2, -1)) * scale + m
    w = torch.softmax(scores, dim=-1)
    return torch.matmul(w, vf).transpose(0, 1).contiguous()


def attention_varlen_fp32(q, k, v, cu_seqlens, scale: float, si: int, sl: int,
                          last: int, mode: int, max_full: int = 4096):
    """Varlen adapter: loop over sequences, one reference-op call per sequence
    (the reference has no varlen).  q: [T, Hq, D].  Returns [T, Hq, D] fp32."""
    import torch
    out = torch.empty(q.shape, dtype=torch.float32)
    for b in range(len(cu_seqlens) - 1):
        s, e = int(cu_seqlens[b]), int(cu_seqlens[b + 1])
        n = e - s
        if n == 0:
            continue
        for r0 in range(0, n, max_full):
            r1 = min(n, r0 + max_full)
            out[s + r0:s + r1] = attention_rows_fp32(
                q[s:e], k[s:e], v[s:e], scale, n, si, sl, last, mode, r0, r1)
    return out


def error_report(out, ref, atol: float = 1.5e-2, rtol: float = 1e-2) -> dict:
    """max-abs, Frobenius-relative error and tolerance-violation count."""
    import torch
    o, r = out.float(), ref.float()
    diff = (o - r).abs()
    viol = int((diff > atol + rtol * r.abs()).sum())
    return {"max_abs": float(diff.max()) if diff.numel() else 0.0,
            "fro_rel": float(torch.linalg.vector_norm(o - r) /
                             max(float(torch.linalg.vector_norm(r)), 1e-30)),
            "violations": viol, "numel": o.numel()}


# --------------------------------------------------------------------------
# Middle Q-K gradient (what the reference's TriangleMixAnalyzer consumes, trianglemix_attention.py:358-373;
# the reference has no producer -- this is the definition the GPU producer is checked against)
# --------------------------------------------------------------------------
def middle_gate_gradient_autograd(q, k, v, d_out, scale: float, si: int, sl: int, last: int):
    """dL/dg per (row, head) by autograd, for ONE sequence: the reference op's arithmetic
    (ascend_trianglemix.py:86-106, fp32) with the additive mask  0 on the triangle piece, log(g) on the
    middle piece (_create_middle_mask, :133-171), -inf elsewhere, L = sum(out * d_out), at g = 1.
    One gate per (row, head) so that the per-row gradients come out of a single backward.
    q, d_out [N, Hq, D]; k, v [N, Hkv, D].  Returns float32 [N, Hq]."""
    import torch
    n, hq, _ = q.shape
    hkv = k.shape[1]
    tri = torch.from_numpy(piece_closed("triangle", n, si, sl, last).astype(np.float32))
    mid = torch.from_numpy(piece_closed("middle", n, si, sl, last).astype(np.float32))
    g = torch.ones(hq, n, 1, dtype=torch.float32, requires_grad=True)
    qf = q.float().transpose(0, 1)
    kf = k.float().repeat_interleave(hq // hkv, dim=1).transpose(0, 1)
    vf = v.float().repeat_interleave(hq // hkv, dim=1).transpose(0, 1)
    scores = torch.matmul(qf, kf.transpose(-2, -1)) * scale
    add = torch.where(tri > 0, torch.zeros(()), torch.full((), -float("inf"))) .unsqueeze(0) .expand(hq, n, n)
    add = torch.where(mid.unsqueeze(0) > 0, torch.log(g).expand(hq, n, n), add)
    w = torch.softmax(scores + add, dim=-1)
    out = torch.matmul(w, vf).transpose(0, 1)
    (out * d_out.float()).sum().backward()
    return g.grad[:, :, 0].transpose(0, 1).contiguous()
