"""The allocation-free entry point tmx_prefill_attn_static (include/tmx.h; no counterpart in the reference):
called through the C ABI with ctypes, its output must equal the existing entry
point's (tmx_prefill_attn with a caller workspace) within the tolerance of the parity tests -- on two calls in a row,
because every call shares ONE static scratch array that keeps the previous call's contents.

Named test_zz_* so that pytest collects it last.  Self-contained: runs alone in a
fresh process and after the other GPU tests.
"""
import ctypes

import pytest
import torch

from oracle import trianglemix_oracle as orc

pytestmark = pytest.mark.gpu
ATOL, RTOL = 1.5e-2, 1e-2          # tests/test_gpu_parity.py
DEV = "cuda:0"
STATIC_BYTES = 16 << 20            # TMX_STATIC_SCRATCH_BYTES
ERR_STATIC_SCRATCH = 8


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _inputs(lens, hq, hkv, seed):
    t = sum(lens)
    g = torch.Generator().manual_seed(seed)
    q = torch.randn(t, hq, 128, generator=g).bfloat16().to(DEV)
    k = torch.randn(t, hkv, 128, generator=g).bfloat16().to(DEV)
    v = torch.randn(t, hkv, 128, generator=g).bfloat16().to(DEV)
    cu = torch.tensor([0] + torch.tensor(lens).cumsum(0).tolist(), dtype=torch.int32, device=DEV)
    return q, k, v, cu


def _both(lib, lens, hq, hkv, si, sl, last, mode, kernel, seed):
    """(status, output, lse) of the static entry point and of the existing one on the same inputs."""
    q, k, v, cu = _inputs(lens, hq, hkv, seed)
    t, b, mx = sum(lens), len(lens), max(lens)
    scale = 128 ** -0.5
    need = lib.tmx_attn_workspace_bytes(b, t, mx, hq, hkv, 128, si, sl, last, mode, kernel)
    ws = torch.empty(max(need, 256), dtype=torch.uint8, device=DEV)
    outs = []
    for static in (True, False):
        out = torch.full_like(q, float("nan"))
        lse = torch.full((hq, t), float("nan"), dtype=torch.float32, device=DEV)
        head = (_ptr(out), _ptr(q), _ptr(k), _ptr(v), _ptr(cu), b, t, mx, hq, hkv, 128, hq * 128, hkv * 128, hq * 128,
                scale, si, sl, last, mode, None, 0, _ptr(lse))
        if static:
            st = lib.tmx_prefill_attn_static(*head, kernel, None)
        else:
            st = lib.tmx_prefill_attn(*head, _ptr(ws), ws.numel(), kernel, None)
        torch.cuda.synchronize()
        outs.append((st, out, lse))
    return need, outs


# two calls in a row per case list: the second call finds the first one's block index and split-KV partials in the
# shared array (different batch shape, mask mode and kernel family on purpose)
CASES = [
    # lens, hq, hkv, si, sl, last, mode, kernel
    ([1500, 333, 2304], 8, 2, 4, 32, 64, 1, 0),     # AUTO triangle: streaming kernel + split-KV tail (partials in scratch)
    ([700, 129], 8, 2, 4, 32, 64, 0, 0),            # AUTO dense-causal on CTA pairs: block index only
    ([2304], 4, 2, 20, 150, 300, 1, 1),             # mma.sync kernel, off-default triangle
    ([900, 64, 1], 4, 1, 4, 32, 64, 1, 2),          # single-CTA tcgen05 kernel
]


def test_static_scratch_matches_the_workspace_entry_point_twice_in_a_row():
    from vllm_trianglemix_b200 import _lib
    lib = _lib.load()
    for rnd in range(2):                            # the whole list twice: also the same shape after other shapes
        for n, (lens, hq, hkv, si, sl, last, mode, kernel) in enumerate(CASES):
            need, ((st_s, out_s, lse_s), (st_w, out_w, lse_w)) = _both(lib, lens, hq, hkv, si, sl, last, mode, kernel,
                                                                      seed=10 * rnd + n)
            assert need <= STATIC_BYTES
            assert st_w == 0, lib.tmx_strerror(st_w).decode()
            assert st_s == 0, lib.tmx_strerror(st_s).decode()
            assert not torch.isnan(out_s.float()).any()
            rep = orc.error_report(out_s.cpu(), out_w.float().cpu(), ATOL, RTOL)
            assert rep["violations"] == 0, (rnd, n, rep)
            assert float((lse_s - lse_w).abs().max()) <= 2e-2, (rnd, n)


def test_static_scratch_rejects_requests_beyond_its_size():
    from vllm_trianglemix_b200 import _lib
    lib = _lib.load()
    hq, hkv, n = 32, 8, 32768                       # AUTO triangle at 32K: 32.5 MB of split-KV partials
    need = lib.tmx_attn_workspace_bytes(1, n, n, hq, hkv, 128, 4, 32, 64, 1, 0)
    assert need > STATIC_BYTES
    q = torch.zeros(n, hq, 128, dtype=torch.bfloat16, device=DEV)
    k = torch.zeros(n, hkv, 128, dtype=torch.bfloat16, device=DEV)
    cu = torch.tensor([0, n], dtype=torch.int32, device=DEV)
    out = torch.full_like(q, 7.0)
    st = lib.tmx_prefill_attn_static(_ptr(out), _ptr(q), _ptr(k), _ptr(k), _ptr(cu), 1, n, n, hq, hkv, 128, hq * 128,
                                     hkv * 128, hq * 128, 128 ** -0.5, 4, 32, 64, 1, None, 0, None, 0, None)
    torch.cuda.synchronize()
    assert st == ERR_STATIC_SCRATCH
    assert "TMX_STATIC_SCRATCH_BYTES" in lib.tmx_strerror(st).decode()
    assert lib.tmx_last_launch_count() == 0
    assert bool((out == 7.0).all())                 # nothing was launched
    # the same layer on the block-index tcgen05 kernel needs only the block index: served
    st = lib.tmx_prefill_attn_static(_ptr(out), _ptr(q), _ptr(k), _ptr(k), _ptr(cu), 1, n, n, hq, hkv, 128, hq * 128,
                                     hkv * 128, hq * 128, 128 ** -0.5, 4, 32, 64, 1, None, 0, None, 3, None)
    torch.cuda.synchronize()
    assert st == 0, lib.tmx_strerror(st).decode()
    assert bool((out == 0.0).all())                 # V = 0 -> O = 0
