"""One process driving TWO GPUs (needs >= 2 visible devices; skipped otherwise).

include/tmx.h promises "re-entrant across devices": kernel attributes (the > 48 KB dynamic shared-memory
opt-in of every tcgen05 kernel), the SM count, the arch check and the static scratch address of
tmx_prefill_attn_static are all PER DEVICE.  Round 1 cached them per process (ADVICE: a launch on cuda:1 after
cuda:0 returned cudaErrorInvalidValue); this test runs every kernel family on cuda:0, then cuda:1, then cuda:0
again in one process and requires bit-identical outputs (same inputs, deterministic kernels, same GPU model)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
SCALE = 128 ** -0.5


def _inputs(lens, hq, hkv, seed):
    t = sum(lens)
    g = torch.Generator().manual_seed(seed)
    q = torch.randn(t, hq, 128, generator=g).bfloat16()
    k = torch.randn(t, hkv, 128, generator=g).bfloat16()
    v = torch.randn(t, hkv, 128, generator=g).bfloat16()
    cu = torch.tensor([0] + torch.tensor(lens).cumsum(0).tolist(), dtype=torch.int32)
    return q, k, v, cu


def _run_all(dev):
    """Every kernel family once on `dev`; returns CPU copies of the results."""
    from vllm_trianglemix_b200 import _lib, ops
    res = {}
    lens = [2304, 700, 129]
    q, k, v, cu = (x.to(dev) for x in _inputs(lens, 8, 2, seed=7))
    # AUTO triangle (streaming kernel + split-KV tail + merge), AUTO dense (CTA pairs), single-CTA tcgen05, mma.sync
    for name, mode, kernel in (("tri_auto", 1, "auto"), ("dense_auto", 0, "auto"), ("dense_v3", 0, "tcgen05"),
                               ("tri_v3", 1, "tcgen05"), ("tri_mma", 1, "mma")):
        res[name] = ops.prefill_attn(q, k, v, cu, max(lens), SCALE, 4, 32, 64, mode, kernel=kernel).cpu()
    # off-default triangle (block-index pass on tcgen05 + tail)
    res["tri_offdefault"] = ops.prefill_attn(q, k, v, cu, max(lens), SCALE, 20, 150, 300, 1).cpu()
    # masks
    res["bits"] = ops.mask_bits(3000, 4, 32, 64, "triangle", dev).cpu()
    res["additive"] = ops.mask_additive(777, 4, 32, 64, 1, dev, torch.bfloat16).cpu()
    # prologue: q/k RMSNorm + RoPE + KV-cache write
    import vllm_trianglemix_b200 as tmx
    t, hq, hkv = 1000, 8, 2
    g = torch.Generator().manual_seed(11)
    qkv = torch.randn(t, (hq + 2 * hkv) * 128, generator=g).bfloat16().to(dev)
    pq, pk, pv = qkv.split([hq * 128, hkv * 128, hkv * 128], dim=-1)
    w = (1 + 0.1 * torch.randn(128, generator=g)).bfloat16().to(dev)
    cs = tmx.compute_cos_sin_cache(128, 2048, 1000000.0, device=dev)
    kvc = torch.zeros(2, 64, 16, hkv, 128, dtype=torch.bfloat16, device=dev)
    ops.qk_norm_rope(torch.arange(t, device=dev), pq, pk, w, w, cs, 1e-6, 128, inplace=True, v=pv,
                     key_cache=kvc[0], value_cache=kvc[1], slot_mapping=torch.arange(t, device=dev))
    res["prologue_qkv"] = qkv.cpu()
    res["prologue_cache"] = kvc.cpu()
    # o_proj GEMM
    x = torch.randn(512, 1024, generator=g).bfloat16().to(dev)
    wt = torch.randn(768, 1024, generator=g).bfloat16().to(dev)
    res["linear"] = ops.linear(x, wt).cpu()
    # the allocation-free entry point: its scratch address is taken per device
    lib = _lib.load()
    out = torch.full_like(q, float("nan"))
    with torch.cuda.device(dev):
        st = lib.tmx_prefill_attn_static(
            ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(q.data_ptr()), ctypes.c_void_p(k.data_ptr()),
            ctypes.c_void_p(v.data_ptr()), ctypes.c_void_p(cu.data_ptr()), len(lens), sum(lens), max(lens), 8, 2, 128,
            8 * 128, 2 * 128, 8 * 128, SCALE, 4, 32, 64, 1, None, 0, None, 0,
            ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    assert st == 0, lib.tmx_strerror(st).decode()
    torch.cuda.synchronize(dev)
    res["static"] = out.cpu()
    return res


def test_one_process_two_devices_bit_identical():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs")
    first = _run_all(torch.device("cuda:0"))
    second = _run_all(torch.device("cuda:1"))
    third = _run_all(torch.device("cuda:0"))
    for name, want in first.items():
        for which, got in (("cuda:1", second[name]), ("cuda:0 again", third[name])):
            same = torch.equal(got.view(torch.int16) if got.dtype == torch.bfloat16 else got,
                               want.view(torch.int16) if want.dtype == torch.bfloat16 else want)
            assert same, (name, which)
    assert not torch.isnan(first["static"].float()).any()
    assert torch.equal(first["static"].view(torch.int16), first["tri_auto"].view(torch.int16))
