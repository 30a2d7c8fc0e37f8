"""CUDA-graph capture of the attention path.

vLLM replays its model step from CUDA graphs (piecewise graphs around `unified_attention_with_output`,
vllm/attention/layer.py:598-645; full graphs for decode).  A library call is capturable only if it allocates
nothing, never synchronises and reads no host state that changes between capture and replay: include/tmx.h
promises exactly that.  Here a three-layer step (dense, triangle, dense) is captured once -- through the registered
torch ops with a caller workspace, and through the allocation-free C entry point tmx_prefill_attn_static -- then
replayed on NEW inputs written into the captured buffers, and must reproduce the eager results bit for bit."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SCALE = 128 ** -0.5
SI, SL, LAST = 4, 32, 64
MODES = (0, 1, 0)          # dense-causal, triangle, dense-causal


def _fill(q, k, v, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    for t in (q, k, v):
        t.copy_(torch.randn(t.shape, generator=g, device=DEV, dtype=torch.float32))


def _setup():
    lens = [2304, 700, 129]
    t, hq, hkv = sum(lens), 8, 2
    q = torch.empty(t, hq, 128, dtype=torch.bfloat16, device=DEV)
    k = torch.empty(t, hkv, 128, dtype=torch.bfloat16, device=DEV)
    v = torch.empty(t, hkv, 128, dtype=torch.bfloat16, device=DEV)
    cu = torch.tensor([0] + torch.tensor(lens).cumsum(0).tolist(), dtype=torch.int32, device=DEV)
    outs = [torch.empty_like(q) for _ in MODES]
    return lens, q, k, v, cu, outs


def _eager(ops, lens, q, k, v, cu):
    return [ops.prefill_attn(q, k, v, cu, max(lens), SCALE, SI, SL, LAST, m) for m in MODES]


def _capture(step):
    step()                                   # warm-up outside capture: lazy module load, kernel attributes
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    return graph


def test_registered_ops_capture_and_replay():
    from vllm_trianglemix_b200 import ops, torch_ops
    t = torch_ops.load()
    lens, q, k, v, cu, outs = _setup()
    b, tt, mx, hq, hkv = len(lens), sum(lens), max(lens), q.shape[1], k.shape[1]
    ws = [torch.empty(max(256, t.attn_workspace_bytes(b, tt, mx, hq, hkv, 128, SI, SL, LAST, m, 0)),
                      dtype=torch.uint8, device=DEV) for m in MODES]

    def step():
        for o, w, m in zip(outs, ws, MODES):
            t.prefill_attn(o, q, k, v, cu, mx, SCALE, SI, SL, LAST, m, None, None, w, 0)

    _fill(q, k, v, 1)
    graph = _capture(step)
    for seed in (2, 3):
        _fill(q, k, v, seed)                 # new inputs in the captured buffers
        for o in outs:
            o.fill_(float("nan"))
        graph.replay()
        torch.cuda.synchronize()
        for got, want in zip(outs, _eager(ops, lens, q, k, v, cu)):
            assert torch.equal(got.view(torch.int16), want.view(torch.int16)), seed


def test_static_entry_point_capture_and_replay():
    """The entry point made for callers that cannot allocate: captured, its kernels keep reading the static
    device array (block index, split-KV partials), which every replay rebuilds."""
    from vllm_trianglemix_b200 import _lib, ops
    lib = _lib.load()
    lens, q, k, v, cu, outs = _setup()
    b, tt, mx, hq, hkv = len(lens), sum(lens), max(lens), q.shape[1], k.shape[1]

    def step():
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        for o, m in zip(outs, MODES):
            st = lib.tmx_prefill_attn_static(
                ctypes.c_void_p(o.data_ptr()), ctypes.c_void_p(q.data_ptr()), ctypes.c_void_p(k.data_ptr()),
                ctypes.c_void_p(v.data_ptr()), ctypes.c_void_p(cu.data_ptr()), b, tt, mx, hq, hkv, 128,
                hq * 128, hkv * 128, hq * 128, SCALE, SI, SL, LAST, m, None, 0, None, 0, stream)
            assert st == 0, lib.tmx_strerror(st).decode()

    _fill(q, k, v, 4)
    graph = _capture(step)
    for seed in (5, 6):
        _fill(q, k, v, seed)
        for o in outs:
            o.fill_(float("nan"))
        graph.replay()
        torch.cuda.synchronize()
        for got, want in zip(outs, _eager(ops, lens, q, k, v, cu)):
            assert torch.equal(got.view(torch.int16), want.view(torch.int16)), seed
