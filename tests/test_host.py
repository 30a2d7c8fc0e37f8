"""CPU-only tests of the host side: the C-ABI library loads and exports every symbol
include/tmx.h declares, the product path fails loudly without a GPU (no CPU fallback),
the reference-facing classes keep the reference's signatures/behaviour, and the
head-sharded N>1 path (partition + output all-gather) is exact under gloo, world_size 2.
No compute entry point of libtmx.so is called here."""
import ctypes
import inspect
import os
import re
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import vllm_trianglemix_b200 as tmx                      # noqa: E402
from vllm_trianglemix_b200 import _lib, head_shard, ops   # noqa: E402
from oracle import trianglemix_oracle as orc              # noqa: E402


# ------------------------------------------------------------------ C ABI
def _declared_symbols():
    with open(os.path.join(ROOT, "include", "tmx.h")) as f:
        src = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(tmx_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_abi():
    syms = _declared_symbols()
    for must in ("tmx_mask_bits", "tmx_mask_dense01", "tmx_mask_additive", "tmx_block_index",
                 "tmx_prefill_attn", "tmx_attn_workspace_bytes", "tmx_strerror"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declared_symbols():
        assert hasattr(raw, name), "libtmx.so does not export %s" % name


def test_ctypes_table_matches_header():
    assert sorted(_lib.SIGNATURES) == _declared_symbols()


def test_torch_library_ops_are_registered():
    """The TORCH_LIBRARY TU loads without a GPU and registers every op with a mutating schema."""
    from vllm_trianglemix_b200 import torch_ops
    t = torch_ops.load()
    for name in ("prefill_attn", "chunked_prefill_attn", "qk_norm_rope", "reshape_and_cache_flash", "mask_bits",
                 "mask_additive", "attn_workspace_bytes"):
        assert hasattr(t, name), name
    schema = str(torch.ops.trianglemix.prefill_attn.default._schema)
    assert "Tensor(a!) out" in schema and "Tensor? bitmask" in schema
    # host-only op: same number as the C ABI
    lib = _lib.load()
    assert t.attn_workspace_bytes(1, 4096, 4096, 32, 8, 128, 4, 32, 64, 1, 0) == \
        lib.tmx_attn_workspace_bytes(1, 4096, 4096, 32, 8, 128, 4, 32, 64, 1, 0)
    # CPU tensors are refused by the op itself
    x = torch.zeros(4, 2, 128, dtype=torch.bfloat16)
    with pytest.raises((RuntimeError, NotImplementedError)):
        t.prefill_attn(x, x, x, x, torch.tensor([0, 4], dtype=torch.int32), 4, 0.1, 4, 32, 64, 1, None, None,
                       torch.zeros(256, dtype=torch.uint8), 0)


def test_status_strings_and_version():
    lib = _lib.load()
    assert lib.tmx_version() >= 100
    seen = set()
    for code in range(0, 9):
        s = lib.tmx_strerror(code).decode()
        assert s and s not in seen
        seen.add(s)
    assert lib.tmx_last_launch_count() == 0          # nothing was launched in this process


def test_workspace_bytes_is_host_only_and_monotonic():
    lib = _lib.load()
    small = lib.tmx_attn_workspace_bytes(1, 4096, 4096, 32, 8, 128, 4, 32, 64, 1, 0)
    big = lib.tmx_attn_workspace_bytes(1, 131072, 131072, 32, 8, 128, 4, 32, 64, 1, 0)
    assert 0 < small <= big < (1 << 32)


# ------------------------------------------------------------------ no CPU fallback
def test_cpu_tensors_are_rejected_loudly():
    q = torch.zeros(64, 4, 128, dtype=torch.bfloat16)
    cu = torch.tensor([0, 64], dtype=torch.int32)
    with pytest.raises(tmx.TmxError, match="no CPU path|CUDA"):
        ops.prefill_attn(q, q[:, :2], q[:, :2], cu, 64, 0.1, 4, 32, 64, ops.TRIANGLE)
    with pytest.raises(tmx.TmxError):
        ops.mask_bits(64, 4, 32, 64, "triangle", device="cpu")
    gen = tmx.TriangleMixMaskGenerator(tmx.TriangleMixConfig())
    with pytest.raises(tmx.TmxError):
        gen.get_attention_mask(64, 0, torch.device("cpu"), torch.float32, use_triangle=True)


def test_missing_library_raises(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", os.path.join(ROOT, "does_not_exist", "libtmx.so"))
    with pytest.raises(tmx.TmxError, match="no CPU fallback"):
        _lib.load()


def test_product_package_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "vllm_trianglemix_b200")
    for name in os.listdir(pkg):
        if name.endswith(".py"):
            with open(os.path.join(pkg, name)) as f:
                src = f.read()
            assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), name


# ------------------------------------------------------------------ reference-facing API
def test_config_defaults_and_switch():
    c = tmx.TriangleMixConfig()
    assert (c.num_sink_tokens, c.sliding_window_size, c.num_last_tokens) == (4, 32, 64)
    assert c.num_triangle_layers is None and c.triangle_layer_indices == []
    gen = tmx.TriangleMixMaskGenerator(tmx.TriangleMixConfig(triangle_layer_indices=[2, 5]))
    assert [gen.should_use_triangle(i) for i in range(6)] == [False, False, True, False, False, True]
    # first-K default mutates the config (reference trianglemix_attention.py:305-307)
    cfg = tmx.TriangleMixConfig(num_triangle_layers=4)
    att = tmx.TriangleMixAttention(cfg, 8)
    assert cfg.triangle_layer_indices == [0, 1, 2, 3]
    assert [i for i in range(8) if att.mask_generator.should_use_triangle(i)] == [0, 1, 2, 3]
    # <= 2048 -> None without touching the device (reference :330-332)
    assert att.get_attn_mask(2048, 0, torch.device("cpu")) is None
    for i in range(8):
        want = orc.should_use_triangle(i, 4, [0, 1, 2, 3])
        assert att.mask_generator.should_use_triangle(i) == want


def test_inference_config_presets():
    f = tmx.TriangleMixInferenceConfig.for_qwen3
    assert f("8B").num_triangle_layers == 8 and f("14B").num_triangle_layers == 10
    assert f("32B").num_triangle_layers == 12
    assert tmx.TriangleMixInferenceConfig().to_trianglemix_config() is None     # disabled by default
    cfg = f("8B").to_trianglemix_config()
    assert isinstance(cfg, tmx.TriangleMixConfig)


def test_hf_adaptor():
    class HF:
        sliding_window = 128
        num_hidden_layers = 36
    assert tmx.create_trianglemix_config_from_hf(HF(), use_trianglemix=False) is None
    cfg = tmx.create_trianglemix_config_from_hf(HF(), use_trianglemix=True, num_triangle_layers=8)
    assert cfg.sliding_window_size == 128 and cfg.num_last_tokens == 256 and cfg.num_triangle_layers == 8


def test_analyzer_known_answer():
    # the reference's test script prints [4, 5, 6, 7] for this input (tests/trianglemix_attention_test.py)
    an = tmx.TriangleMixAnalyzer(8)
    for layer in range(8):
        an.record_middle_qk_gradient(layer, torch.full((4,), 1.0 / (layer + 1)))
    assert an.get_triangle_layers(4) == [4, 5, 6, 7]
    an.reset()


def test_backend_signature_matches_vllm_slot():
    b = tmx.TriangleMixBackend
    assert b.accept_output_buffer is True
    assert b.get_impl_cls() is tmx.TriangleMixImpl
    assert b.get_metadata_cls() is tmx.TriangleMixMetadata
    assert b.get_builder_cls() is tmx.TriangleMixMetadataBuilder
    assert b.get_kv_cache_shape(10, 16, 8, 128) == (2, 10, 16, 8, 128)
    with pytest.raises(ValueError):
        b.get_kv_cache_shape(10, 7, 8, 128)
    init = list(inspect.signature(tmx.TriangleMixImpl.__init__).parameters)
    assert init[1:11] == ["num_heads", "head_size", "scale", "num_kv_heads", "alibi_slopes", "sliding_window",
                          "kv_cache_dtype", "logits_soft_cap", "attn_type", "kv_sharing_target_layer_name"]
    fwd = list(inspect.signature(tmx.TriangleMixImpl.forward).parameters)
    assert fwd == ["self", "layer", "query", "key", "value", "kv_cache", "attn_metadata", "output",
                   "output_scale", "output_block_scale"]


def test_backend_forward_contract_without_gpu():
    impl = tmx.TriangleMixImpl(8, 128, 0.1, num_kv_heads=2)
    x = torch.zeros(4, 8 * 128, dtype=torch.bfloat16)
    with pytest.raises(AssertionError):
        impl.forward(None, x, x, x, None, None, output=None)
    with pytest.raises(NotImplementedError):
        impl.forward(None, x, x, x, None, None, output=x, output_scale=torch.ones(1))
    assert impl.forward(None, x, x, x, None, None, output=x) is x            # profiling run: untouched
    with pytest.raises(NotImplementedError):
        tmx.TriangleMixImpl(8, 128, 0.1, alibi_slopes=[1.0])
    # dense when no TriangleMix helper; per-layer switch otherwise
    assert impl.mask_mode()[0] == ops.DENSE_CAUSAL
    cfg = tmx.TriangleMixConfig(triangle_layer_indices=[1])
    modes = [tmx.TriangleMixImpl(8, 128, 0.1, 2, trianglemix_config=cfg, layer_idx=i, num_layers=2).mask_mode()
             for i in range(2)]
    assert modes[0][0] == ops.DENSE_CAUSAL and modes[1] == (ops.TRIANGLE, 4, 32, 64)


# ------------------------------------------------------------------ the step in front of attention
def test_qwen3_block_heads_per_rank_and_no_cpu_path():
    """Heads per tensor-parallel rank as reference qwen3.py:85-98; CPU tensors are refused."""
    for tp, want in ((1, (32, 8)), (2, (16, 4)), (8, (4, 1)), (16, (2, 1))):
        blk = tmx.Qwen3TriangleMixAttention(32, 8, 128, max_position=64, tp_size=tp)
        assert (blk.num_heads, blk.num_kv_heads) == want
        assert blk.q_size == want[0] * 128 and blk.kv_size == want[1] * 128 and blk.scaling == 128 ** -0.5
    with pytest.raises(AssertionError):
        tmx.Qwen3TriangleMixAttention(32, 8, tp_size=3)
    blk = tmx.Qwen3TriangleMixAttention(4, 2, 128, max_position=64, rope_theta=10000.0)
    assert blk.cos_sin_cache.shape == (64, 128) and blk.cos_sin_cache.dtype == torch.bfloat16
    assert torch.equal(blk.cos_sin_cache[0, :64], torch.ones(64, dtype=torch.bfloat16))      # cos(0) | sin(0)
    assert not blk.cos_sin_cache[0, 64:].any()
    qkv = torch.zeros(3, (4 + 2 * 2) * 128, dtype=torch.bfloat16)
    meta = tmx.TriangleMixMetadataBuilder.from_seq_lens([3], "cpu")
    with pytest.raises(_lib.TmxError):
        blk(torch.arange(3), qkv, None, meta)


def test_cache_write_signature_matches_reference_op():
    """ops.reshape_and_cache_flash keeps the reference's argument list (vllm/_custom_ops.py:1605-1614)."""
    names = list(inspect.signature(ops.reshape_and_cache_flash).parameters)
    assert names == ["key", "value", "key_cache", "value_cache", "slot_mapping", "kv_cache_dtype", "k_scale",
                     "v_scale"]
    k = torch.zeros(2, 1, 128, dtype=torch.bfloat16)
    kc = torch.zeros(1, 16, 1, 128, dtype=torch.bfloat16)
    with pytest.raises(_lib.TmxError):          # CPU tensors: no fallback
        ops.reshape_and_cache_flash(k, k, kc, kc, torch.zeros(2, dtype=torch.int64), "auto", None, None)
    # metadata carries slot_mapping through the builder (flash_attn.py:127)
    import types
    m = types.SimpleNamespace(num_actual_tokens=2, max_query_len=2, query_start_loc=torch.tensor([0, 2]),
                              max_seq_len=2, seq_lens=None, causal=True, slot_mapping=torch.tensor([5, 6]))
    md = tmx.TriangleMixMetadataBuilder().build(common_attn_metadata=m)
    assert md.slot_mapping.tolist() == [5, 6]


# ------------------------------------------------------------------ head partition
@pytest.mark.parametrize("hq,hkv", [(32, 8), (64, 8), (16, 8)])
@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_head_partition_covers_all_heads(hq, hkv, world):
    parts = [head_shard.head_partition(r, world, hq, hkv) for r in range(world)]
    q_owned = sorted(h for p in parts for h in range(p.q_head_start, p.q_head_start + p.num_heads))
    assert q_owned == list(range(hq))
    kv_owned = sorted(h for p in parts for h in range(p.kv_head_start, p.kv_head_start + p.num_kv_heads))
    assert kv_owned == list(range(hkv))
    for p in parts:       # each rank's query heads map onto its own KV heads
        g = hq // hkv
        for h in range(p.q_head_start, p.q_head_start + p.num_heads):
            assert p.kv_head_start <= h // g < p.kv_head_start + p.num_kv_heads
        assert p.num_heads == hq // world and p.num_kv_heads == max(1, hkv // world)   # qwen3.py:85-98


def test_head_partition_kv_replication_and_errors():
    p = head_shard.head_partition(5, 16, 64, 8)          # world > kv heads: replicate
    assert (p.num_heads, p.num_kv_heads, p.kv_head_start, p.kv_replicas) == (4, 1, 2, 2)
    assert p.q_head_start // (64 // 8) == p.kv_head_start
    with pytest.raises(ValueError):
        head_shard.head_partition(0, 3, 32, 8)
    with pytest.raises(ValueError):
        head_shard.head_partition(2, 2, 32, 8)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _gloo_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)                              # same global tensors on every rank
        lens, hq, hkv, d = [96, 200], 8, 4, 128
        t = sum(lens)
        cu = [0, 96, 296]
        si, sl, last = 2, 8, 16
        q = torch.randn(t, hq, d).bfloat16()
        k = torch.randn(t, hkv, d).bfloat16()
        v = torch.randn(t, hkv, d).bfloat16()
        part = head_shard.head_partition(rank, world, hq, hkv)
        scale = d ** -0.5
        # the checker stands in for the device kernel: each rank computes ITS heads only
        mine = orc.attention_varlen_fp32(head_shard.shard_heads(q, part, "q"), head_shard.shard_heads(k, part, "kv"),
                                         head_shard.shard_heads(v, part, "kv"), cu, scale, si, sl, last, orc.TRIANGLE)
        mine = torch.as_tensor(np.asarray(mine)).to(torch.bfloat16)
        gathered, _ = head_shard.all_gather_heads(mine)
        full = head_shard.heads_from_gathered(gathered)
        ref = torch.as_tensor(np.asarray(
            orc.attention_varlen_fp32(q, k, v, cu, scale, si, sl, last, orc.TRIANGLE))).to(torch.bfloat16)
        ok = torch.equal(full, ref) and gathered.shape == (world, t, hq // world, d)
        # the double-buffered asynchronous gather bench.py uses: 5 "layers", each result must be that layer's
        og = head_shard.OverlappedGather((4, 3), torch.float32, "cpu", depth=2)
        for i in range(5):
            og.buffer(i).fill_(100.0 * i + rank)
            og.launch(i)
            if i >= 1:          # layer i-1's gather is complete and intact while layer i's is in flight
                prev = og.result(i - 1)
                ok = ok and all(bool((prev[r] == 100.0 * (i - 1) + r).all()) for r in range(world))
        og.finish()
        ok = ok and all(w is None for w in og.pending)
        ok = ok and all(bool((og.result(4)[r] == 400.0 + r).all()) for r in range(world))
        # max-over-ranks timing reduction used by bench.py
        ms = torch.tensor([10.0 + rank])
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ok = ok and float(ms) == 10.0 + world - 1
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_head_sharded_path_gloo_world2():
    world = 2
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
        assert p.exitcode == 0
    assert dict(ret) == {0: True, 1: True}


# ------------------------------------------------------------------ backend host logic (no kernel runs: ops are recorded)
class _Recorder:
    def __init__(self):
        self.calls = []

    def prefill_attn(self, q, k, v, cu, max_seqlen, scale, si, sl, last, mode, out=None, kernel="auto", **kw):
        self.calls.append(("prefill", q.shape[0], cu.tolist(), max_seqlen, mode))
        return out

    def chunked_prefill_attn(self, q, cu, seq_lens, max_query_len, scale, si, sl, last, mode, **kw):
        self.calls.append(("chunked", q.shape[0], cu.tolist(), seq_lens.tolist(), max_query_len, mode,
                           tuple(kw["block_table"].shape)))
        return kw.get("out")


def _impl_with_recorder(monkeypatch, layer_idx=0):
    from vllm_trianglemix_b200 import backend
    rec = _Recorder()
    monkeypatch.setattr(backend.ops, "prefill_attn", rec.prefill_attn)
    monkeypatch.setattr(backend.ops, "chunked_prefill_attn", rec.chunked_prefill_attn)
    cfg = tmx.TriangleMixConfig(4, 32, 64, triangle_layer_indices=[0])
    impl = tmx.TriangleMixImpl(4, 128, 0.1, num_kv_heads=2, trianglemix_config=cfg, layer_idx=layer_idx, num_layers=2)
    return impl, rec


def _fwd(impl, md, kv_cache=None):
    t = md.num_actual_tokens
    x = torch.zeros(t, 4 * 128, dtype=torch.bfloat16)
    return impl.forward(None, x, x[:, :256], x[:, :256], kv_cache, md, output=torch.zeros_like(x))


def test_backend_threshold_splits_mixed_batches(monkeypatch):
    """Triangle layer: sequences <= 2048 tokens are dense-causal (reference trianglemix_attention.py:330-332), longer
    ones triangle; runs of one class become one kernel call over a contiguous token range."""
    impl, rec = _impl_with_recorder(monkeypatch)
    _fwd(impl, tmx.TriangleMixMetadataBuilder.from_seq_lens([700, 100, 2600, 3000, 2048], "cpu"))
    assert rec.calls == [("prefill", 800, [0, 700, 800], 700, ops.DENSE_CAUSAL),
                         ("prefill", 5600, [0, 2600, 5600], 3000, ops.TRIANGLE),
                         ("prefill", 2048, [0, 2048], 2048, ops.DENSE_CAUSAL)]
    rec.calls.clear()
    _fwd(impl, tmx.TriangleMixMetadataBuilder.from_seq_lens([2049, 5000], "cpu"))        # one class: one call
    assert rec.calls == [("prefill", 7049, [0, 2049, 7049], 5000, ops.TRIANGLE)]
    rec.calls.clear()
    dense, rec2 = _impl_with_recorder(monkeypatch, layer_idx=1)                            # dense layer: never split
    _fwd(dense, tmx.TriangleMixMetadataBuilder.from_seq_lens([700, 2600], "cpu"))
    assert rec2.calls == [("prefill", 3300, [0, 700, 3300], 2600, ops.DENSE_CAUSAL)]


def test_backend_context_path_and_checks(monkeypatch):
    """seq_lens > query lengths (chunked prefill / prefix hit / decode) -> keys from the paged cache; without a
    cache to read it raises instead of attending only the new tokens."""
    impl, rec = _impl_with_recorder(monkeypatch)
    from vllm_trianglemix_b200 import backend
    monkeypatch.setattr(backend.ops, "reshape_and_cache_flash", lambda *a, **k: None)
    bt = torch.zeros(3, 400, dtype=torch.int32)
    md = tmx.TriangleMixMetadataBuilder.from_seq_lens([300, 5000, 2100], "cpu", query_lens=[1, 1000, 2100],
                                                      block_table=bt, slot_mapping=torch.zeros(3101, dtype=torch.int64))
    assert md.context_present()
    with pytest.raises(NotImplementedError, match="context keys"):
        _fwd(impl, md, kv_cache=None)
    kv = torch.zeros(2, 4, 16, 2, 128, dtype=torch.bfloat16)
    _fwd(impl, md, kv_cache=kv)
    assert rec.calls == [("chunked", 1, [0, 1], [300], 1, ops.DENSE_CAUSAL, (1, 400)),
                         ("chunked", 3100, [0, 1000, 3100], [5000, 2100], 2100, ops.TRIANGLE, (2, 400))]
    # whole-sequence metadata is not "context"
    assert not tmx.TriangleMixMetadataBuilder.from_seq_lens([300, 5000], "cpu").context_present()
    # max_query_len below the longest query is refused (grids are sized from it)
    bad = tmx.TriangleMixMetadataBuilder.from_seq_lens([300, 5000], "cpu")
    bad.max_query_len = 4000
    with pytest.raises(ValueError, match="max_query_len"):
        _fwd(impl, bad)


def test_metadata_builder_reads_common_attention_metadata():
    """Field names of vllm/v1/attention/backends/utils.py:45-84 (duck-typed here; the reference's real class is
    exercised in tests/test_reference_wiring.py)."""
    import types
    cu = torch.tensor([0, 5, 6], dtype=torch.int32)
    common = types.SimpleNamespace(
        query_start_loc=cu, query_start_loc_cpu=cu, seq_lens=torch.tensor([5, 40], dtype=torch.int32),
        seq_lens_cpu=torch.tensor([5, 40], dtype=torch.int32), num_computed_tokens_cpu=torch.tensor([0, 39]),
        num_reqs=2, num_actual_tokens=6, max_query_len=5, max_seq_len=40,
        block_table_tensor=torch.zeros(2, 3, dtype=torch.int32), slot_mapping=torch.arange(6), causal=True)
    md = tmx.TriangleMixBackend.get_builder_cls()(None, ["l0"], None, "cpu").build(0, common)
    assert (md.num_actual_tokens, md.max_query_len, md.max_seq_len) == (6, 5, 40)
    assert md.block_table is common.block_table_tensor and md.has_context is True and md.context_present()
    common.num_computed_tokens_cpu = torch.tensor([0, 0])
    common.seq_lens_cpu = torch.tensor([5, 1], dtype=torch.int32)
    common.max_seq_len = 5
    assert not tmx.TriangleMixBackend.get_builder_cls()(None, ["l0"], None, "cpu").build(0, common).context_present()


def test_layer_selector_feeds_the_analyzer(monkeypatch):
    from vllm_trianglemix_b200 import layer_selection
    vals = {0: 3.0, 1: 0.5, 2: 2.0, 3: 0.1}
    monkeypatch.setattr(layer_selection, "middle_qk_gradient",
                        lambda q, *a, **k: torch.full((4, 2), -vals[int(q[0, 0, 0])]))
    sel = tmx.TriangleLayerSelector(4)
    for layer in range(4):
        q = torch.full((4, 2, 128), float(layer))
        sel.observe(layer, q, q, q, q, [4], 0.1)
    assert sel.select(2) == [1, 3]                                   # lowest mean |gradient|
    assert sel.to_config(2).triangle_layer_indices == [1, 3]
