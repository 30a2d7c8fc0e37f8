"""Wiring proof against the REFERENCE's own classes.

Run with the reference's `vllm` package first on sys.path (tests/test_reference_wiring.py does that in a
subprocess: PYTHONPATH=baseline/_ref, the offline pip install of the unmodified reference, or /root/reference):

  * builds the reference's `Attention(..., attn_backend=TriangleMixBackend, trianglemix_config=..., layer_idx=...,
    num_layers=...)` (vllm/attention/layer.py:102,186-201) and checks that the layer's `impl` is a TriangleMixImpl
    carrying the per-layer dense-vs-triangle switch;
  * hands `TriangleMixMetadataBuilder.build` a real `CommonAttentionMetadata`
    (vllm/v1/attention/backends/utils.py:45-84);
  * with --gpu: runs the reference layer's own `Attention.forward(q, k, v)` under `set_forward_context`
    (vllm/attention/layer.py:263-345 -> torch.ops.vllm.unified_attention_with_output -> impl.forward) with a real
    paged kv_cache, as a whole-sequence prefill and as a second chunk with context, and checks the output against
    the CPU oracle.
Prints one JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(1, ROOT)          # after the reference root, before site-packages

import types  # noqa: E402

import torch  # noqa: E402

if "--gpu" in sys.argv:
    # On a GPU box the reference selects its CUDA platform, whose module starts with a bare `import vllm._C`
    # (vllm/platforms/cuda.py:18) -- the compiled extension, which the `VLLM_TARGET_DEVICE=empty` install does not
    # contain.  Nothing on this path calls into it (our backend replaces the kernels), so an empty stand-in
    # module lets the unmodified Python of the reference load.
    sys.modules.setdefault("vllm._C", types.ModuleType("vllm._C"))
import vllm  # noqa: E402
from vllm.attention.layer import Attention  # noqa: E402
from vllm.config import CacheConfig, DeviceConfig, VllmConfig, set_current_vllm_config  # noqa: E402
from vllm.v1.attention.backends.utils import CommonAttentionMetadata  # noqa: E402

import vllm_trianglemix_b200 as tmx  # noqa: E402


def main():
    gpu = "--gpu" in sys.argv
    res = {"vllm_version": vllm.__version__, "vllm_file": vllm.__file__}
    assert vllm.__version__.startswith("0.11"), "this must run against the reference's vllm, got %s" % vllm.__file__
    hq, hkv, d, n_layers = 8, 2, 128, 3
    cfg = tmx.TriangleMixConfig(4, 32, 64, triangle_layer_indices=[1, 2])
    vc = VllmConfig(device_config=DeviceConfig(device="cuda" if gpu else "cpu"))
    with set_current_vllm_config(vc):
        layers = [Attention(hq, d, d ** -0.5, num_kv_heads=hkv, cache_config=CacheConfig(block_size=16),
                            prefix="model.layers.%d.self_attn.attn" % i, attn_backend=tmx.TriangleMixBackend,
                            trianglemix_config=cfg, layer_idx=i, num_layers=n_layers) for i in range(n_layers)]
    for i, layer in enumerate(layers):
        assert isinstance(layer.impl, tmx.TriangleMixImpl), type(layer.impl)
        assert layer.attn_backend is tmx.TriangleMixBackend and layer.use_output
        assert layer.impl.num_heads == hq and layer.impl.num_kv_heads == hkv and layer.impl.layer_idx == i
        assert layer.impl.mask_mode() == ((tmx.ops.TRIANGLE if i in (1, 2) else tmx.ops.DENSE_CAUSAL), 4, 32, 64)
    res["impl"] = [type(x.impl).__name__ for x in layers]
    res["modes"] = [x.impl.mask_mode()[0] for x in layers]

    # a real CommonAttentionMetadata: one whole-sequence prefill + one chunk with 100 context tokens
    dev = torch.device("cuda:0" if gpu else "cpu")
    q_lens, seq_lens = [2300, 50], [2300, 150]
    cu = torch.tensor([0, 2300, 2350], dtype=torch.int32)
    lens = torch.tensor(seq_lens, dtype=torch.int32)
    page = 16
    nblk = sum((s + page - 1) // page for s in seq_lens)
    bt = torch.zeros(2, 160, dtype=torch.int32)
    nxt = 0
    for b, s in enumerate(seq_lens):
        k = (s + page - 1) // page
        bt[b, :k] = torch.arange(nxt, nxt + k)
        nxt += k
    slots = torch.cat([bt[0].long()[torch.arange(2300) // page] * page + torch.arange(2300) % page,
                       bt[1].long()[torch.arange(100, 150) // page] * page + torch.arange(100, 150) % page])
    common = CommonAttentionMetadata(
        query_start_loc=cu.to(dev), query_start_loc_cpu=cu, seq_lens=lens.to(dev), seq_lens_cpu=lens,
        num_computed_tokens_cpu=torch.tensor([0, 100], dtype=torch.int32), num_reqs=2, num_actual_tokens=2350,
        max_query_len=2300, max_seq_len=2300, block_table_tensor=bt.to(dev), slot_mapping=slots.to(dev), causal=True)
    builder = tmx.TriangleMixBackend.get_builder_cls()(None, [x.layer_name for x in layers], vc, dev)
    md = builder.build(common_prefix_len=0, common_attn_metadata=common)
    assert isinstance(md, tmx.TriangleMixMetadata)
    assert md.num_actual_tokens == 2350 and md.max_query_len == 2300 and md.max_seq_len == 2300
    assert md.block_table is common.block_table_tensor and md.slot_mapping is common.slot_mapping
    assert md.context_present() is True
    res["metadata"] = "ok"

    if gpu:
        from vllm.forward_context import set_forward_context
        from oracle import trianglemix_oracle as orc
        g = torch.Generator().manual_seed(0)
        full = [(torch.randn(s, hq, d, generator=g).bfloat16(), torch.randn(s, hkv, d, generator=g).bfloat16(),
                 torch.randn(s, hkv, d, generator=g).bfloat16()) for s in seq_lens]
        for layer in layers:
            kv = torch.zeros(tmx.TriangleMixBackend.get_kv_cache_shape(nblk + 1, page, hkv, d), dtype=torch.bfloat16,
                             device=dev)
            # the context of sequence 1 (its first 100 tokens) is already in the cache
            ctx_slots = bt[1].long()[torch.arange(100) // page] * page + torch.arange(100) % page
            kv[0].view(-1, hkv, d)[ctx_slots.to(dev)] = full[1][1][:100].to(dev)
            kv[1].view(-1, hkv, d)[ctx_slots.to(dev)] = full[1][2][:100].to(dev)
            layer.kv_cache = [kv]
        q = torch.cat([full[0][0], full[1][0][100:]]).to(dev).view(2350, hq * d)
        k = torch.cat([full[0][1], full[1][1][100:]]).to(dev).view(2350, hkv * d)
        v = torch.cat([full[0][2], full[1][2][100:]]).to(dev).view(2350, hkv * d)
        errs = []
        with set_forward_context({x.layer_name: md for x in layers}, vc):
            for i, layer in enumerate(layers):
                out = layer(q, k, v)                       # the REFERENCE's Attention.forward
                torch.cuda.synchronize()
                # sequence 0 is 2300 > 2048 tokens: triangle in layers 1, 2; sequence 1 (150 tokens so far) is below
                # the reference helper's threshold: the batch mixes both, which the context path refuses -> per
                # layer type: dense layers accept it
                mode = layer.impl.mask_mode()[0]
                ref0 = orc.attention_varlen_fp32(*full[0], [0, 2300], d ** -0.5, 4, 32, 64, mode)
                ref1 = orc.attention_rows_fp32(*full[1], d ** -0.5, 150, 4, 32, 64, orc.DENSE_CAUSAL, 100, 150)
                rep0 = orc.error_report(out[:2300].cpu().view(-1, hq, d), ref0)
                rep1 = orc.error_report(out[2300:].cpu().view(-1, hq, d), ref1)
                errs.append((i, rep0["violations"], rep1["violations"], rep0["max_abs"], rep1["max_abs"]))
        res["forward"] = errs
        assert all(e[1] == 0 and e[2] == 0 for e in errs), errs
    print(json.dumps(res))


if __name__ == "__main__":
    main()
