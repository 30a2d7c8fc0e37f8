"""Generate golden vectors by running the UNMODIFIED reference in the build container.

Usage (build container only; /root/reference does not exist on the GPU box):
    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

Loads the reference modules by file path (stub parent packages so that
``vllm/attention/__init__.py`` -- which drags in the whole engine -- never runs)
and writes
    tests/golden/masks.json      counts, sha256 of LSB-first packed bits, switch/threshold answers
    tests/golden/masks_small.npz packed uint32 bitmasks of the five pieces at small N
    tests/golden/attn_small.npz  inputs + fp32 outputs of the reference attention op
"""
import hashlib
import importlib.util
import json
import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("TMX_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


def load_reference():
    for name in ("vllm", "vllm.attention"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__path__ = []
            sys.modules[name] = m
    mods = {}
    for short in ("trianglemix_attention", "ascend_trianglemix", "trianglemix_config"):
        full = "vllm.attention." + short
        spec = importlib.util.spec_from_file_location(
            full, os.path.join(REF, "vllm", "attention", short + ".py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[full] = mod
        spec.loader.exec_module(mod)
        mods[short] = mod
    return mods


def pack(mask01: torch.Tensor) -> np.ndarray:
    m = (mask01 > 0).numpy().astype(np.uint8)
    r, n = m.shape
    words = (n + 31) // 32
    pad = np.zeros((r, words * 32), np.uint8)
    pad[:, :n] = m
    return np.packbits(pad, axis=1, bitorder="little").view("<u4").reshape(r, words)


PARAMS = [  # (N, si, sl, last)
    (12, 2, 4, 6), (16, 3, 6, 8), (24, 4, 8, 16), (97, 4, 32, 64), (300, 4, 32, 64),
    (130, 0, 0, 1), (64, 7, 3, 200), (64, 100, 3, 5), (33, 4, 32, 64), (1, 4, 32, 64),
    (2, 0, 0, 0), (512, 4, 32, 64), (257, 1, 5, 9), (384, 16, 64, 128),
]
BIG = [(2048, 4, 32, 64), (4096, 4, 8, 16)]   # triangle only (loops take seconds)


def main():
    mods = load_reference()
    tm = mods["trianglemix_attention"]
    asc = mods["ascend_trianglemix"]
    cfgm = mods["trianglemix_config"]
    dev = torch.device("cpu")
    out = {"cases": [], "big": []}
    small = {}
    for (n, si, sl, last) in PARAMS:
        gen = tm.TriangleMixMaskGenerator(tm.TriangleMixConfig(si, sl, last))
        rec = {"N": n, "si": si, "sl": sl, "num_last": last}
        for piece in ("streaming", "last", "middle", "triangle", "dense"):
            m = getattr(gen, "_create_%s_mask" % piece)(n, dev, torch.float32)
            bits = pack(m)
            rec[piece] = int(m.sum().item())
            rec[piece + "_sha"] = hashlib.sha256(bits.tobytes()).hexdigest()
            if n <= 300:
                small["%s_%d_%d_%d_%d" % (piece, n, si, sl, last)] = bits
        add = gen.get_attention_mask(n, 0, dev, torch.float32, use_triangle=True)
        rec["additive_neg_inf"] = int(torch.isinf(add).sum().item())
        rec["additive_zero"] = int((add == 0).sum().item())
        out["cases"].append(rec)
        print(rec["N"], rec["triangle"], rec["streaming"], rec["last"], rec["middle"], rec["dense"])
    for (n, si, sl, last) in BIG:
        gen = tm.TriangleMixMaskGenerator(tm.TriangleMixConfig(si, sl, last))
        m = gen._create_triangle_mask(n, dev, torch.float32)
        out["big"].append({"N": n, "si": si, "sl": sl, "num_last": last,
                           "triangle": int(m.sum().item()),
                           "triangle_sha": hashlib.sha256(pack(m).tobytes()).hexdigest()})
        print(n, out["big"][-1]["triangle"])

    # layer switch / threshold / analyzer / -inf rewrite (reference test script :132-220)
    cfg = tm.TriangleMixConfig(num_triangle_layers=4)
    helper = tm.TriangleMixAttention(cfg, num_layers=8)
    out["switch_first4_of_8"] = [helper.mask_generator.should_use_triangle(i) for i in range(8)]
    out["config_indices_after_helper"] = list(cfg.triangle_layer_indices)
    out["threshold_none_at"] = {str(n): helper.get_attn_mask(n, 0, dev) is None
                                for n in (1, 512, 2048)}
    cfg2 = tm.TriangleMixConfig(triangle_layer_indices=[5, 7], num_triangle_layers=2)
    g2 = tm.TriangleMixMaskGenerator(cfg2)
    out["switch_explicit_5_7"] = [g2.should_use_triangle(i) for i in range(8)]
    g3 = tm.TriangleMixMaskGenerator(tm.TriangleMixConfig())
    out["switch_unset"] = [g3.should_use_triangle(i) for i in range(4)]
    an = tm.TriangleMixAnalyzer(num_layers=8)
    for layer in range(8):
        for _ in range(3):
            an.record_middle_qk_gradient(layer, torch.full((4, 4), 1.0 / (layer + 1)))
    out["analyzer_pick4"] = an.get_triangle_layers(4)
    opt = asc.AscendNPUTriangleMixOptimizer()
    add16 = tm.TriangleMixMaskGenerator(tm.TriangleMixConfig(3, 6, 8)).get_attention_mask(
        16, 0, dev, torch.float32, True)
    rew = opt.optimize_mask_for_npu(add16)
    out["rewrite16"] = {"any_inf": bool(torch.isinf(rew).any()),
                        "count_-1e6": int((rew == -1e6).sum()), "count_0": int((rew == 0).sum())}
    for size in ("8B", "14B", "32B"):
        ic = cfgm.TriangleMixInferenceConfig.for_qwen3(size)
        tc = ic.to_trianglemix_config()
        out["for_qwen3_" + size] = {"num_triangle_layers": tc.num_triangle_layers,
                                    "si": tc.num_sink_tokens, "sl": tc.sliding_window_size,
                                    "last": tc.num_last_tokens}

    class HF:
        sliding_window = 128
    hc = asc.create_trianglemix_config_from_hf(HF(), use_trianglemix=True, num_triangle_layers=6)
    out["from_hf_sw128"] = None if hc is None else {
        "si": hc.num_sink_tokens, "sl": hc.sliding_window_size, "last": hc.num_last_tokens,
        "k": hc.num_triangle_layers}
    out["from_hf_disabled_is_none"] = asc.create_trianglemix_config_from_hf(HF()) is None

    # attention op outputs (fp32 on bf16-rounded inputs), GQA via repeat_interleave
    attn = {}
    torch.manual_seed(0)
    for tag, (s, hq, hkv, d, si, sl, last) in {
            "a": (96, 4, 2, 128, 2, 8, 12), "b": (288, 2, 1, 128, 4, 32, 64),
            "c": (130, 2, 1, 128, 0, 0, 1)}.items():
        q = torch.randn(1, s, hq, d).bfloat16()
        k = torch.randn(1, s, hkv, d).bfloat16()
        v = torch.randn(1, s, hkv, d).bfloat16()
        scale = d ** -0.5
        op = asc.AscendTriangleMixAttention(hq, d, scale)
        gen = tm.TriangleMixMaskGenerator(tm.TriangleMixConfig(si, sl, last))
        kk = k.float().repeat_interleave(hq // hkv, dim=2)
        vv = v.float().repeat_interleave(hq // hkv, dim=2)
        for mode, tri in (("tri", True), ("dense", False)):
            mask = gen.get_attention_mask(s, 0, dev, torch.float32, use_triangle=tri)
            o = op.forward(q.float(), kk, vv, attn_mask=mask)
            attn["%s_%s_out" % (tag, mode)] = o.numpy()
        if tag == "a":
            o = op.forward(q.float(), kk, vv)      # no mask at all: non-causal quirk
            attn[tag + "_nomask_out"] = o.numpy()
        attn[tag + "_q"] = q.view(torch.int16).numpy()
        attn[tag + "_k"] = k.view(torch.int16).numpy()
        attn[tag + "_v"] = v.view(torch.int16).numpy()
        attn[tag + "_params"] = np.array([s, hq, hkv, d, si, sl, last])

    with open(os.path.join(HERE, "masks.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    np.savez_compressed(os.path.join(HERE, "masks_small.npz"), **small)
    np.savez_compressed(os.path.join(HERE, "attn_small.npz"), **attn)


if __name__ == "__main__":
    main()
