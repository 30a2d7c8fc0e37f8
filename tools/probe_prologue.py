"""Developer probe: time the fused q/k RMSNorm + RoPE (+ KV-cache write) kernel (CUDA events)
at Qwen3-8B shape, in place on a packed qkv buffer, against its algorithmic HBM bytes."""
import argparse
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vllm_trianglemix_b200 import ops, compute_cos_sin_cache
from tools.probe import time_ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, nargs="+", default=[32768, 131072])
    ap.add_argument("--hq", type=int, default=32)
    ap.add_argument("--hkv", type=int, default=8)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    dev, d, bs = "cuda:0", 128, 16
    res = []
    for n in args.n:
        hq, hkv = args.hq, args.hkv
        qkv = torch.randn(n, (hq + 2 * hkv) * d, device=dev, dtype=torch.bfloat16)
        q, k, v = qkv.split([hq * d, hkv * d, hkv * d], dim=-1)
        w = torch.ones(d, device=dev, dtype=torch.bfloat16)
        cache = compute_cos_sin_cache(d, n, 1000000.0, device=dev)
        pos = torch.arange(n, device=dev)
        kv = torch.zeros(2, (n + bs - 1) // bs, bs, hkv, d, device=dev, dtype=torch.bfloat16)
        slots = torch.arange(n, device=dev)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        for name, fn, by in (
            ("norm_rope", lambda: ops.qk_norm_rope(pos, q, k, w, w, cache, 1e-6, d, inplace=True),
             2 * n * (hq + hkv) * d * 2 + n * d * 2 + n * 8),
            ("norm_rope_cache", lambda: ops.qk_norm_rope(pos, q, k, w, w, cache, 1e-6, d, inplace=True, v=v,
                                                         key_cache=kv[0], value_cache=kv[1], slot_mapping=slots),
             2 * n * (hq + hkv) * d * 2 + 3 * n * hkv * d * 2 + n * d * 2 + 2 * n * 8),
            ("cache_only", lambda: ops.reshape_and_cache_flash(k.view(n, hkv, d), v.view(n, hkv, d), kv[0], kv[1],
                                                               slots),
             4 * n * hkv * d * 2 + n * 8),
        ):
            def step():
                flush.zero_()          # 256 MiB > 126 MB L2: every timed launch starts cold
                fn()
            t_flush = time_ms(lambda: flush.zero_(), 2, 10)
            ms = time_ms(step, 3, 10) - t_flush
            res.append({"kernel": name, "n": n, "ms": ms, "algorithmic_bytes": by, "gbs": by / ms / 1e6})
            print("%s n=%d: %.4f ms  %.1f GB/s algorithmic (L2 flushed)" % (name, n, ms, by / ms / 1e6), flush=True)
    if args.json:
        with open(args.json, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
