"""Developer comparator (never part of the product or of bench.py): the dense-causal layer of the
workload through the kernels the reference delegates dense attention to -- flash-attn 2.8
(flash_attn_varlen_func, vllm/v1/attention/backends/flash_attn.py:526) and FlashInfer's
trtllm-gen Blackwell context kernel (vllm/v1/attention/backends/flashinfer.py:943) -- on the same
box, same shape, CUDA events, next to libtmx's tcgen05 kernel.  Library code: comparison only.
Also cross-checks the outputs against each other."""
import argparse
import json
import os
import sys
import time
import traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vllm_trianglemix_b200 import ops
from tools.probe import time_ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, nargs="+", default=[32768, 131072])
    ap.add_argument("--hq", type=int, default=32)
    ap.add_argument("--hkv", type=int, default=8)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    dev, d = "cuda:0", 128
    scale = d ** -0.5
    res = []
    for n in args.n:
        torch.manual_seed(0)
        hq, hkv = args.hq, args.hkv
        q = torch.randn(n, hq, d, device=dev, dtype=torch.bfloat16)
        k = torch.randn(n, hkv, d, device=dev, dtype=torch.bfloat16)
        v = torch.randn(n, hkv, d, device=dev, dtype=torch.bfloat16)
        cu = torch.tensor([0, n], dtype=torch.int32, device=dev)
        flops = 4.0 * d * hq * (n * (n + 1) // 2)
        it = 3 if n > 40000 else 10

        def record(name, fn, ref=None):
            t0 = time.time()
            try:
                out = fn()
                torch.cuda.synchronize()
                ms = time_ms(fn, 1, it)
                r = {"kernel": name, "n": n, "ms": ms, "tflops_eff": flops / ms / 1e9,
                     "first_call_s": round(time.time() - t0, 1)}
                if ref is not None:
                    r["max_abs_vs_tmx"] = float((out.float().view(n, hq, d) - ref.float()).abs().max().item())
                res.append(r)
                print(json.dumps(r), flush=True)
                return out
            except Exception as e:  # noqa: BLE001 -- comparators are optional
                r = {"kernel": name, "n": n, "error": repr(e)[:400]}
                res.append(r)
                print(json.dumps(r), flush=True)
                traceback.print_exc(limit=2)
                return None

        ours = torch.empty_like(q)
        record("libtmx (this repo, kernel=auto: CTA-pair tcgen05 kernel)",
               lambda: ops.prefill_attn(q, k, v, cu, n, scale, 0, 0, 0, ops.DENSE_CAUSAL, out=ours))

        def fa2():
            from flash_attn import flash_attn_varlen_func
            return flash_attn_varlen_func(q, k, v, cu, cu, n, n, softmax_scale=scale, causal=True)
        record("flash-attn 2.8.3 flash_attn_varlen_func", fa2, ours)

        def fi_trtllm():
            from flashinfer import prefill
            page = 64
            npages = n // page
            kvc = torch.stack([k.view(npages, page, hkv, d).permute(0, 2, 1, 3),
                               v.view(npages, page, hkv, d).permute(0, 2, 1, 3)], dim=1).contiguous()
            ws = getattr(fi_trtllm, "ws", None)
            if ws is None:
                ws = fi_trtllm.ws = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
            bt = torch.arange(npages, dtype=torch.int32, device=dev).view(1, -1)
            sl = torch.tensor([n], dtype=torch.int32, device=dev)
            state = getattr(fi_trtllm, "state", None)
            if state is None or state[0] != n:
                fi_trtllm.state = (n, kvc, bt, sl)
            _, kvc, bt, sl = fi_trtllm.state
            return prefill.trtllm_batch_context_with_kv_cache(q, kvc, ws, bt, sl, n, n, scale, 1.0, 1, cu, cu,
                                                              kv_layout="HND")
        record("FlashInfer 0.6.11 trtllm-gen context (paged KV, page 64)", fi_trtllm, ours)

        def fi_single():
            import flashinfer
            return flashinfer.single_prefill_with_kv_cache(q, k, v, causal=True, sm_scale=scale)
        record("FlashInfer 0.6.11 single_prefill_with_kv_cache (auto backend)", fi_single, ours)
    if args.json:
        with open(args.json, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
