"""Developer probe: time the attention / mask kernels at a few shapes (CUDA events)."""
import argparse
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vllm_trianglemix_b200 import _lib
if "--dev" in sys.argv:            # developer library (make DEV=1): honours TMX_SM100_VARIANT
    sys.argv.remove("--dev")
    _lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), os.environ.get("TMX_DEV_LIB", "libtmx_dev.so"))
from vllm_trianglemix_b200 import ops


def time_ms(fn, warm=2, it=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


def pairs(n, si, sl, last, mode):
    if mode == 0:
        return n * (n + 1) // 2
    lrow = max(0, n - last + 1)
    tot = 0
    for i in (range(n)):
        tot += (i + 1) if i >= lrow else (min(i, si) + 1 + min(i, sl) + 1 - max(0, min(i, si) - max(0, i - sl) + 1))
    return tot


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, nargs="+", default=[8192, 32768, 131072])
    ap.add_argument("--hq", type=int, default=32)
    ap.add_argument("--hkv", type=int, default=8)
    ap.add_argument("--kernels", nargs="+", default=["mma"])
    ap.add_argument("--modes", type=int, nargs="+", default=[1, 0])
    ap.add_argument("--params", type=int, nargs=3, default=[4, 32, 64])
    ap.add_argument("--iters", type=int, default=0, help="timed iterations (default: 3 for long dense runs, else 10)")
    args = ap.parse_args()
    dev = "cuda:0"
    si, sl, last = args.params
    for n in args.n:
        q = torch.randn(n, args.hq, 128, device=dev, dtype=torch.bfloat16)
        k = torch.randn(n, args.hkv, 128, device=dev, dtype=torch.bfloat16)
        v = torch.randn(n, args.hkv, 128, device=dev, dtype=torch.bfloat16)
        out = torch.empty_like(q)
        cu = torch.tensor([0, n], dtype=torch.int32, device=dev)
        for kern in args.kernels:
            for mode in args.modes:
                it = args.iters or (3 if (mode == 0 and n > 40000) else 10)
                ms = time_ms(lambda: ops.prefill_attn(q, k, v, cu, n, 128 ** -0.5, si, sl, last, mode, out=out, kernel=kern), 2, it)
                p = pairs(n, si, sl, last, mode)
                fl = 4 * 128 * args.hq * p
                by = 2 * n * 128 * 2 * (args.hq + args.hkv)
                print("attn n=%d kernel=%s mode=%d: %.3f ms  %.1f TFLOP/s eff  %.1f GB/s alg" %
                      (n, kern, mode, ms, fl / ms / 1e9, by / ms / 1e6), flush=True)
        if n <= 131072:
            bits = ops.mask_bits(n, si, sl, last, "triangle", dev)
            ms = time_ms(lambda: ops.mask_bits(n, si, sl, last, "triangle", dev, out=bits))
            print("mask_bits n=%d: %.3f ms %.1f GB/s" % (n, ms, bits.numel() * 4 / ms / 1e6), flush=True)
            del bits
        if n <= 32768:
            ms = time_ms(lambda: ops.mask_additive(n, si, sl, last, 1, dev, torch.bfloat16), 1, 3)
            print("mask_additive bf16 n=%d: %.3f ms %.1f GB/s" % (n, ms, n * n * 2 / ms / 1e6), flush=True)


if __name__ == "__main__":
    main()
