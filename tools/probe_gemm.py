"""Developer probe: the o_proj GEMM (ops.linear, tcgen05) against torch.matmul (cuBLAS) at Qwen3 shapes."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vllm_trianglemix_b200 import ops


def time_ms(fn, warm=3, it=10):
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


def main():
    dev = "cuda:0"
    for (m, n, k, what) in [(131072, 4096, 4096, "Qwen3-8B o_proj, 1 GPU"), (131072, 4096, 512, "Qwen3-8B o_proj shard at P=8"),
                            (131072, 5120, 1024, "Qwen3-32B o_proj shard at P=8"), (8192, 4096, 4096, "8K-token chunk")]:
        a = torch.randn(m, k, device=dev, dtype=torch.bfloat16)
        w = (torch.randn(n, k, device=dev, dtype=torch.bfloat16) * k ** -0.5)
        out = torch.empty(m, n, device=dev, dtype=torch.bfloat16)
        ours = time_ms(lambda: ops.linear(a, w, out=out))
        ref = torch.empty_like(out)
        cublas = time_ms(lambda: torch.matmul(a, w.t(), out=ref))
        fl = 2.0 * m * n * k
        by = 2.0 * (m * k + n * k + m * n)
        print("linear %-32s M=%d N=%d K=%d: ours %.3f ms %.0f TFLOP/s %.0f GB/s | cuBLAS %.3f ms %.0f TFLOP/s | max diff %.3g" %
              (what, m, n, k, ours, fl / ours / 1e9, by / ours / 1e6, cublas, fl / cublas / 1e9,
               (out.float() - ref.float()).abs().max().item()), flush=True)


if __name__ == "__main__":
    main()
