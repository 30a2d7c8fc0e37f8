"""Developer check: which rows of the tcgen05 kernel differ from the mma.sync kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vllm_trianglemix_b200 import ops

dev = "cuda:0"
torch.manual_seed(0)
for (n, hq, hkv, mode, si, sl, last) in [(1000, 4, 2, 1, 4, 32, 64), (1024, 2, 1, 1, 4, 32, 64), (512, 2, 1, 1, 4, 32, 64),
                                         (1000, 2, 1, 1, 4, 32, 0), (1000, 2, 1, 1, 200, 32, 0)]:
    q = torch.randn(n, hq, 128, device=dev).bfloat16()
    k = torch.randn(n, hkv, 128, device=dev).bfloat16()
    v = torch.randn(n, hkv, 128, device=dev).bfloat16()
    cu = torch.tensor([0, n], dtype=torch.int32, device=dev)
    a = ops.prefill_attn(q, k, v, cu, n, 128 ** -0.5, si, sl, last, mode, kernel="mma").float()
    b = ops.prefill_attn(q, k, v, cu, n, 128 ** -0.5, si, sl, last, mode, kernel="tcgen05").float()
    torch.cuda.synchronize()
    d = (a - b).abs().amax(dim=2)     # [n, hq]
    bad = (d > 2e-2).nonzero()
    rows = sorted(set(bad[:, 0].tolist()))
    print((n, hq, hkv, mode, si, sl, last), "bad rows", len(rows), rows[:40], "heads", sorted(set(bad[:, 1].tolist())), flush=True)
    ranges, counts, full = ops.block_index(cu, n, 128, 64, si, sl, last, mode)
    print(" ranges", ranges.tolist(), flush=True)
