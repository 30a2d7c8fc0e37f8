import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vllm_trianglemix_b200 import ops
DEV = "cuda:0"
n, hq, hkv = int(sys.argv[1]) if len(sys.argv) > 1 else 32768, 32, 8
g = torch.Generator(device=DEV).manual_seed(17)
q = torch.randn(n, hq, 128, generator=g, device=DEV).bfloat16()
k = torch.randn(n, hkv, 128, generator=g, device=DEV).bfloat16()
v = torch.randn(n, hkv, 128, generator=g, device=DEV).bfloat16()
cu = torch.tensor([0, n], dtype=torch.int32, device=DEV)
for kern in ("tcgen05", "mma"):
    a = ops.prefill_attn(q, k, v, cu, n, 128 ** -0.5, 4, 32, 64, 0, kernel=kern)
    b = ops.prefill_attn(q, k, v, cu, n, 128 ** -0.5, 4, 32, 64, 0, kernel=kern)
    c = ops.prefill_attn(q, k, v * 4, cu, n, 128 ** -0.5, 4, 32, 64, 0, kernel=kern)
    torch.cuda.synchronize()
    print(kern, "deterministic:", torch.equal(a, b), "linear:", torch.equal(c, a * 4))
    for name, x, y in (("rerun", a, b), ("x4", c, a * 4)):
        d = (x.float() - y.float()).abs()
        bad = (d > 0).nonzero()
        print(" ", name, "ndiff", bad.shape[0], "maxdiff", float(d.max()))
        if bad.shape[0]:
            rows = bad[:, 0].unique()
            print("   rows", rows[:20].tolist(), "... count", rows.numel(), "heads", bad[:, 1].unique().tolist()[:40])
            print("   rows mod 128 hist:", torch.bincount(rows % 128, minlength=128).tolist())
