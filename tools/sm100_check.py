"""Developer check: tcgen05 kernel variants vs the mma.sync kernel on a few shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vllm_trianglemix_b200 import ops

dev = "cuda:0"
torch.manual_seed(0)
for (n, hq, hkv, mode) in [(128, 2, 1, 0), (256, 2, 1, 0), (1000, 4, 2, 0), (1000, 4, 2, 1), (4096, 8, 2, 0), (333, 2, 1, 2)]:
    q = torch.randn(n, hq, 128, device=dev).bfloat16()
    k = torch.randn(n, hkv, 128, device=dev).bfloat16()
    v = torch.randn(n, hkv, 128, device=dev).bfloat16()
    cu = torch.tensor([0, n], dtype=torch.int32, device=dev)
    a = ops.prefill_attn(q, k, v, cu, n, 128 ** -0.5, 4, 32, 64, mode, kernel="mma").float()
    b = ops.prefill_attn(q, k, v, cu, n, 128 ** -0.5, 4, 32, 64, mode, kernel="tcgen05").float()
    torch.cuda.synchronize()
    d = (a - b).abs()
    print("variant", os.environ.get("TMX_SM100_VARIANT", "0"), (n, hq, hkv, mode), "max diff %.4e" % d.max().item(),
          "nan", int(torch.isnan(b).sum()), "bad rows", int((d.amax(dim=(1, 2)) > 2e-2).sum()), flush=True)
