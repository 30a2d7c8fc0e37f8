"""Developer check: a TMX_SM100_VARIANT of the pair kernel (dev library) against the mma.sync kernel on ragged
batches, dense / triangle-by-block-index / unmasked, incl. peaky scores (exercises the refresh path)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vllm_trianglemix_b200 import _lib
_lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libtmx_dev.so")
from vllm_trianglemix_b200 import ops  # noqa: E402

dev = "cuda:0"
bad = 0
for seed, (lens, hq, hkv, qs) in enumerate([([1000], 4, 1, 1.0), ([3000, 129, 1, 700], 8, 2, 1.0), ([2048], 16, 2, 4.0),
                                            ([5000], 8, 1, 8.0), ([130, 131], 4, 1, 0.3)]):
    t = sum(lens)
    g = torch.Generator().manual_seed(seed)
    q = (torch.randn(t, hq, 128, generator=g) * qs).bfloat16().to(dev)
    k = torch.randn(t, hkv, 128, generator=g).bfloat16().to(dev)
    v = torch.randn(t, hkv, 128, generator=g).bfloat16().to(dev)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=dev)
    for mode, prm in ((0, (0, 0, 0)), (1, (20, 150, 40)), (2, (0, 0, 0))):
        a, la = ops.prefill_attn(q, k, v, cu, max(lens), 128 ** -0.5, *prm, mode, kernel="pair", return_lse=True)
        b, lb = ops.prefill_attn(q, k, v, cu, max(lens), 128 ** -0.5, *prm, mode, kernel="mma", return_lse=True)
        d = (a.float() - b.float()).abs()
        viol = int((d > 1.5e-2 + 1e-2 * b.float().abs()).sum())
        lse_d = float((la - lb).abs().max())
        bad += viol + (lse_d > 2e-2)
        print("variant %s lens %s hq %d mode %d: max diff %.3e viol %d lse diff %.3e" %
              (os.environ.get("TMX_SM100_VARIANT", "0"), lens, hq, mode, float(d.max()), viol, lse_d), flush=True)
print("VARIANT CHECK", "OK" if bad == 0 else "FAILED")
